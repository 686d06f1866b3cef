"""Achieved HBM bandwidth of the memory-bound helper kernels (SURVEY 8 f1 / f2 and the ring helpers) against the measured copy
peak (MEASURED_PEAKS.json hbm_gbs): algorithmic bytes / time, CUDA events, L2 flushed between repetitions.
    python tools/helper_bandwidth.py > profiles/r02_helper_bandwidth.jsonl"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
import flash_attn_cuda as fc

dev = "cuda"
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def timed(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps


def emit(kernel, what, nbytes, ms):
    gbs = nbytes / ms / 1e6
    print(json.dumps({"kernel": kernel, "case": what, "algorithmic_bytes": nbytes, "ms": ms, "GB/s": gbs, "frac_of_copy_peak": gbs / peak,
                      "peak_GB/s": peak}), flush=True)


# BERT-large-ish varlen batch: b = 64 sequences of up to 512 tokens, hidden 1024 (fp16), 'third' padding
b, s, hid = 64, 512, 1024
x = torch.randn(b, s, 3 * hid, device=dev, dtype=torch.float16)          # a QKV activation: rows of 6 KB
lens = torch.randint(s // 3, s + 1, (b,), device=dev)
mask = torch.arange(s, device=dev)[None] < lens[:, None]
idx = torch.nonzero(mask.flatten()).flatten()
flat = x.reshape(b * s, -1)
nnz, row = idx.numel(), flat.shape[1] * 2
packed = fc.gather_rows(flat, idx)
emit("gather_rows", f"unpad_input of ({b},{s},{3*hid}) fp16, {nnz} of {b*s} rows", 2 * nnz * row + nnz * 8, timed(lambda: fc.gather_rows(flat, idx)))
out = torch.empty_like(flat)
emit("scatter_rows", "pad_input (rows only; the zero fill of the padded rows is a separate memset)", 2 * nnz * row + nnz * 8,
     timed(lambda: fc.scatter_rows(packed, idx, b * s, out=out)))
emit("torch index (reference's own path: x[indices])", "same gather through torch advanced indexing", 2 * nnz * row + nnz * 8, timed(lambda: flat[idx]))

# rotary on the q and k thirds of a (b*s, 3, h, d) activation, in place
h, d = 16, 64
qkv = torch.randn(b * s, 3, h, d, device=dev, dtype=torch.float16)
cos = torch.randn(s, d // 2, device=dev)
sin = torch.randn(s, d // 2, device=dev)
emit("rotary_qk", f"in place on q, k of ({b*s},3,{h},{d}) fp16", 2 * (2 * b * s * h * d * 2) + 2 * s * (d // 2) * 4,
     timed(lambda: fc.rotary_qk(qkv, qkv, cos, sin, s)))

# ring helpers at the 64K / 8-GPU chunk size: rows = 2c = 8192, h = 12, d = 128
rows, h, d = 8192, 12, 128
o_acc = torch.randn(rows, h, d, device=dev)
lse_acc = torch.randn(1, h, rows, device=dev)
o_blk = torch.randn(rows, h, d, device=dev, dtype=torch.bfloat16)
lse_blk = torch.randn(1, h, rows, device=dev)
emit("merge_partials", f"({rows},{h},{d}) fp32 accumulator (read + write) + bf16 block + 2 lse", rows * h * d * (4 + 4 + 2) + 3 * rows * h * 4,
     timed(lambda: fc.merge_partials(o_acc, lse_acc, o_blk, lse_blk, 1, rows, False)))
emit("accumulate_f32", f"({rows},{h},{d}) fp32 += bf16", rows * h * d * (4 + 4 + 2), timed(lambda: fc.accumulate_f32(o_acc, o_blk, 1, rows * h * d, rows * h * d, rows * h * d)))

# backward preprocess at the headline size (reads O and dO, writes D / lse copies)
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
bb, ss, hh, dd = 16, 4096, 12, 128
qkv = torch.randn(bb * ss, 3, hh, dd, device=dev, dtype=torch.bfloat16, requires_grad=True)
cu = torch.arange(0, (bb + 1) * ss, ss, device=dev, dtype=torch.int32)
o = flash_attn_unpadded_qkvpacked_func(qkv, cu, ss, 0.0, causal=True)
g = torch.randn_like(o)
fc.kernel_timing(True)
e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.autograd.grad(o, qkv, g, retain_graph=True)
torch.cuda.synchronize()
fc.kernel_timing_collect()
tot, reps = 0.0, 5
for _ in range(reps):
    flush.zero_()
    e[0].record()
    torch.autograd.grad(o, qkv, g, retain_graph=True)
    e[1].record()
    torch.cuda.synchronize()
    tot += e[0].elapsed_time(e[1])
kt = fc.kernel_timing_collect()
fc.kernel_timing(False)
pre_ms = tot / reps - kt["bwd_ms"] / kt["bwd_launches"]
emit("bwd_preprocess (+ launch gaps: whole backward minus its main kernel)", f"b{bb} h{hh} d{dd} S{ss}: reads O, dO, lse; writes D, -lse*log2e",
     2 * bb * ss * hh * dd * 2 + 4 * bb * ss * hh * 4, pre_ms)
