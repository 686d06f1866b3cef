"""Kernel-level timing of one configuration: python tools/time_cfg.py b h d s causal p dtype(bf16|fp16) [reps] [det]
(det: bitwise reproducible dQ, FA_FLAG_DETERMINISTIC)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
import flash_attn_cuda
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func

b, h, d, s = [int(x) for x in sys.argv[1:5]]
causal, p = sys.argv[5] == "1", float(sys.argv[6])
dtype = torch.bfloat16 if sys.argv[7] == "bf16" else torch.float16
reps = int(sys.argv[8]) if len(sys.argv) > 8 else 10
if len(sys.argv) > 9 and sys.argv[9] == "det":
    flash_attn_cuda.set_deterministic(True)
qkv = torch.randn(b * s, 3, h, d, device="cuda", dtype=dtype, requires_grad=True)
dout = torch.randn(b * s, h, d, device="cuda", dtype=dtype)
cu = torch.arange(0, (b + 1) * s, s, device="cuda", dtype=torch.int32)
for _ in range(3):
    out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=causal)
    torch.autograd.grad(out, qkv, dout)
torch.cuda.synchronize()
flash_attn_cuda.kernel_timing(True)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
tf = tb = 0.0
for _ in range(reps):
    ev[0].record()
    out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=causal)
    ev[1].record()
    torch.autograd.grad(out, qkv, dout)
    ev[2].record()
    torch.cuda.synchronize()
    tf += ev[0].elapsed_time(ev[1]); tb += ev[1].elapsed_time(ev[2])
flash_attn_cuda.kernel_timing(False)
kt = flash_attn_cuda.kernel_timing_collect()
f = 4.0 * b * h * s * s * d / (2 if causal else 1)
kf, kb = kt["fwd_ms"] / kt["fwd_launches"], kt["bwd_ms"] / kt["bwd_launches"]
print(f"b{b} h{h} d{d} S{s} causal={int(causal)} p={p} {sys.argv[7]}: fwd {tf/reps:.3f} ms (kernel {kf:.3f} ms, {f/kf/1e9:.0f} TF/s)  "
      f"bwd {tb/reps:.3f} ms (kernel {kb:.3f} ms, {2.5*f/kb/1e9:.0f} TF/s)  fwd+bwd {3.5*f/((tf+tb)/reps)/1e9:.0f} TF/s")
