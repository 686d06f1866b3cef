"""bench.py -- attention fwd+bwd throughput on B200 (BASELINE.json metric), one JSON line on stdout.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload d128|d64|bert|ring64k] [--sweep]

A "step" is one forward+backward pass of the attention hot path over one synthetic batch, through the public API
(flash_attn_unpadded_qkvpacked_func -> flash_attn_cuda -> libfa_b200.so).  N > 1: one process per GPU (torchrun),
the (batch x heads) units are sharded with no data-path collective (weak scaling: every rank runs the full
per-GPU workload); time = max over ranks.
FLOP model: F_fwd = 4*b*h*sq*sk*d (/2 causal), F_bwd = 2.5*F_fwd, F_fwd+bwd = 3.5*F_fwd.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "hazyresearch_flash-attention_b200"), ROOT):
    if _p in sys.path:
        sys.path.remove(_p)
    sys.path.insert(0, _p)

import torch  # noqa: E402

WORKLOADS = {
    # BASELINE.json configs[2] / north_star target: head dim 128 bf16 b16 h12, seqlen >= 4K, causal
    "d128": dict(b=16, h=12, d=128, s=4096, causal=True, p=0.0, dtype="bf16",
                 name="b16 h12 d128 S4096 causal bf16 fwd+bwd (BASELINE configs[2], north_star target)"),
    # BASELINE.json configs[1]: GPT-2 causal fp16 b8 h12 d64, dropout 0.1
    "d64": dict(b=8, h=12, d=64, s=4096, causal=True, p=0.1, dtype="fp16",
                name="b8 h12 d64 S4096 causal fp16 dropout 0.1 fwd+bwd (BASELINE configs[1])"),
    # BASELINE.json configs[0]: BERT-base non-causal S512
    "bert": dict(b=8, h=12, d=64, s=512, causal=False, p=0.0, dtype="fp16",
                 name="b8 h12 d64 S512 non-causal fp16 fwd+bwd (BASELINE configs[0])"),
    # ONE 64K-token causal problem, sequence-sharded (zig-zag) over the GPUs with the K/V ring
    "ring64k": dict(b=1, h=12, d=128, s=65536, causal=True, p=0.0, dtype="bf16", ring=True,
                    name="b1 h12 d128 S65536 causal bf16 fwd+bwd, zig-zag K/V ring over the GPUs (SURVEY 8e C5)"),
}


def flops_fwd(b, h, sq, sk, d, causal):
    f = 4.0 * b * h * sq * sk * d
    return f / 2 if causal else f


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        pk = json.load(open(path))
        return dict(burst=pk["bf16_tflops"], sustained=pk["bf16_tflops_sustained"], hbm=pk["hbm_gbs"], src="measured")
    return dict(burst=1590.0, sustained=1400.0, hbm=6650.0, src="fallback")


class ClockSampler:
    """nvidia-smi clock / throttle-reason samples during the timed region (B200_PROFILING.md recipe)."""
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def make_batch(w, device, seed):
    dtype = torch.bfloat16 if w["dtype"] == "bf16" else torch.float16
    g = torch.Generator(device=device).manual_seed(seed)
    total = w["b"] * w["s"]
    qkv = torch.randn(total, 3, w["h"], w["d"], device=device, dtype=dtype, generator=g)
    dout = torch.randn(total, w["h"], w["d"], device=device, dtype=dtype, generator=g)
    cu = torch.arange(0, (w["b"] + 1) * w["s"], w["s"], device=device, dtype=torch.int32)
    return qkv, dout, cu


def cpu_reference_step(w, sample_b, sample_h):
    """The reference's 'PyTorch standard attention' (oracle port of attention_ref, fp32) fwd+bwd on host cores."""
    from oracle import attention_ref as orc
    g = torch.Generator().manual_seed(0)
    q, k, v = [torch.randn(sample_b, w["s"], sample_h, w["d"], generator=g).requires_grad_() for _ in range(3)]
    dout = torch.randn(sample_b, w["s"], sample_h, w["d"], generator=g)

    def step():
        keep = None
        if w["p"] > 0:
            keep = torch.rand(sample_b, sample_h, w["s"], w["s"]) >= w["p"]
        out, _ = orc.attention_ref(q, k, v, None, None, w["p"], keep, causal=w["causal"])
        torch.autograd.grad(out, (q, k, v), dout)
    return step, 3.5 * flops_fwd(sample_b, sample_h, w["s"], w["s"], w["d"], w["causal"])


def run_reference(args, w):
    """--impl reference: the reference's own CPU implementation of the path (oracle port; the reference's CUDA
    kernels do not run on cc 10.0), all host threads, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    torch.set_num_threads(os.cpu_count() or 1)
    sample_b, sample_h = 1, 2
    step, flops = cpu_reference_step(w, sample_b, sample_h)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = flops / dt / 1e12
    sample = (f"b={sample_b} h={sample_h} slice (the first sequence, the first two heads; 1/{w['b'] * w['h'] // (sample_b * sample_h)} of the "
              f"workload's (b{w['b']}, h{w['h']}) independent units -- the fp32 scores of one unit are 64 MiB, the full set would "
              f"not fit a step of seconds) per step, fp32; value = FLOPs of the slice / its time, so the ratio to the GPU arm "
              f"is a throughput ratio (FLOP-normalised), not a same-work time ratio")
    line = {"impl": "reference", "metric": "attn fwd+bwd TFLOP/s, whole job (sum over the GPUs; per GPU: config.per_gpu_tflops)",
            "value": val, "unit": "TFLOP/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": {"workload": w["name"], "sample": sample},
            "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": torch.get_num_threads(), "kind": "port",
                             "sample": sample},
            "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def ring_parity_check(world, rank, device, dist):
    """Before any ring number is reported: one small causal problem through the ring on all ranks, gathered and compared
    on rank 0 with the single-GPU kernels on the unsharded sequence (O, dQ, dK, dV).  Raises on mismatch."""
    from flash_attn import ring_attention as ring
    from flash_attn.flash_attn_interface import flash_attn_unpadded_func
    b, h, d, c = 1, 12, 128, 512
    S = 2 * world * c
    g = torch.Generator(device=device).manual_seed(4242)      # same seed on every rank: everyone holds the full problem
    q, k, v, dout = [torch.randn(b, S, h, d, device=device, dtype=torch.bfloat16, generator=g) for _ in range(4)]
    ql, kl, vl = [ring.zigzag_shard(t, world, rank).requires_grad_() for t in (q, k, v)]
    out = ring.ring_flash_attn_func(ql, kl, vl)
    grads = torch.autograd.grad(out, (ql, kl, vl), ring.zigzag_shard(dout, world, rank))
    full = []
    for t in (out.detach(),) + tuple(grads):
        parts = [torch.empty_like(t) for _ in range(world)]
        if world > 1:
            dist.all_gather(parts, t.contiguous())
        else:
            parts = [t]
        full.append(ring.zigzag_unshard(parts))
    errs = None
    if rank == 0:
        q1, k1, v1 = [t.reshape(b * S, h, d).clone().requires_grad_() for t in (q, k, v)]
        cu = torch.arange(0, (b + 1) * S, S, device=device, dtype=torch.int32)
        o1 = flash_attn_unpadded_func(q1, k1, v1, cu, cu, S, S, 0.0, causal=True)
        g1 = torch.autograd.grad(o1, (q1, k1, v1), dout.reshape(b * S, h, d))
        errs = {}
        for name, got, ref in zip(("O", "dQ", "dK", "dV"), full, (o1.detach(),) + tuple(g1)):
            ref = ref.reshape(b, S, h, d).float()
            err = (got.float() - ref).abs().max().item()
            tol = 2e-2 * max(1.0, ref.abs().max().item())      # a few 16-bit ulps of the largest entry
            errs[name] = err
            if not err <= tol:
                raise AssertionError(f"ring parity check failed at world {world}: {name} max error {err:.3e} > {tol:.3e}")
    return errs


def ring_record(args, w, world, rank, local_rank, device, dist, steps, warmup):
    """ONE 64K-token causal problem sequence-sharded (zig-zag) over the ranks: strong scaling.  -> dict (rank 0) or None."""
    import flash_attn_cuda
    from flash_attn import ring_attention as ring
    errs = ring_parity_check(world, rank, device, dist)
    b, h, d, s = w["b"], w["h"], w["d"], w["s"]
    c = s // (2 * world)
    g = torch.Generator(device=device).manual_seed(99 + rank)
    q, k, v = [torch.randn(b, 2, c, h, d, device=device, dtype=torch.bfloat16, generator=g).requires_grad_() for _ in range(3)]
    dout = torch.randn(b, 2, c, h, d, device=device, dtype=torch.bfloat16, generator=g)

    def step():
        out = ring.ring_flash_attn_func(q, k, v)
        return torch.autograd.grad(out, (q, k, v), dout)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    barrier()
    launches0 = flash_attn_cuda._lib.fa_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    barrier()
    launches = flash_attn_cuda._lib.fa_launch_count() - launches0
    t = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank != 0:
        return None
    ms_step = t.item() / steps
    value = 3.5 * flops_fwd(b, h, s, s, d, True) / (ms_step * 1e-3) / 1e12
    peaks = load_peaks()
    return {"metric": "attn fwd+bwd TFLOP/s, whole job (ONE sequence sharded over the GPUs)", "value": value, "unit": "TFLOP/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms_step, "scaling": "strong",
            "workload": w["name"], "per_gpu_tflops": value / world,
            "frac_of_bf16_peak_per_gpu": value / world / peaks["burst"],
            "kv_block_bytes_per_hop": 2 * 2 * b * c * h * d * 2, "dkv_block_bytes_per_hop": 2 * 2 * b * c * h * d * 4,
            "exchange": "K/V (bf16) and, in the backward, the fp32 dK/dV accumulator move rank->rank+1 as copy-engine pulls out of "
                        "the predecessor's CUDA-IPC-mapped buffer over NVLink, ordered by flag kernels; each transfer overlaps an "
                        "attention kernel (K/V: the same step's, dK/dV: the next step's)",
            "transport": ring.default_transport or "peer (world > 1)",
            "parity_check_max_abs_err": errs, "gpu_launches_per_step": launches / steps,
            "l2": "per-GPU inputs exceed L2 at every N <= 8 (no flush needed)"}


def run_ring(args, w):
    """--workload ring64k: the ring record as the bench line itself."""
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("nccl" if world > 1 else "gloo", rank=rank, world_size=world,
                            **({"device_id": device} if world > 1 else {}))
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    rec = ring_record(args, w, world, rank, local_rank, device, dist, args.steps, args.warmup)
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        rec.update({"higher_is_better": True, "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                    "config": {"workload": w["name"], "global_batch": w["b"], "seq_len": w["s"],
                               "parallelism": f"sequence sharded zig-zag over {world} GPU(s)"},
                    "gpu_launches": int(rec["gpu_launches_per_step"] * args.steps), "clocks": clocks})
        print(json.dumps(rec), flush=True)
    dist.destroy_process_group()


def sub_record(wname, device, steps=10, warmup=3):
    """A further BASELINE.json configuration timed like the headline (inputs resident, CUDA events on the launching
    stream, L2 flushed between steps when the inputs fit in it), reported as a sub-record of the bench line."""
    import flash_attn_cuda
    from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
    w = WORKLOADS[wname]
    qkv, dout, cu = make_batch(w, device, seed=77)
    qkv.requires_grad_()
    s, p, causal = w["s"], w["p"], w["causal"]
    f_fwd = flops_fwd(w["b"], w["h"], s, s, w["d"], causal)
    in_bytes = qkv.numel() * qkv.element_size() + dout.numel() * dout.element_size()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device) if in_bytes < 2 * 126e6 else None

    def step():
        out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=causal)
        torch.autograd.grad(out, qkv, dout)

    for _ in range(warmup):
        step()
    torch.cuda.synchronize()
    flash_attn_cuda.kernel_timing(True)
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(steps)]
    for i in range(steps):
        if flush is not None:
            flush.zero_()
        ev[i][0].record()
        step()
        ev[i][1].record()
    torch.cuda.synchronize()
    flash_attn_cuda.kernel_timing(False)
    kt = flash_attn_cuda.kernel_timing_collect()
    ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    # the same step replayed from a CUDA graph (dropout included: device-side Philox counter): what a training loop
    # that graphs its step sees, free of the Python launch path
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        step()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        step()
    gev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(steps)]
    for i in range(steps):
        if flush is not None:
            flush.zero_()
        gev[i][0].record()
        graph.replay()
        gev[i][1].record()
    torch.cuda.synchronize()
    gms = sum(a.elapsed_time(b) for a, b in gev) / steps
    peaks = load_peaks()
    kf, kb = kt["fwd_ms"] / max(kt["fwd_launches"], 1), kt["bwd_ms"] / max(kt["bwd_launches"], 1)
    return {"workload": w["name"], "dtype": w["dtype"], "steps": steps, "ms_per_step": ms, "fwd_bwd_tflops": 3.5 * f_fwd / ms / 1e9,
            "frac_of_bf16_burst_peak": 3.5 * f_fwd / ms / 1e9 / peaks["burst"],
            "graph_ms_per_step": gms, "graph_fwd_bwd_tflops": 3.5 * f_fwd / gms / 1e9,
            "fwd_kernel_ms": kf, "bwd_kernel_ms": kb, "fwd_kernel_tflops": f_fwd / kf / 1e9, "bwd_kernel_tflops": 2.5 * f_fwd / kb / 1e9,
            "l2": "flushed between steps" if flush is not None else "inputs exceed L2"}


def run_ours(args, w):
    import flash_attn_cuda
    from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    qkv, dout, cu = make_batch(w, device, seed=1234 + rank)   # every rank owns its own (batch x heads) units
    qkv.requires_grad_()
    s, p, causal = w["s"], w["p"], w["causal"]
    f_fwd = flops_fwd(w["b"], w["h"], s, s, w["d"], causal)
    f_step = 3.5 * f_fwd
    in_bytes = qkv.numel() * qkv.element_size() + dout.numel() * dout.element_size()
    # timing rule: inputs larger than L2, or L2 flushed between timed iterations (small workloads: bert, ...)
    flush = None
    l2_note = f"inputs {in_bytes / 2**20:.0f} MiB per GPU > 126 MB L2 (no flush needed)"
    if in_bytes < 2 * 126e6:
        flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
        l2_note = (f"inputs {in_bytes / 2**20:.0f} MiB per GPU fit in the 126 MB L2: a 256 MiB buffer is overwritten before "
                   "every timed step, outside the per-step CUDA events that are summed")

    def step():
        out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=causal)
        dqkv, = torch.autograd.grad(out, qkv, dout)
        return out, dqkv

    for _ in range(args.warmup):
        step()
    barrier()

    # ---- timed region: K steps, inputs resident in HBM; per-kernel-group events on the launching stream ----
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    launches0 = flash_attn_cuda._lib.fa_launch_count()
    flash_attn_cuda.kernel_timing(True)   # cudaEvents around the two main kernels, on the launching stream
    barrier()
    t_start = torch.cuda.Event(enable_timing=True)
    t_end = torch.cuda.Event(enable_timing=True)
    t_start.record()
    for i in range(args.steps):
        if flush is not None:
            flush.zero_()
        ev[i][0].record()
        out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=causal)
        ev[i][1].record()
        torch.autograd.grad(out, qkv, dout)
        ev[i][2].record()
    t_end.record()
    barrier()
    launches = flash_attn_cuda._lib.fa_launch_count() - launches0
    flash_attn_cuda.kernel_timing(False)
    kt = flash_attn_cuda.kernel_timing_collect()
    clocks = sampler.stop() if sampler else None
    ms_total = t_start.elapsed_time(t_end)
    if flush is not None:
        ms_total = sum(e[0].elapsed_time(e[2]) for e in ev)
    fwd_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps
    bwd_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps

    if args.profile:
        if rank == 0:
            print(json.dumps({"profile_run": True, "ms_per_step": ms_total / args.steps, "fwd_kernel_ms": kt["fwd_ms"] / max(kt["fwd_launches"], 1),
                              "bwd_kernel_ms": kt["bwd_ms"] / max(kt["bwd_launches"], 1)}), flush=True)
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- end-to-end: host buffers in, host buffers out, copies inside the timed region ----
    # The (batch x heads) units are independent, so the user-level call sequence splits the batch into chunks and
    # overlaps the pinned-memory H2D copy of chunk i+1 and the D2H copy of chunk i-1 with the kernels of chunk i
    # (three streams); every byte of input and output still crosses PCIe inside the timed region.
    b_tot = w["b"]
    n_chunks = next(c for c in (8, 4, 2, 1) if b_tot % c == 0)
    rows = (b_tot // n_chunks) * s
    cu_c = cu[: b_tot // n_chunks + 1].contiguous()
    qkv_h = qkv.detach().cpu().pin_memory()
    dout_h = dout.cpu().pin_memory()
    out_h = torch.empty_like(dout_h).pin_memory()
    dqkv_h = torch.empty_like(qkv_h).pin_memory()
    qkv_dc = [torch.empty_like(qkv_h[:rows], device=device).requires_grad_() for _ in range(n_chunks)]
    dout_dc = [torch.empty_like(dout_h[:rows], device=device) for _ in range(n_chunks)]
    s_in, s_out = torch.cuda.Stream(device), torch.cuda.Stream(device)
    ev_in = [torch.cuda.Event() for _ in range(n_chunks)]
    ev_done = [torch.cuda.Event() for _ in range(n_chunks)]

    def e2e_step():
        cur = torch.cuda.current_stream(device)
        for i in range(n_chunks):
            with torch.cuda.stream(s_in), torch.no_grad():
                s_in.wait_event(ev_done[i])          # the previous step is done with this chunk's device buffers
                qkv_dc[i].copy_(qkv_h[i * rows:(i + 1) * rows], non_blocking=True)
                dout_dc[i].copy_(dout_h[i * rows:(i + 1) * rows], non_blocking=True)
                ev_in[i].record(s_in)
        for i in range(n_chunks):
            cur.wait_event(ev_in[i])
            o = flash_attn_unpadded_qkvpacked_func(qkv_dc[i], cu_c, s, p, causal=causal)
            g, = torch.autograd.grad(o, qkv_dc[i], dout_dc[i])
            ev_done[i].record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_done[i])
                out_h[i * rows:(i + 1) * rows].copy_(o.detach(), non_blocking=True)
                dqkv_h[i * rows:(i + 1) * rows].copy_(g, non_blocking=True)
                o.record_stream(s_out)
                g.record_stream(s_out)
        cur.wait_stream(s_out)

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(e2e_steps):
        e2e_step()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1) / e2e_steps
    if rank == 0 and os.environ.get("FA_BENCH_CHECK_E2E"):
        torch.cuda.synchronize()
        ref_o, ref_g = step()
        assert torch.equal(out_h.to(device), ref_o.detach()), "e2e output differs from the resident-input run"
        assert torch.allclose(dqkv_h.to(device).float(), ref_g.float(), rtol=2e-2, atol=2e-2)

    times = torch.tensor([ms_total, e2e_ms], device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_step = times[0].item() / args.steps
    e2e_ms = times[1].item()

    # what the host<->device path of this box delivers to this rank with nothing else on the GPU (all ranks copy at
    # the same time): explains the end-to-end number, which is PCIe-bound
    copy_gbs = []
    for src, dst in ((qkv_h, qkv.detach()), (dqkv_h, qkv.detach())):
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        c0.record()
        for _ in range(3):
            if src is qkv_h:
                dst.copy_(src, non_blocking=True)
            else:
                src.copy_(dst, non_blocking=True)
        c1.record()
        torch.cuda.synchronize()
        copy_gbs.append(3 * qkv_h.numel() * qkv_h.element_size() / (c0.elapsed_time(c1) * 1e-3) / 1e9)
    copy_t = torch.tensor(copy_gbs, device=device, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(copy_t, op=dist.ReduceOp.SUM)
    # free the headline buffers before the other workloads
    del qkv_dc, dout_dc, qkv_h, dout_h, out_h, dqkv_h
    ring_rec = ring_record(args, WORKLOADS["ring64k"], world, rank, local_rank, device, dist, steps=max(3, min(args.steps, 5)),
                           warmup=max(args.warmup, 3))

    if rank == 0:
        peaks = load_peaks()
        long_region = times[0].item() > 1000.0
        peak = peaks["sustained"] if long_region else peaks["burst"]
        # dominant kernel = the one with the larger summed device time; its duration is the average of the
        # CUDA-event spans the library recorded around that kernel alone during the timed region
        kf_ms = kt["fwd_ms"] / max(kt["fwd_launches"], 1)
        kb_ms = kt["bwd_ms"] / max(kt["bwd_launches"], 1)
        dom = "bwd" if kb_ms >= kf_ms else "fwd"
        dom_flops = 2.5 * f_fwd if dom == "bwd" else f_fwd
        dom_ms = kb_ms if dom == "bwd" else kf_ms
        achieved = dom_flops / (dom_ms * 1e-3) / 1e12
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
                traffic = json.load(fh).get(args.workload, {}).get(dom)
        except (OSError, ValueError):
            pass
        value = world * f_step / (ms_step * 1e-3) / 1e12
        line = {
            "metric": "attn fwd+bwd TFLOP/s, whole job (sum over the GPUs; per GPU: config.per_gpu_tflops)", "value": value,
            "unit": "TFLOP/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": w["dtype"], "data": "synthetic",
            "config": {"workload": w["name"], "global_batch": w["b"] * world, "seq_len": s,
                       "parallelism": f"(batch x heads) units sharded over {world} GPU(s), no collective",
                       "l2": l2_note,
                       "per_gpu_tflops": value / world, "fwd_ms": fwd_ms, "bwd_ms": bwd_ms,
                       "fwd_tflops": f_fwd / (fwd_ms * 1e-3) / 1e12, "bwd_tflops": 2.5 * f_fwd / (bwd_ms * 1e-3) / 1e12,
                       "fwd_kernel_ms": kf_ms, "bwd_kernel_ms": kb_ms,
                       "fwd_kernel_tflops": f_fwd / (kf_ms * 1e-3) / 1e12,
                       "bwd_kernel_tflops": 2.5 * f_fwd / (kb_ms * 1e-3) / 1e12,
                       "frac_of_bf16_peak_fwd_bwd": value / world / peak,
                       # the same number against the sustained cuBLAS rate and the nominal dense bf16 peak
                       "frac_of_sustained_bf16": value / world / peaks["sustained"],
                       "frac_of_nominal_2250": value / world / 2250.0},
            "roofline": {"bound": "tensor", "kernel": f"{dom}_sm100_kernel (tcgen05/TMA), algorithmic FLOPs "
                                                         f"{dom_flops / 1e12:.3f} T per launch, {dom_ms:.3f} ms per launch",
                         "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                         "peak_source": f"{peaks['src']} {'sustained' if long_region else 'burst'} bf16 matmul",
                         "traffic": traffic},
            "e2e": {"value": world * f_step / (e2e_ms * 1e-3) / 1e12, "unit": "TFLOP/s", "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": in_bytes, "d2h_bytes_per_step": in_bytes,
                    "pipeline": f"{n_chunks} batch chunks; H2D, kernels and D2H on three streams",
                    "host_link_gbs_all_ranks": {"h2d": copy_t[0].item(), "d2h": copy_t[1].item(),
                                                "note": "pinned-memory copies alone, all ranks at once, summed over ranks"}},
            "gpu_launches": int(launches), "clocks": clocks,
            "ring64k": ring_rec,
        }
        if world == 1:
            line["sub_records"] = {"d64_dropout": sub_record("d64", device), "bert_s512": sub_record("bert", device)}
        if world == 1 and not args.no_cpu_baseline:
            torch.set_num_threads(os.cpu_count() or 1)
            sb, sh = 1, 2
            cstep, cflops = cpu_reference_step(w, sb, sh)
            cstep()
            best = 1e30
            t_all = time.perf_counter()
            while True:
                t0 = time.perf_counter()
                cstep()
                best = min(best, time.perf_counter() - t0)
                if time.perf_counter() - t_all > 10.0:
                    break
            line["cpu_baseline"] = {"value": cflops / best / 1e12, "unit": "TFLOP/s", "cores": torch.get_num_threads(),
                                    "kind": "port",
                                    "sample": f"b={sb} h={sh} slice of the workload, fp32 attention_ref fwd+bwd, best step in 10 s"}
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def _clock_snapshot():
    """one nvidia-smi sample (SM clock, max, active throttle reasons): recorded with every sweep point, because the same
    binary differs by several percent between boxes of the pool and with the power state"""
    try:
        out = subprocess.run(["nvidia-smi", "-i", "0", f"--query-gpu={ClockSampler.FIELDS}", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=10).stdout.strip().split(",")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]),
                "reasons": [n for n, v in zip(names, out[2:6]) if v.strip().lower().startswith("active")]}
    except Exception:      # noqa: BLE001
        return None


def run_sweep(args):
    """fwd / bwd / fwd+bwd TFLOP/s vs seqlen for both BASELINE shapes (one line per point, to stdout)."""
    from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
    device = torch.device("cuda", 0)
    peaks = load_peaks()
    pts = [("d64", dict(b=8, h=12, d=64, dtype="fp16", p=0.1), [512, 1024, 2048, 4096]),
           ("d64_p0", dict(b=8, h=12, d=64, dtype="fp16", p=0.0), [512, 1024, 2048, 4096]),
           ("d128", dict(b=16, h=12, d=128, dtype="bf16", p=0.0), [512, 1024, 2048, 4096, 8192, 16384])]
    for tag, base, seqs in pts:
        for s in seqs:
            if s > args.max_seqlen:
                continue
            for causal in (False, True):
                w = dict(base, s=s, causal=causal)
                qkv, dout, cu = make_batch(w, device, 0)
                qkv.requires_grad_()
                f = flops_fwd(w["b"], w["h"], s, s, w["d"], causal)

                def fwd_only():
                    return flash_attn_unpadded_qkvpacked_func(qkv, cu, s, w["p"], causal=causal)

                def fwd_bwd():
                    torch.autograd.grad(fwd_only(), qkv, dout)

                def timed(fn, reps):
                    """`reps` calls enqueued back to back (no host sync in between: launch latency overlaps device work)."""
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    torch.cuda.synchronize()
                    e0.record()
                    for _ in range(reps):
                        fn()
                    e1.record()
                    torch.cuda.synchronize()
                    return e0.elapsed_time(e1) / reps

                for _ in range(3):
                    fwd_bwd()
                est = timed(fwd_bwd, 2)
                reps = int(min(400, max(10, 150.0 / max(est, 1e-3))))     # ~150 ms of device time per measurement
                tf = timed(fwd_only, reps)
                tfb = tf if args.fwd_only else timed(fwd_bwd, reps)
                tb = max(tfb - tf, 1e-9)
                row = {"sweep": tag, "seqlen": s, "causal": causal, "dropout": w["p"], "reps": reps, "clocks": _clock_snapshot(),
                       "fwd_ms": tf, "bwd_ms": tb, "fwd_tflops": f / tf / 1e9, "bwd_tflops": 2.5 * f / tb / 1e9,
                       "fwd_bwd_tflops": 3.5 * f / tfb / 1e9, "frac_burst_peak": 3.5 * f / tfb / 1e9 / peaks["burst"]}
                if not args.fwd_only:
                    # the same step replayed from a CUDA graph (dropout included: device-side Philox counter)
                    side = torch.cuda.Stream()
                    side.wait_stream(torch.cuda.current_stream())
                    with torch.cuda.stream(side):
                        fwd_bwd()
                    torch.cuda.current_stream().wait_stream(side)
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph):
                        fwd_bwd()
                    tg = timed(graph.replay, reps)
                    row["graph_fwd_bwd_ms"] = tg
                    row["graph_fwd_bwd_tflops"] = 3.5 * f / tg / 1e9
                    del graph
                print(json.dumps(row), flush=True)
                del qkv, dout


def _sweep_time(fwd, leaves, dout, graph_ok):
    """fwd / fwd+bwd time of one point: calls enqueued back to back, optional CUDA-graph replay of the step."""
    def fwd_bwd():
        torch.autograd.grad(fwd(), leaves, dout)

    def timed(fn, reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    for _ in range(3):
        fwd_bwd()
    reps = int(min(400, max(10, 150.0 / max(timed(fwd_bwd, 2), 1e-3))))
    tf, tfb, tg = timed(fwd, reps), timed(fwd_bwd, reps), None
    if graph_ok:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            fwd_bwd()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            fwd_bwd()
        tg = timed(graph.replay, reps)
    return reps, tf, tfb, tg


def run_sweep_extra(args):
    """The remaining measurement inputs of SURVEY 8(d): C4 (varlen BERT batch, 32K tokens of mixed lengths 64-512,
    d = 64 / 80 / 96, and cross-attention with independent q / k lengths) and C5's blocksparse butterfly mask."""
    from flash_attn.flash_attn_interface import flash_attn_unpadded_kvpacked_func, flash_attn_unpadded_qkvpacked_func
    from flash_attn.flash_blocksparse_attn_interface import convert_blockmask, flash_blocksparse_attn_func
    from flash_attn.flash_blocksparse_attention import ButterflySparsityConfig
    device = torch.device("cuda", 0)
    peaks = load_peaks()
    h = 12

    def draw_lengths(seed, total=32768):
        g = torch.Generator().manual_seed(seed)
        out, acc = [], 0
        while acc < total:
            n = min(int(torch.randint(64, 513, (1,), generator=g)), total - acc)
            out.append(n)
            acc += n
        return torch.tensor(out, dtype=torch.int64)

    def emit(row, f, reps, tf, tfb, tg):
        tb = max(tfb - tf, 1e-9)
        row.update({"reps": reps, "clocks": _clock_snapshot(), "fwd_ms": tf, "bwd_ms": tb, "fwd_tflops": f / tf / 1e9, "bwd_tflops": 2.5 * f / tb / 1e9,
                    "fwd_bwd_tflops": 3.5 * f / tfb / 1e9, "frac_burst_peak": 3.5 * f / tfb / 1e9 / peaks["burst"]})
        if tg is not None:
            row.update({"graph_fwd_bwd_ms": tg, "graph_fwd_bwd_tflops": 3.5 * f / tg / 1e9})
        print(json.dumps(row), flush=True)

    lens_q, lens_k = draw_lengths(0), draw_lengths(1)
    cu_q = torch.nn.functional.pad(lens_q.cumsum(0), (1, 0)).to(device=device, dtype=torch.int32)
    for d in (64, 80, 96):
        for causal in (False, True):
            qkv = torch.randn(32768, 3, h, d, device=device, dtype=torch.float16, requires_grad=True)
            dout = torch.randn(32768, h, d, device=device, dtype=torch.float16)
            mx = int(lens_q.max())
            f = 4.0 * h * d * float((lens_q * lens_q).sum()) / (2 if causal else 1)
            r = _sweep_time(lambda: flash_attn_unpadded_qkvpacked_func(qkv, cu_q, mx, 0.0, causal=causal), qkv, dout, True)
            emit({"sweep": "varlen32k", "d": d, "seqs": len(lens_q), "causal": causal, "dropout": 0.0}, f, *r)
    # cross-attention: the same number of sequences, independent q / k lengths (tests/test_flash_attn.py:446-447)
    nseq = min(len(lens_q), len(lens_k))
    lq, lk = lens_q[:nseq], lens_k[:nseq]
    cq = torch.nn.functional.pad(lq.cumsum(0), (1, 0)).to(device=device, dtype=torch.int32)
    ck = torch.nn.functional.pad(lk.cumsum(0), (1, 0)).to(device=device, dtype=torch.int32)
    for d in (64, 96):
        q = torch.randn(int(lq.sum()), h, d, device=device, dtype=torch.float16, requires_grad=True)
        kv = torch.randn(int(lk.sum()), 2, h, d, device=device, dtype=torch.float16, requires_grad=True)
        dout = torch.randn_like(q)
        f = 4.0 * h * d * float((lq * lk).sum())
        r = _sweep_time(lambda: flash_attn_unpadded_kvpacked_func(q, kv, cq, ck, int(lq.max()), int(lk.max()), 0.0), (q, kv), dout, True)
        emit({"sweep": "cross_varlen", "d": d, "seqs": nseq, "causal": False, "dropout": 0.0}, f, *r)
    # blocksparse, butterfly layout at the kernel's (16 x 256) granularity, d = 64 fp16
    for s, b in ((4096, 8), (16384, 2)):
        layout = ButterflySparsityConfig().make_layout(s).to(device)
        mask = convert_blockmask(layout, causal=False)
        density = float(layout.float().mean())
        qkv = torch.randn(b * s, 3, h, 64, device=device, dtype=torch.float16, requires_grad=True)
        dout = torch.randn(b * s, h, 64, device=device, dtype=torch.float16)
        cu = torch.arange(0, (b + 1) * s, s, device=device, dtype=torch.int32)
        f = 4.0 * b * h * s * s * 64 * density
        r = _sweep_time(lambda: flash_blocksparse_attn_func(qkv, cu, mask, 0.0, s, convert_mask=False), qkv, dout, True)
        emit({"sweep": "blocksparse_butterfly", "d": 64, "seqlen": s, "b": b, "density": density, "causal": False,
              "dropout": 0.0}, f, *r)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="d128", choices=sorted(WORKLOADS))
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--sweep-extra", action="store_true", help="only the varlen / cross-attention / blocksparse points of the sweep")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--profile", action="store_true", help="runs under ncu: the headline steps only (no e2e / ring / sub-records / CPU legs; "
                                                           "the line it prints is not a bench value)")
    ap.add_argument("--fwd-only", action="store_true", help="sweep: time the forward only")
    ap.add_argument("--max-seqlen", type=int, default=16384, help="sweep: largest seqlen")
    args = ap.parse_args()
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference(args, w)
    elif args.sweep_extra:
        run_sweep_extra(args)
    elif args.sweep:
        run_sweep(args)
        run_sweep_extra(args)
    elif w.get("ring"):
        run_ring(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
