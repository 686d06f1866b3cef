"""Kernel-level timing of several configurations in one process.
   python tools/time_many.py "b h d s causal p dtype [ENV=VAL ...]" ...   (each argument is one configuration)"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
import flash_attn_cuda
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func


def run(cfg, reps=10):
    tok = cfg.split()
    b, h, d, s = [int(x) for x in tok[:4]]
    causal, p = tok[4] == "1", float(tok[5])
    dtype = torch.bfloat16 if tok[6] == "bf16" else torch.float16
    env = dict(t.split("=", 1) for t in tok[7:])
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
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
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    try:
        import subprocess
        smi = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,temperature.gpu", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=10).stdout.strip().replace(", ", "/")
    except Exception:
        smi = "?"
    f = 4.0 * b * h * s * s * d / (2 if causal else 1)
    kf, kb = kt["fwd_ms"] / kt["fwd_launches"], kt["bwd_ms"] / kt["bwd_launches"]
    print(f"{cfg}: fwd {tf/reps:.3f} ms (kernel {kf:.3f} ms, {f/kf/1e9:.0f} TF/s)  "
          f"bwd {tb/reps:.3f} ms (kernel {kb:.3f} ms, {2.5*f/kb/1e9:.0f} TF/s)  fwd+bwd {3.5*f/((tf+tb)/reps)/1e9:.0f} TF/s  [idle-after MHz/W/C {smi}]", flush=True)
    del qkv, dout, out
    torch.cuda.empty_cache()


import time
for attempt in range(5):        # a fresh box occasionally refuses the first CUDA initialisation
    try:
        torch.zeros(1, device="cuda")
        break
    except RuntimeError as e:
        print("CUDA init failed, retrying:", e, flush=True)
        time.sleep(5)
for c in sys.argv[1:]:
    run(c)
