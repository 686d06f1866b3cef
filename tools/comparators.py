"""On-box peers for the headline configuration (b16 h12 d128 S4096 causal bf16 fwd+bwd): what ALREADY ships on this image.
Context only -- none of this is imported by the product, and nothing of this repository is imported here (the upstream
wheel is also called `flash_attn`, so this script must run WITHOUT the repo's package directory on sys.path).
    python tools/comparators.py > profiles/r02_comparators.jsonl"""
import json
import subprocess
import sys

import torch
import torch.nn.functional as F

b, h, d, s, causal = 16, 12, 128, 4096, True
f_fwd = 4.0 * b * h * s * s * d / 2
dev = "cuda"


def clocks():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.sw_power_cap", "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=10).stdout.strip().split(",")
        return {"sm_mhz": float(out[0]), "sm_max_mhz": float(out[1]), "sw_power_cap": out[2].strip()}
    except Exception:      # noqa: BLE001
        return None


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def report(name, fwd, fwd_bwd, note):
    row = {"comparator": name, "config": f"b{b} h{h} d{d} S{s} causal bf16", "note": note}
    try:
        tf = timed(fwd)
        tfb = timed(fwd_bwd)
        row.update(fwd_ms=tf, fwd_bwd_ms=tfb, fwd_tflops=f_fwd / tf / 1e9, bwd_tflops=2.5 * f_fwd / max(tfb - tf, 1e-9) / 1e9,
                   fwd_bwd_tflops=3.5 * f_fwd / tfb / 1e9, clocks=clocks())
    except Exception as e:      # noqa: BLE001
        row["unavailable"] = f"{type(e).__name__}: {str(e)[:200]}"
    print(json.dumps(row), flush=True)


# --- torch SDPA, cuDNN and flash backends: (b, h, s, d) layout ---
q, k, v = [torch.randn(b, h, s, d, device=dev, dtype=torch.bfloat16, requires_grad=True) for _ in range(3)]
g = torch.randn(b, h, s, d, device=dev, dtype=torch.bfloat16)
from torch.nn.attention import SDPBackend, sdpa_kernel
for name, backend in (("torch SDPA / cuDNN backend", SDPBackend.CUDNN_ATTENTION), ("torch SDPA / flash backend (sm80-style mma.sync kernels)", SDPBackend.FLASH_ATTENTION)):
    def fwd():
        with sdpa_kernel(backend):
            return F.scaled_dot_product_attention(q, k, v, is_causal=causal)

    def fwd_bwd():
        torch.autograd.grad(fwd(), (q, k, v), g)
    report(name, fwd, fwd_bwd, f"torch {torch.__version__}")
del q, k, v, g

# --- upstream flash-attn wheel: (b, s, h, d) layout ---
try:
    import flash_attn as up
    assert "site-packages" in up.__file__, f"not the upstream wheel: {up.__file__}"
    q, k, v = [torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16, requires_grad=True) for _ in range(3)]
    g = torch.randn(b, s, h, d, device=dev, dtype=torch.bfloat16)
    report(f"upstream flash-attn {up.__version__} wheel", lambda: up.flash_attn_func(q, k, v, causal=causal),
           lambda: torch.autograd.grad(up.flash_attn_func(q, k, v, causal=causal), (q, k, v), g), up.__file__)
except Exception as e:      # noqa: BLE001
    print(json.dumps({"comparator": "upstream flash-attn wheel", "unavailable": f"{type(e).__name__}: {str(e)[:200]}"}), flush=True)
