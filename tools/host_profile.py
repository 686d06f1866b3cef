"""cProfile of the host path of one forward+backward at a launch-bound size (b8 h12 d64 S512): where the Python time goes."""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import time
import torch
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func

b, s, h, d = 8, 512, 12, 64
qkv = torch.randn(b * s, 3, h, d, device="cuda", dtype=torch.float16, requires_grad=True)
dout = torch.randn(b * s, h, d, device="cuda", dtype=torch.float16)
cu = torch.arange(0, (b + 1) * s, s, device="cuda", dtype=torch.int32)


def step():
    out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, 0.0, causal=False)
    torch.autograd.grad(out, qkv, dout)


for _ in range(20):
    step()
torch.cuda.synchronize()
n = 2000
t0 = time.perf_counter()
for _ in range(n):
    step()
t1 = time.perf_counter()
torch.cuda.synchronize()
print(f"host time per fwd+bwd (enqueue only): {(t1 - t0) / n * 1e6:.1f} us; with the device drained: {(time.perf_counter() - t0) / n * 1e6:.1f} us")
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    step()
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("tottime").print_stats(22)
