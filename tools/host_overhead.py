"""Host-side cost of one forward / backward call at a size where the kernels are negligible (cProfile)."""
import cProfile, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func

b, h, d, s = 2, 4, 64, 128
p = float(sys.argv[1]) if len(sys.argv) > 1 else 0.0
qkv = torch.randn(b * s, 3, h, d, device="cuda", dtype=torch.float16, requires_grad=True)
dout = torch.randn(b * s, h, d, device="cuda", dtype=torch.float16)
cu = torch.arange(0, (b + 1) * s, s, device="cuda", dtype=torch.int32)
def it():
    out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=True)
    torch.autograd.grad(out, qkv, dout)
for _ in range(20):
    it()
torch.cuda.synchronize()
n = 2000
t0 = time.perf_counter()
for _ in range(n):
    it()
torch.cuda.synchronize()
print(f"fwd+bwd host-bound time per iteration: {(time.perf_counter() - t0) / n * 1e6:.1f} us (p={p})")
pr = cProfile.Profile()
pr.enable()
for _ in range(n):
    it()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
