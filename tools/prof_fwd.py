"""Launches the forward (and optionally backward) kernels a few times on one fixed config: the command ncu wraps."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch  # noqa: E402
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func  # noqa: E402

b, h, d, s = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (16, 12, 128, 4096))]
causal = (sys.argv[5] == "1") if len(sys.argv) > 5 else True
bwd = (sys.argv[6] == "1") if len(sys.argv) > 6 else False
p = float(sys.argv[7]) if len(sys.argv) > 7 else 0.0
dtype = torch.bfloat16
qkv = torch.randn(b * s, 3, h, d, device="cuda", dtype=dtype, requires_grad=bwd)
cu = torch.arange(0, (b + 1) * s, s, device="cuda", dtype=torch.int32)
for _ in range(4):
    out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=causal)
    if bwd:
        torch.autograd.grad(out, qkv, torch.ones_like(out))
torch.cuda.synchronize()
print("done")
