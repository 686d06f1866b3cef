"""One FlashMHA training step (Wqkv -> rotary -> attention -> out_proj, forward + backward), for a kernel launch list:
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file profiles/...csv python tools/flashmha_step.py
SURVEY 8 (f3): the (b, s, 3*h*d) GEMM output is consumed in place through strided TMA views and the (T, h, d) attention output is
what out_proj reads -- the list must show GEMMs, one rotary pass each way and the attention kernels, and no copy / cat / stack."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
from flash_attn.flash_attention import FlashMHA

torch.manual_seed(0)
b, s, e, h = 8, 1024, 1024, 16
mha = FlashMHA(e, h, causal=True, use_rotary_emb="1d", device="cuda", dtype=torch.bfloat16)
x = torch.randn(b, s, e, device="cuda", dtype=torch.bfloat16, requires_grad=True)
g = torch.randn(b, s, e, device="cuda", dtype=torch.bfloat16)
for it in range(3):          # the last iteration is the one to read in the list
    torch.cuda.nvtx.range_push(f"step{it}")
    y, _ = mha(x)
    y.backward(g)
    mha.zero_grad(set_to_none=True)
    x.grad = None
    torch.cuda.nvtx.range_pop()
torch.cuda.synchronize()
print("ok")
