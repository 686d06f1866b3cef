"""Which operator launches the FillFunctor kernel in a FlashMHA training step?  (torch profiler, python stacks)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
from torch.profiler import ProfilerActivity, profile
from flash_attn.flash_attention import FlashMHA

torch.manual_seed(0)
b, s, e, h = 8, 1024, 1024, 16
mha = FlashMHA(e, h, causal=True, use_rotary_emb="1d", device="cuda", dtype=torch.bfloat16)
x = torch.randn(b, s, e, device="cuda", dtype=torch.bfloat16, requires_grad=True)
g = torch.randn(b, s, e, device="cuda", dtype=torch.bfloat16)
for _ in range(2):
    y, _ = mha(x)
    y.backward(g)
    mha.zero_grad(set_to_none=True)
    x.grad = None
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
    y, _ = mha(x)
    y.backward(g)
    torch.cuda.synchronize()
for ev in prof.events():
    if ev.name in ("aten::zero_", "aten::fill_", "aten::zeros", "aten::zeros_like", "aten::new_zeros"):
        print(ev.name, ev.input_shapes, "cuda_us", getattr(ev, "device_time", getattr(ev, "cuda_time", None)))
        for fr in (ev.stack or [])[:8]:
            print("     ", fr)
        p = ev.cpu_parent
        chain = []
        while p is not None and len(chain) < 6:
            chain.append(p.name)
            p = p.cpu_parent
        print("      parents:", " <- ".join(chain))
