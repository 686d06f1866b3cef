"""Two small shapes through forward + backward, for compute-sanitizer (memcheck / racecheck):
    compute-sanitizer --tool memcheck python tools/sanitize_case.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
from flash_attn.flash_attn_interface import flash_attn_unpadded_func

torch.manual_seed(0)
for (lens_q, lens_k, h, d, dtype, causal, p) in (([200, 77], [200, 130], 2, 128, torch.bfloat16, True, 0.0),
                                                 ([300, 129, 64], [300, 129, 64], 2, 64, torch.float16, False, 0.1)):
    cu_q = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), device="cuda", dtype=torch.int32)
    cu_k = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), device="cuda", dtype=torch.int32)
    q = torch.randn(sum(lens_q), h, d, device="cuda", dtype=dtype, requires_grad=True)
    k = torch.randn(sum(lens_k), h, d, device="cuda", dtype=dtype, requires_grad=True)
    v = torch.randn(sum(lens_k), h, d, device="cuda", dtype=dtype, requires_grad=True)
    out = flash_attn_unpadded_func(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), p, causal=causal)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), torch.randn_like(out))
    torch.cuda.synchronize()
    assert all(torch.isfinite(t.float()).all() for t in (out, dq, dk, dv))
    print(f"ok: d={d} causal={causal} p={p} |O|max={out.abs().max().item():.3f} |dQ|max={dq.abs().max().item():.3f}", flush=True)
