"""Timeline of one backward CTA (needs `make -C hazyresearch_flash-attention_b200/csrc trace`).
   FA_B200_LIB=.../libfa_b200_trace.so python tools/trace_bwd.py [b h d s causal]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
os.environ.setdefault("FA_B200_LIB", os.path.join(ROOT, "hazyresearch_flash-attention_b200", "csrc", "libfa_b200_trace.so"))
import torch
import flash_attn_cuda
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func

b, h, d, s = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (16, 12, 128, 4096))]
causal = (sys.argv[5] == "1") if len(sys.argv) > 5 else False
qkv = torch.randn(b * s, 3, h, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
cu = torch.arange(0, (b + 1) * s, s, device="cuda", dtype=torch.int32)
for _ in range(2):
    out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, 0.0, causal=causal)
    torch.autograd.grad(out, qkv, torch.ones_like(out))
torch.cuda.synchronize()
n = 64 * 32
buf = (ctypes.c_longlong * n)()
flash_attn_cuda._lib.fa_debug_read_bwd_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
rc = flash_attn_cuda._lib.fa_debug_read_bwd_trace(buf, n)
assert rc == 0, rc
T = [[buf[i * 32 + j] for j in range(32)] for i in range(64)]
t0 = min(x for row in T[:4] for x in row if x > 0)
names = {6: "M:dq issued", 7: "M:dq committed", 0: "M:ds_full", 1: "M:dq,dk issued", 2: "M:p_full(i+1)", 3: "M:dq_empty", 4: "M:dP(i+1) issued", 5: "M:S(i+2) issued",
         8: "C0:s_full", 9: "C0:P done", 10: "C0:dp_full", 11: "C0:dS done", 12: "C1:s_full", 13: "C1:P done", 14: "C1:dp_full",
         15: "C1:dS done", 16: "D:dq_full", 17: "D:dq_empty arr", 18: "D:tile done", 20: "L:q_empty", 21: "L:do_empty"}
for it in range(2, 10):
    ev = sorted((T[it][j] - t0, names[j]) for j in names if T[it][j] > 0)
    print(f"--- iteration {it}")
    for t, nme in ev:
        print(f"   {t:8d}  {nme}")
per = [T[i + 1][0] - T[i][0] for i in range(4, 24) if T[i + 1][0] > 0 and T[i][0] > 0]
print("period (M:ds_full to next):", per)
# ---- per-CTA lifetimes ----
ncta = min(8192, (s // 128) * h * b)
cb = (ctypes.c_longlong * (ncta * 12))()
flash_attn_cuda._lib.fa_debug_read_bwd_cta.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert flash_attn_cuda._lib.fa_debug_read_bwd_cta(cb, ncta * 12) == 0
import statistics
C = [[cb[i * 12 + j] for j in range(12)] for i in range(ncta)]
C = [c for c in C if c[5] > 0]
print("CTAs", len(C))
print("median clk: setup %d  loop %d  epilogue %d  total %d" % (statistics.median(c[3] - c[2] for c in C), statistics.median(c[4] - c[3] for c in C),
      statistics.median(c[5] - c[4] for c in C), statistics.median(c[5] - c[2] for c in C)))
print("epilogue median clk after loop end: dkv_full +%d  dK/dV stored +%d  drain last issue +%d  drain reads done +%d  CTA end +%d" % tuple(
    statistics.median(c[j] - c[4] for c in C) for j in (7, 8, 9, 10, 5)))
by_sm = {}
for c in C:
    by_sm.setdefault(c[0], []).append(c)
gaps, lives = [], []
for sm, l in by_sm.items():
    l.sort(key=lambda c: c[1])
    for a, bb in zip(l, l[1:]):
        gaps.append(bb[1] - a[6])
    lives += [c[6] - c[1] for c in l]
print("per-SM: CTAs/SM %.1f  median CTA life %.1f us  median gap between CTAs %.2f us  max gap %.2f us" %
      (len(C) / len(by_sm), statistics.median(lives) / 1e3, statistics.median(gaps) / 1e3, max(gaps) / 1e3))
t_first = min(c[1] for c in C); t_last = max(c[6] for c in C)
print("kernel span %.1f us; sum of CTA lives / SMs = %.1f us" % ((t_last - t_first) / 1e3, sum(lives) / len(by_sm) / 1e3))
