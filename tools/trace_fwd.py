"""Timeline of one forward CTA (needs `make -C hazyresearch_flash-attention_b200/csrc trace`).
   FA_B200_LIB=.../libfa_b200_trace.so python tools/trace_fwd.py [b h d s causal]   (traces the last query block of (0, 0))"""
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
dtype = torch.bfloat16 if d > 64 else torch.float16
qkv = torch.randn(b * s, 3, h, d, device="cuda", dtype=dtype)
cu = torch.arange(0, (b + 1) * s, s, device="cuda", dtype=torch.int32)
for _ in range(3):
    flash_attn_unpadded_qkvpacked_func(qkv, cu, s, 0.0, causal=causal)
torch.cuda.synchronize()
n = 64 * 32
buf = (ctypes.c_longlong * n)()
flash_attn_cuda._lib.fa_debug_read_fwd_trace.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert flash_attn_cuda._lib.fa_debug_read_fwd_trace(buf, n) == 0
T = [[buf[i * 32 + j] for j in range(32)] for i in range(64)]
t0 = min(x for row in T[:2] for x in row if x > 0)
names = {0: "A:s_full", 1: "A:S loaded", 2: "A:max done", 3: "A:turn", 4: "A:exp done", 5: "A:P handed",
         8: "B:s_full", 9: "B:S loaded", 10: "B:max done", 11: "B:turn", 12: "B:exp done", 13: "B:P handed",
         16: "M:A p0", 17: "M:A PV issued", 18: "M:A S(u+1) issued", 20: "M:B p0", 21: "M:B PV issued", 22: "M:B S(u+1) issued"}
for u in range(6, 12):
    ev = sorted((T[u][j] - t0, names[j]) for j in names if T[u][j] > 0)
    print(f"--- K block {u}")
    for t, nme in ev:
        print(f"   {t:8d}  {nme}")
per = [T[u + 1][0] - T[u][0] for u in range(4, 28) if T[u + 1][0] > 0 and T[u][0] > 0]
print("period (A:s_full to next):", per)
for nm, a, bb in (("ld", 0, 1), ("max+rescale", 1, 2), ("turn wait", 2, 3), ("exp", 3, 4), ("st+handoff", 4, 5)):
    vals = [T[u][bb] - T[u][a] for u in range(4, 28) if T[u][bb] > 0 and T[u][a] > 0]
    if vals:
        print(f"A {nm}: median {sorted(vals)[len(vals) // 2]}")
vals = [T[u + 1][0] - T[u][5] for u in range(4, 28) if T[u + 1][0] > 0 and T[u][5] > 0]
print("A P handed -> next s_full: median", sorted(vals)[len(vals) // 2])

# ---- per-CTA lifetimes ----
import statistics
ncta = min(8192, ((s + 255) // 256) * h * b)
cb = (ctypes.c_longlong * (ncta * 8))()
flash_attn_cuda._lib.fa_debug_read_fwd_cta.argtypes = [ctypes.c_void_p, ctypes.c_int]
assert flash_attn_cuda._lib.fa_debug_read_fwd_cta(cb, ncta * 8) == 0
C = [[cb[i * 8 + j] for j in range(8)] for i in range(ncta)]
C = [c for c in C if c[5] > 0 and c[4] > 0 and c[7] > 0]
print("CTAs", len(C))
print("median clk: setup %d  loop %d (%.0f per K block, median %d blocks)  epilogue %d  total %d" % (
    statistics.median(c[3] - c[2] for c in C), statistics.median(c[4] - c[3] for c in C),
    statistics.median((c[4] - c[3]) / c[7] for c in C), statistics.median(c[7] for c in C),
    statistics.median(c[5] - c[4] for c in C), statistics.median(c[5] - c[2] for c in C)))
# fixed cost per CTA from a linear fit of loop time against the number of K blocks
xs, ys = [c[7] for c in C], [c[5] - c[2] for c in C]
if len(set(xs)) > 1:
    mx, my = sum(xs) / len(xs), sum(ys) / len(ys)
    slope = sum((x - mx) * (y - my) for x, y in zip(xs, ys)) / sum((x - mx) ** 2 for x in xs)
    print("linear fit: CTA clk = %.0f + %.0f * K blocks" % (my - slope * mx, slope))
by_sm = {}
for c in C:
    by_sm.setdefault(c[0], []).append(c)
gaps = []
for sm, l in by_sm.items():
    l.sort(key=lambda c: c[1])
    for a, bb in zip(l, l[1:]):
        gaps.append(bb[1] - a[6])
if gaps:
    print("per-SM: CTAs/SM %.1f  median gap between CTAs %.2f us  max %.2f us" % (len(C) / len(by_sm), statistics.median(gaps) / 1e3, max(gaps) / 1e3))
