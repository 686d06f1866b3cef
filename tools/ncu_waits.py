"""Summarise an `ncu --page source --csv` dump of the backward kernel: samples per mbarrier wait and per role."""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[1], rows[2:]
iS, iSrc = hdr.index("# Samples"), hdr.index("Source")
names = ['k_full', 'v_full', 'q_full0', 'q_full1', 'q_empty0', 'q_empty1', 'do_full', 'do_empty', 's_full', 'dp_full',
         'p_full', 'ds_full', 'ds_empty', 'dq_full', 'dq_empty', 'dkv_full']
base = None
tot = sum(int(r[iS] or 0) for r in data)
print("total samples", tot)
agg = {}
for i, r in enumerate(data):
    m = re.search(r'TRYWAIT P\d, \[(\S+?)\+?(0x[0-9a-f]+)\]', r[iSrc])
    if m:
        off = int(m.group(2), 16)
        if base is None or off < base:
            base = off if base is None else min(base, off)
offs = []
for i, r in enumerate(data):
    m = re.search(r'TRYWAIT P\d, \[(\S+?)\+?(0x[0-9a-f]+)\]', r[iSrc])
    if m:
        offs.append((i, int(m.group(2), 16)))
base = min(o for _, o in offs) & ~0x7f
for i, off in offs:
    idx = (off - base) // 8
    n = names[idx] if 0 <= idx < 16 else hex(off)
    s = int(data[i][iS] or 0) + int(data[i + 1][iS] or 0)
    agg.setdefault(n, []).append((i, s))
for n, l in agg.items():
    print(f"{n:10s} total {sum(s for _, s in l):7d}   " + " ".join(f"@{i}:{s}" for i, s in l if s > 50))
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 400
print("--- lines with >= %d samples" % thr)
for i, r in enumerate(data):
    s = int(r[iS] or 0)
    if s >= thr:
        print(i, s, r[iSrc][:100])
