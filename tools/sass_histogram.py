"""SASS opcode histogram of the built library (profiles/r02_sass_histogram.txt): python tools/sass_histogram.py [lib.so]"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "hazyresearch_flash-attention_b200", "csrc", "libfa_b200.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
ops = collections.Counter()
n_kernels = 0
for line in sass.splitlines():
    if re.match(r"\s*Function : ", line):
        n_kernels += 1
    m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m:
        ops[m.group(1)] += 1
print(f"# SASS opcode histogram of csrc/libfa_b200.so (cuobjdump -sass, all {n_kernels} kernels; tools/sass_histogram.py).  The tcgen05 / TMA /")
print("# TMEM opcodes that prove the hot path is Blackwell-native (B200_PROFILING.md): UTCHMMA = tcgen05.mma, UTMALDG =")
print("# cp.async.bulk.tensor, UBLKCP / UBLKRED = cp.async.bulk / cp.reduce.async.bulk, LDTM / STTM = tcgen05.ld / st.  HMMA (mma.sync) must be 0.")
print()
for fam in ("UTCHMMA", "UTMALDG", "UBLKCP", "UBLKRED", "LDTM", "STTM", "UTCBAR", "FFMA2", "FADD2", "MUFU", "HMMA", "SYNCS", "ATOM", "RED", "LDG", "STG"):
    sub = {k: v for k, v in ops.items() if k == fam or k.startswith(fam + ".")}
    print(f"{fam:<10} {sum(sub.values())}")
    if len(sub) > 1 or (sub and next(iter(sub)) != fam):
        for k in sorted(sub):
            print(f"    {k:<40} {sub[k]}")
