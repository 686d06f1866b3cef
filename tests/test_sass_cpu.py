"""What the built library is made of (no GPU needed: cuobjdump reads the cubins).  B200_PROFILING.md names the SASS mnemonics
that prove a kernel is Blackwell-native: UTCHMMA (tcgen05.mma), UTMALDG (TMA tensor loads), LDTM / STTM (tcgen05.ld / st),
UBLKRED (cp.reduce.async.bulk) -- and HMMA (mma.sync) must not appear: the recompiled mma.sync kernels are the baseline."""
import collections
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "hazyresearch_flash-attention_b200", "csrc", "libfa_b200.so")

pytestmark = pytest.mark.skipif(shutil.which("cuobjdump") is None, reason="cuobjdump not on PATH")


@pytest.fixture(scope="module")
def kernels():
    assert os.path.exists(LIB), "build the library first: python __graft_entry__.py"
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    out, name = {}, None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            name = m.group(1)
            out[name] = collections.Counter()
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_]*)([A-Z0-9_.]*)", line)
        if m and name:
            out[name][m.group(1)] += 1                      # opcode family ...
            if m.group(2):
                out[name][m.group(1) + m.group(2)] += 1     # ... and the full mnemonic
    return out


def test_hot_kernels_are_tcgen05_tma_kernels(kernels):
    fwd = {k: v for k, v in kernels.items() if "fwd_sm100_kernel" in k}
    bwd = {k: v for k, v in kernels.items() if "bwd_sm100_kernel" in k}
    assert len(fwd) == 40 and len(bwd) == 80, (len(fwd), len(bwd))      # bwd: 48 plain / trimmed + 32 ordered-dQ
    for name, ops in {**fwd, **bwd}.items():
        assert ops["UTCHMMA"] > 0 and ops["UTMALDG"] > 0 and ops["LDTM"] > 0, name       # tcgen05.mma, TMA, TMEM loads
        assert ops["HMMA"] == 0 and ops["IMMA"] == 0, name                                # no mma.sync
    for name, ops in fwd.items():
        assert ops["STTM"] > 0, name                                                      # P goes back to TMEM for the TS MMA
    for name, ops in bwd.items():
        assert ops["UBLKRED"] > 0, name                                                   # dQ leaves through bulk reductions


def test_no_kernel_of_the_library_uses_mma_sync(kernels):
    assert len(kernels) > 100
    assert sum(ops["HMMA"] + ops["IMMA"] + ops["DMMA"] for ops in kernels.values()) == 0


def test_ordered_dq_kernels_take_turns_with_acquire_release(kernels):
    """the 32 FA_FLAG_DETERMINISTIC instantiations (last template argument 2) are the only main kernels with the turn hand-over:
    one gpu-scope acquire load (the spin), one release store, the async-proxy fences around them; the 48 others have none of it"""
    det = [ops for k, ops in kernels.items() if re.search(r"bwd_sm100_kernelI.*ELi2EE", k)]
    plain = [ops for k, ops in kernels.items() if re.search(r"bwd_sm100_kernelI.*ELi[01]EE", k)]
    assert len(det) == 32 and len(plain) == 48
    for ops in det:
        assert ops["LDG.E.STRONG.GPU"] == 1 and ops["STG.E.STRONG.GPU"] == 1 and ops["FENCE.VIEW.ASYNC.G"] == 2
    for ops in plain:
        assert ops["LDG.E.STRONG.GPU"] == 0 and ops["STG.E.STRONG.GPU"] == 0 and ops["FENCE.VIEW.ASYNC.G"] == 0
    assert all(ops["BPT.TRAP"] > 0 for ops in det + plain)     # every mbarrier / turn wait traps on timeout instead of hanging
