"""CPU tests (no GPU needed): the oracle against fixtures produced by the reference's own Python code
(tests/golden/make_golden.py), the host-side logic of the drop-in package, and the C-ABI library's exports."""
import ctypes
import math
import os
import re

import pytest
import torch

import fa_test_utils as U
from oracle import attention_ref as orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "hazyresearch_flash-attention_b200", "csrc", "libfa_b200.so")
HEADER = os.path.join(ROOT, "include", "fa_b200.h")

ATTN_CASES = U.load_golden("attention_ref_cases.pt")
BS_CASES = U.load_golden("blocksparse_cases.pt")


@pytest.mark.parametrize("idx", range(len(ATTN_CASES)))
def test_oracle_matches_reference_attention_ref(idx):
    c = ATTN_CASES[idx]
    dmask = U.unpack_bool(c["dmask"])
    q, k, v = [c[n].clone().requires_grad_() for n in ("q", "k", "v")]
    out, attn = orc.attention_ref(q, k, v, c["qmask"], c["kmask"], c["p"], dmask, causal=c["causal"])
    # same math, same op order, same dtype -> bitwise equal on the same CPU build; allow 1 ulp of the dtype
    tol = 2e-3 if c["dtype"] == "torch.float16" else 1.6e-2
    assert (out.float() - c["out_ref"].float()).abs().max().item() <= tol * max(1.0, c["out_ref"].float().abs().max().item())
    if c["attn_ref"] is not None:
        assert (attn.float() - c["attn_ref"].float()).abs().max().item() <= tol
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), c["dout"])
    for got, ref in ((dq, c["dq"]), (dk, c["dk"]), (dv, c["dv"])):
        assert (got.float() - ref.float()).abs().max().item() <= tol * max(1.0, ref.float().abs().max().item())
    out_pt, _ = orc.attention_ref(c["q"], c["k"], c["v"], c["qmask"], c["kmask"], c["p"], dmask, causal=c["causal"],
                                  upcast=False, reorder_ops=True)
    pt_err = (out_pt.float() - out.float()).abs().max().item()
    assert abs(pt_err - c["out_pt_maxerr"]) <= 0.5 * c["out_pt_maxerr"] + 1e-6
    frac = orc.get_dropout_fraction(dmask, c["qmask"], c["kmask"], causal=c["causal"]).item()
    assert abs(frac - c["dropout_fraction"]) < 1e-6


def test_oracle_none_dropout_mask_is_keep_all():
    c = ATTN_CASES[0]
    a, _ = orc.attention_ref(c["q"], c["k"], c["v"], c["qmask"], c["kmask"], 0.0, None, causal=c["causal"])
    ones = torch.ones(c["b"], c["h"], c["sq"], c["sk"], dtype=torch.bool)
    b, _ = orc.attention_ref(c["q"], c["k"], c["v"], c["qmask"], c["kmask"], 0.0, ones, causal=c["causal"])
    assert torch.equal(a, b)


def test_oracle_lse_consistent_with_attention():
    c = ATTN_CASES[1]
    lse = orc.softmax_lse_ref(c["q"], c["k"], c["qmask"], c["kmask"], c["causal"])
    q, k = c["q"].float(), c["k"].float()
    scores = torch.einsum("bthd,bshd->bhts", q / math.sqrt(c["d"]), k)
    _, attn = orc.attention_ref(c["q"], c["k"], c["v"], None, c["kmask"], 0.0, None, causal=c["causal"])
    recon = torch.exp(scores - lse[..., None])
    vis = attn.float() > 0
    assert (recon[vis] - attn.float()[vis]).abs().max().item() < 2e-3


@pytest.mark.parametrize("idx", range(len(BS_CASES["attn"])))
def test_oracle_matches_reference_blocksparse_ref(idx):
    c = BS_CASES["attn"][idx]
    dmask = U.unpack_bool(c["dmask"])
    out, attn = orc.attention_blocksparse_ref(c["qkv"], c["blockmask"], c["amask"], c["p"], dmask)
    assert (out.float() - c["out"].float()).abs().max().item() <= 2e-3
    if c["attn"] is not None:
        assert (attn.float() - c["attn"].float()).abs().max().item() <= 1e-3


@pytest.mark.parametrize("idx", range(len(BS_CASES["convert"])))
def test_convert_blockmask_matches_reference(idx):
    from flash_attn.flash_blocksparse_attn_interface import convert_blockmask
    c = BS_CASES["convert"][idx]
    got = convert_blockmask(c["blockmask"], causal=False)
    assert got.dtype == torch.int32
    assert torch.equal(got, c["converted"])
    assert torch.equal(orc.convert_blockmask_ref(c["blockmask"]), c["converted"])


def test_convert_blockmask_rejects_causal():
    from flash_attn.flash_blocksparse_attn_interface import convert_blockmask
    with pytest.raises(AssertionError):
        convert_blockmask(torch.ones(16, 1), causal=True)


def test_butterfly_layout():
    from flash_attn.flash_blocksparse_attention import ButterflySparsityConfig
    lay = ButterflySparsityConfig().make_layout(2048)
    assert tuple(lay.shape) == (128, 8)
    assert torch.equal(lay.bool(), orc.butterfly_blockmask(2048))
    blocks = lay[::16].bool()
    assert torch.equal(blocks, blocks.T) and bool(blocks.diagonal().all())
    assert blocks.sum().item() == 8 * (1 + 3)   # self + log2(8) partners


def test_bert_padding_matches_reference_cpu():
    from flash_attn.bert_padding import pad_input, unpad_input
    for c in U.load_golden("bert_padding_cases.pt"):
        x = c["x"].clone().requires_grad_()
        xu, idx, cu, mx = unpad_input(x, c["mask"])
        assert torch.equal(xu, c["x_unpad"]) and torch.equal(idx, c["indices"])
        assert torch.equal(cu, c["cu_seqlens"]) and cu.dtype == torch.int32 and mx == c["max_seqlen"]
        back = pad_input(xu, idx, x.shape[0], x.shape[1])
        assert torch.equal(back, c["repadded"])
        back.sum().backward()   # gather bwd (scatter) o scatter bwd (gather)
        assert torch.equal(x.grad, c["mask"][..., None].expand_as(x).to(x.dtype))


def test_index_first_axis_residual_cpu():
    from flash_attn.bert_padding import index_first_axis_residual
    x = torch.randn(10, 3, 4, requires_grad=True)
    idx = torch.tensor([7, 2, 5])
    out, res = index_first_axis_residual(x, idx)
    assert torch.equal(out, x[idx]) and torch.equal(res, x)
    (out.sum() + 2 * res.sum()).backward()
    expect = torch.full_like(x, 2.0)
    expect[idx] += 1.0
    assert torch.equal(x.grad, expect)


def test_rotary_matches_reference():
    from flash_attn.rotary import RotaryEmbedding, RotaryEmbedding2D
    c = U.load_golden("rotary_cases.pt")
    q1, k1 = RotaryEmbedding(32)(c["q"], c["k"], seq_dimension=-3)
    q2, k2 = RotaryEmbedding2D(32)(c["q"], c["k"], seq_dimension=-3)
    for got, ref in ((q1, c["q1"]), (k1, c["k1"]), (q2, c["q2"]), (k2, c["k2"])):
        assert torch.allclose(got, ref, atol=1e-6, rtol=1e-6)


def test_rotary_tables_reproduce_reference_modules():
    """The fp32 tables the fused kernel consumes, applied with the kernel's pair formula, against the outputs the
    reference's RotaryEmbedding / RotaryEmbedding2D produced for the same inputs (tests/golden/rotary_cases.pt)."""
    from flash_attn.rotary import rotary_tables
    c = U.load_golden("rotary_cases.pt")
    for mode, names in (("1d", ("q1", "k1")), ("2d", ("q2", "k2"))):
        for x, ref in ((c["q"], c[names[0]]), (c["k"], c[names[1]])):
            b, s, h, d = x.shape
            cos, sin = rotary_tables(mode, s, d, x.device)
            assert cos.shape == (s, d // 2) and cos.dtype == torch.float32
            x0, x1 = x.float()[..., 0::2], x.float()[..., 1::2]
            cs, sn = cos[None, :, None, :], sin[None, :, None, :]
            got = torch.stack((x0 * cs - x1 * sn, x1 * cs + x0 * sn), dim=-1).flatten(-2)
            assert torch.allclose(got, ref.float(), atol=1e-5, rtol=1e-5)


# ------------------------------------------------------------------------------------------------------------
# C-ABI library: loads without a GPU and exports every symbol the header declares
# ------------------------------------------------------------------------------------------------------------
def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fa_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "build the library first: python __graft_entry__.py"
    lib = ctypes.CDLL(LIB)
    names = _declared_symbols()
    assert {"fa_fwd", "fa_bwd", "fa_fwd_block", "fa_bwd_block", "fa_gather_rows", "fa_scatter_rows",
            "fa_rotary_qk", "fa_last_error", "fa_workspace_bytes", "fa_philox_increment"} <= set(names)
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/fa_b200.h but not exported"


def test_shim_struct_mirror_and_host_helpers():
    import flash_attn_cuda
    assert flash_attn_cuda._lib.fa_sizeof_params() == ctypes.sizeof(flash_attn_cuda.FaParams)
    P = flash_attn_cuda.FaParams()
    P.p_dropout = 0.1
    assert flash_attn_cuda._lib.fa_philox_increment(ctypes.byref(P)) == 4
    P.p_dropout = 0.0
    assert flash_attn_cuda._lib.fa_philox_increment(ctypes.byref(P)) == 0
    P.blockmask_nrow, P.blockmask_ncol = 128, 8
    assert flash_attn_cuda._lib.fa_workspace_bytes(ctypes.byref(P), flash_attn_cuda.FA_OP_FWD_BLOCK) == 1024
    assert flash_attn_cuda.seqlen_k_rounded(97, 64) == 128 and flash_attn_cuda.seqlen_k_rounded(200, 64) == 256
    assert flash_attn_cuda.seqlen_k_rounded(257, 64) == 512 and flash_attn_cuda.seqlen_k_rounded(257, 128) == 384


def test_invalid_arguments_are_rejected_without_a_gpu():
    import flash_attn_cuda
    P = flash_attn_cuda.FaParams()
    rc = flash_attn_cuda._lib.fa_fwd(ctypes.byref(P), None)
    assert rc == 1 and b"batch size" in flash_attn_cuda._lib.fa_last_error()
    P.b, P.h, P.d = 1, 1, 12
    rc = flash_attn_cuda._lib.fa_fwd(ctypes.byref(P), None)
    assert rc == 1 and b"head dim" in flash_attn_cuda._lib.fa_last_error()
    q = torch.zeros(4, 1, 8, dtype=torch.float16)
    cu = torch.tensor([0, 4], dtype=torch.int32)
    with pytest.raises(RuntimeError):   # CPU tensors: the product path has no CPU fallback
        flash_attn_cuda.fwd(q, q, q, cu, cu, 4, 4, 0.0, 1.0, False, False, False, None)


def _bwd_params(b=16, h=12, d=128, s=4096):
    import flash_attn_cuda as F
    P = F.FaParams()
    P.b, P.h, P.d, P.total_q, P.total_k = b, h, d, b * s, b * s
    P.max_seqlen_q = P.max_seqlen_k = P.seqlen_q_rounded = P.seqlen_k_rounded = s
    for f in ("q", "k", "v", "o", "dout", "dq", "dk", "dv", "softmax_lse", "softmax_d", "cu_seqlens_q", "cu_seqlens_k"):
        setattr(P, f, 0x10000)        # never dereferenced: argument validation happens on the host
    for f in ("q", "k", "v", "o", "do", "dq", "dk", "dv"):
        setattr(P, f + "_row_stride", h * d)
        setattr(P, f + "_head_stride", d)
    P.softmax_scale = 1.0
    return P


@pytest.mark.parametrize("d,dpad", [(128, 128), (96, 128), (64, 64), (40, 64)])
def test_backward_workspace_layout_is_linear_in_total_q(d, dpad):
    """fp32 dQ accumulator (the range that is zero on entry and on exit, at byte 0), two fp32 statistics per row, 16
    bytes of turn counters per (head, tile slot); slots = total_q / 128 + batch + 1, independent of max_seqlen."""
    import flash_attn_cuda as F
    b, h, s = 16, 12, 4096
    P = _bwd_params(b, h, d, s)
    slots = (b * s >> 7) + b + 1
    accum, stats, turn = h * slots * 128 * dpad * 4, h * slots * 128 * 4, (h * slots * 16 + 255) // 256 * 256
    assert F._lib.fa_workspace_bytes(ctypes.byref(P), F.FA_OP_BWD) == accum + 2 * stats + turn
    off = ctypes.c_uint64(99)
    assert F._lib.fa_workspace_zero_range(ctypes.byref(P), F.FA_OP_BWD, ctypes.byref(off)) == accum and off.value == 0
    P.max_seqlen_q = P.seqlen_q_rounded = 8 * s          # one long sequence among short ones does not grow it
    assert F._lib.fa_workspace_bytes(ctypes.byref(P), F.FA_OP_BWD) == accum + 2 * stats + turn


def test_flag_combinations_are_validated_on_the_host():
    import flash_attn_cuda as F
    P = _bwd_params()
    P.workspace, P.workspace_bytes = 0x10000, F._lib.fa_workspace_bytes(ctypes.byref(P), F.FA_OP_BWD)
    P.flags = F.FA_FLAG_DETERMINISTIC | F.FA_FLAG_KEEP_DQ_ACCUM | F.FA_FLAG_WS_ZEROED
    assert F._lib.fa_bwd(ctypes.byref(P), None) == 1 and b"FA_FLAG_DETERMINISTIC applies to fa_bwd" in F._lib.fa_last_error()
    P.flags = F.FA_FLAG_DETERMINISTIC
    assert F._lib.fa_fwd(ctypes.byref(P), None) == 1 and b"FA_FLAG_DETERMINISTIC" in F._lib.fa_last_error()
    P.flags = F.FA_FLAG_KEEP_DQ_ACCUM
    assert F._lib.fa_bwd(ctypes.byref(P), None) == 1 and b"needs FA_FLAG_WS_ZEROED" in F._lib.fa_last_error()
    P.flags = 8
    assert F._lib.fa_bwd(ctypes.byref(P), None) == 1 and b"unknown flags" in F._lib.fa_last_error()


def test_deterministic_switch_follows_torch_unless_set():
    import flash_attn_cuda as F
    assert F._deterministic is None and F._want_deterministic() == torch.are_deterministic_algorithms_enabled()
    try:
        assert F.set_deterministic(True) is None and F._want_deterministic()
        assert F.set_deterministic(False) is True and not F._want_deterministic()
        torch.use_deterministic_algorithms(True, warn_only=True)
        assert not F._want_deterministic()                      # an explicit setting wins over torch's switch
        assert F.set_deterministic(None) is False and F._want_deterministic()
        with pytest.raises(TypeError):
            F.set_deterministic(1)
    finally:
        torch.use_deterministic_algorithms(False)
        F.set_deterministic(None)
    assert not F._want_deterministic()


def test_dropout_stream_statistics_host():
    """The position-keyed Philox stream (integer work, evaluated on the host) keeps ~(1-p) of positions."""
    import flash_attn_cuda
    p = 0.17
    kept = total = 0
    for row in range(64):
        for g in range(64):
            keep = flash_attn_cuda.dropout_keep8_host(1234, 8, 3, row, g, p)
            kept += sum(keep)
            total += 8
    assert abs((1 - kept / total) / p - 1.0) < 0.03
    a = flash_attn_cuda.dropout_keep8_host(1, 0, 0, 5, 7, 0.5)
    assert a == flash_attn_cuda.dropout_keep8_host(1, 0, 0, 5, 7, 0.5)
    assert a != flash_attn_cuda.dropout_keep8_host(1, 4, 0, 5, 7, 0.5) or a != flash_attn_cuda.dropout_keep8_host(2, 0, 0, 5, 7, 0.5)
    assert all(flash_attn_cuda.dropout_keep8_host(1, 0, 0, 5, 7, 0.0))


def test_package_resolves_into_repo():
    import flash_attn
    import flash_attn_cuda
    assert flash_attn.__file__.startswith(ROOT) and flash_attn_cuda.__file__.startswith(ROOT)
    from flash_attn.flash_attn_interface import (flash_attn_func, flash_attn_unpadded_func,  # noqa: F401
                                                 flash_attn_unpadded_kvpacked_func,
                                                 flash_attn_unpadded_qkvpacked_func, _get_block_size)
    from flash_attn.flash_attention import FlashAttention, FlashMHA   # noqa: F401
    from flash_attn.flash_blocksparse_attn_interface import flash_blocksparse_attn_func   # noqa: F401
    from flash_attn.flash_blocksparse_attention import FlashBlocksparseAttention, FlashBlocksparseMHA   # noqa: F401


def test_product_never_touches_the_oracle_or_a_fallback():
    """The oracle is test infrastructure: nothing under the package may import it, and the package holds no second backend
    (no Triton / torch.compile / SDPA call that could stand in for the CUDA library)."""
    import ast
    pkg = os.path.join(ROOT, "hazyresearch_flash-attention_b200")
    n_files = 0
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if not f.endswith(".py"):
                continue
            n_files += 1
            path = os.path.join(dirpath, f)
            tree = ast.parse(open(path).read())
            for node in ast.walk(tree):
                names = []
                if isinstance(node, ast.Import):
                    names = [a.name for a in node.names]
                elif isinstance(node, ast.ImportFrom):
                    names = [node.module or ""]
                for n in names:
                    assert n.split(".")[0] not in ("oracle", "triton", "tilelang", "flashinfer", "vllm", "cutlass"), (path, n)
                if isinstance(node, ast.Attribute):
                    assert node.attr not in ("scaled_dot_product_attention", "compile"), (path, node.attr)
    assert n_files >= 8
    # an upstream flash_attn wheel is installed in this image: the package must shadow it, not use it
    import flash_attn_cuda
    assert not hasattr(flash_attn_cuda, "flash_attn_2_cuda")


def test_header_is_plain_c_and_links_from_c(tmp_path):
    """The drop-in boundary is a C ABI: include/fa_b200.h has to compile as C99 (cgo / ctypes / JNI bindings read it as C), and a C
    program has to link against the library and get sane answers from the host-side entry points."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("gcc not on PATH")
    hdr = os.path.join(ROOT, "include", "fa_b200.h")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", hdr], check=True)
    src = tmp_path / "probe.c"
    src.write_text(r"""
#include <stdio.h>
#include <string.h>
#include "fa_b200.h"
int main(void) {
  fa_params p;
  memset(&p, 0, sizeof p);
  if (fa_sizeof_params() != sizeof p) return 2;
  if (fa_fwd(&p, 0) == FA_OK) return 3;                       /* an empty block is rejected on the host ... */
  if (strstr(fa_last_error(), "batch size") == 0) return 4;   /* ... with a message */
  p.p_dropout = 0.1f;
  if (fa_philox_increment(&p) <= 0) return 5;
  printf("abi %d\n", fa_abi_version());
  return 0;
}
""")
    exe = tmp_path / "probe"
    libdir = os.path.dirname(LIB)
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe), "-L", libdir, "-lfa_b200",
                    f"-Wl,-rpath,{libdir}"], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0 and out.stdout.strip() == "abi 2", (out.returncode, out.stdout, out.stderr)
