"""fa_bwd_static_scratch (include/fa_b200.h): the backward for callers that cannot allocate device memory -- its scratch
is a static __device__ array inside the library.  No counterpart in the reference.
Called here straight through the C ABI (ctypes) and compared with the existing entry point fa_bwd on the
same inputs: dK / dV bitwise (both paths are deterministic there), dQ within the tolerance the other tests use for it
(its fp32 reduction order may differ run to run).  Two calls in a row with DIFFERENT sizes, so the second call finds
the first call's contents in the shared array.  Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
FA_ERR_WORKSPACE = 4
STATIC_SCRATCH_BYTES = 32 << 20


def _problem(lens_q, lens_k, h, d, dtype, causal, seed):
    import flash_attn_cuda
    g = torch.Generator(device=DEV).manual_seed(seed)
    tq, tk = sum(lens_q), sum(lens_k)
    q = torch.randn(tq, h, d, device=DEV, dtype=dtype, generator=g)
    k = torch.randn(tk, h, d, device=DEV, dtype=dtype, generator=g)
    v = torch.randn(tk, h, d, device=DEV, dtype=dtype, generator=g)
    dout = torch.randn(tq, h, d, device=DEV, dtype=dtype, generator=g)
    cu_q = torch.tensor([0] + torch.tensor(lens_q).cumsum(0).tolist(), device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0] + torch.tensor(lens_k).cumsum(0).tolist(), device=DEV, dtype=torch.int32)
    scale = d ** -0.5
    out, lse = flash_attn_cuda.fwd(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), 0.0, scale, False, causal, False, None)
    return dict(q=q, k=k, v=v, dout=dout, out=out, lse=lse, cu_q=cu_q, cu_k=cu_k, mq=max(lens_q), mk=max(lens_k),
                scale=scale, causal=causal, b=len(lens_q), h=h, d=d)


def _bwd_existing(pr):
    """the existing entry point (fa_bwd behind flash_attn_cuda.bwd, caller-owned workspace)"""
    import flash_attn_cuda
    dq, dk, dv = torch.empty_like(pr["q"]), torch.empty_like(pr["k"]), torch.empty_like(pr["v"])
    flash_attn_cuda.bwd(pr["dout"], pr["q"], pr["k"], pr["v"], pr["out"], pr["lse"], dq, dk, dv, pr["cu_q"], pr["cu_k"],
                        pr["mq"], pr["mk"], 0.0, pr["scale"], False, pr["causal"], None)
    return dq, dk, dv


def _params(pr, dq, dk, dv, softmax_d):
    import flash_attn_cuda as fc
    P = fc.FaParams()
    fc._fill_common(P, pr["q"], pr["k"], pr["v"], pr["cu_q"], pr["cu_k"], pr["b"], pr["h"], pr["d"], pr["q"].shape[0],
                    pr["k"].shape[0], pr["mq"], pr["mk"], 0.0, pr["scale"], pr["causal"])
    P.impl = fc.FA_IMPL_AUTO
    for name, t in (("o", pr["out"]), ("dq", dq), ("dk", dk), ("dv", dv)):
        setattr(P, name, t.data_ptr())
        setattr(P, name + "_row_stride", t.stride(0))
        setattr(P, name + "_head_stride", t.stride(1))
    P.dout, P.do_row_stride, P.do_head_stride = pr["dout"].data_ptr(), pr["dout"].stride(0), pr["dout"].stride(1)
    P.softmax_lse, P.softmax_d = pr["lse"].data_ptr(), softmax_d.data_ptr()
    return P


def _bwd_static(pr):
    """the new entry point, straight through the C ABI; workspace pointer deliberately NULL"""
    import flash_attn_cuda as fc
    dq, dk, dv = [torch.full_like(pr[n], float("nan")) for n in ("q", "k", "v")]      # must be fully overwritten
    softmax_d = torch.empty_like(pr["lse"])
    P = _params(pr, dq, dk, dv, softmax_d)
    assert P.workspace is None and P.workspace_bytes == 0
    rc = fc._lib.fa_bwd_static_scratch(ctypes.byref(P), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, fc._lib.fa_last_error().decode()
    return dq, dk, dv


def _check(pr):
    ref = _bwd_existing(pr)
    got = _bwd_static(pr)
    torch.cuda.synchronize()
    for name, a, b in zip(("dQ", "dK", "dV"), got, ref):
        assert torch.isfinite(a.float()).all(), name
        if name == "dQ":     # same tolerance as tests/test_flash_attn_gpu.py::test_flash_attn_race_condition
            assert torch.allclose(a.float(), b.float(), rtol=2e-2, atol=1e-3 * max(1.0, b.abs().max().item())), name
        else:
            assert torch.equal(a, b), name


def test_static_scratch_backward_matches_fa_bwd_on_two_calls_in_a_row():
    # call 1: ragged varlen batch, d = 128 bf16 causal; call 2: other sizes / dtype / head dim, so the statistics and
    # accumulator areas of the shared array land elsewhere and meet the first call's leftovers
    first = _problem([300, 77, 512, 129], [300, 77, 512, 129], 4, 128, torch.bfloat16, True, seed=11)
    second = _problem([1000, 64, 257], [900, 200, 31], 3, 64, torch.float16, False, seed=12)
    _check(first)
    _check(second)
    _check(first)      # and back again


def test_static_scratch_rejects_sizes_beyond_the_array():
    import flash_attn_cuda as fc
    b, s, h, d = 16, 4096, 12, 128                       # the headline size needs ~400 MB of scratch
    q = torch.empty(b * s, h, d, device=DEV, dtype=torch.bfloat16)
    cu = torch.arange(0, (b + 1) * s, s, device=DEV, dtype=torch.int32)
    lse = torch.empty(b, h, s, device=DEV, dtype=torch.float32)
    pr = dict(q=q, k=q, v=q, dout=q, out=q, lse=lse, cu_q=cu, cu_k=cu, mq=s, mk=s, scale=d ** -0.5, causal=True, b=b, h=h, d=d)
    P = _params(pr, q, q, q, lse)
    assert fc._lib.fa_workspace_bytes(ctypes.byref(P), fc.FA_OP_BWD) > STATIC_SCRATCH_BYTES
    rc = fc._lib.fa_bwd_static_scratch(ctypes.byref(P), torch.cuda.current_stream().cuda_stream)
    msg = fc._lib.fa_last_error().decode()
    assert rc == FA_ERR_WORKSPACE and str(STATIC_SCRATCH_BYTES) in msg and "FA_STATIC_SCRATCH_BYTES" in msg, (rc, msg)
    torch.cuda.synchronize()      # nothing was launched, the context is healthy
