"""GPU parity tests (need a B200).  Port of the reference's tests/test_flash_attn.py with its commented-out
backward asserts enabled: same inputs (seed 0, batch 32, 4 heads, x ~ randn, Wqkv = nn.Linear), same acceptance
rule -- max error vs the fp32 oracle at most 2x the error of the same-precision PyTorch implementation
(tests/test_flash_attn.py:407-418) -- for O, attention probabilities, dQ, dK, dV; dropout fraction within 1 %
(:414); all-keep at p = 0 (:411-412).  Everything goes through the public API -> flash_attn_cuda -> libfa_b200.so.
"""
import math
import os

import pytest
import torch

import fa_test_utils as U
from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu

# the reference's full grid (tests/test_flash_attn.py:332-345, 421-434, 514-527, 610-623) runs by default
DTYPES = [torch.float16, torch.bfloat16]
SEQLENS = [97, 128, 200, 256, 257, 384, 512, 768, 1024, 1025, 2048]
HEADDIMS = [32, 56, 64, 80, 96, 128]
BATCH, NHEADS = 32, 4


def _api():
    from flash_attn import flash_attn_interface as fi
    return fi


def make_inputs(seqlen, d, dtype, device="cuda"):
    """tests/test_flash_attn.py:351-355"""
    torch.random.manual_seed(0)
    x = torch.randn(BATCH, seqlen, NHEADS * d, device=device, dtype=dtype, requires_grad=True)
    Wqkv = torch.nn.Linear(NHEADS * d, 3 * NHEADS * d, device=device, dtype=dtype)
    q, k, v = Wqkv(x).detach().chunk(3, dim=-1)
    return [t.reshape(BATCH, seqlen, NHEADS, d).contiguous() for t in (q, k, v)]


def check_2x(name, got, ref, pt):
    err, pt_err = orc.max_abs_err(got, ref), orc.max_abs_err(pt, ref)
    assert err <= 2 * pt_err + 1e-6, f"{name}: error {err:.3e} > 2 x PyTorch error {pt_err:.3e}"


def check_dropout(dropout_mask, dropout_p, qmask, kmask, causal):
    if dropout_p == 0.0:
        valid = qmask[:, None, :, None] & kmask[:, None, None, :]
        assert bool((dropout_mask | ~valid).all())   # tests/test_flash_attn.py:411-412
    else:
        frac = orc.get_dropout_fraction(dropout_mask, qmask, kmask, causal=causal).item()
        assert 0.99 <= frac / dropout_p <= 1.01, f"dropout fraction {frac}"   # :414


def reference_triplet(q, k, v, qmask, kmask, dropout_p, dropout_mask, causal, g):
    """fp32 oracle + low-precision PyTorch baseline, with gradients w.r.t. padded q, k, v for upstream grad g."""
    qf, kf, vf = [t.detach().clone().requires_grad_() for t in (q, k, v)]
    out_ref, attn_ref = orc.attention_ref(qf, kf, vf, qmask, kmask, dropout_p, dropout_mask, causal=causal)
    grads_ref = torch.autograd.grad(out_ref, (qf, kf, vf), g)
    qp, kp, vp = [t.detach().clone().requires_grad_() for t in (q, k, v)]
    out_pt, attn_pt = orc.attention_ref(qp, kp, vp, qmask, kmask, dropout_p, dropout_mask, causal=causal,
                                        upcast=False, reorder_ops=True)
    grads_pt = torch.autograd.grad(out_pt, (qp, kp, vp), g)
    return out_ref, attn_ref, grads_ref, out_pt, attn_pt, grads_pt


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", HEADDIMS)
@pytest.mark.parametrize("seqlen", SEQLENS)
@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
def test_flash_attn_unpadded_qkvpacked(seqlen, d, dropout_p, causal, dtype):
    """tests/test_flash_attn.py:332-418"""
    fi = _api()
    q, k, v = make_inputs(seqlen, d, dtype)
    mask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    P = U.generate_qkv(q, k, v, mask, mask, qkvpacked=True)
    out_unpad, lse, S_dmask = fi.flash_attn_unpadded_qkvpacked_func(P.qkv_unpad, P.cu_q, P.max_q, dropout_p,
                                                                    return_attn_probs=True, causal=causal)
    out = P.pad_q(out_unpad)
    dropout_mask, attn = U.decode_s_dmask(S_dmask, seqlen, seqlen)
    g = torch.randn_like(out)
    dqkv_unpad, = torch.autograd.grad(out_unpad, P.qkv_unpad, g[mask])
    dq, dk, dv = [P.pad_q(dqkv_unpad[:, i]) for i in range(3)]
    out_ref, attn_ref, gr, out_pt, attn_pt, gp = reference_triplet(q, k, v, mask, mask, dropout_p, dropout_mask,
                                                                   causal, g)
    check_2x("O", out, out_ref, out_pt)
    check_2x("attn", attn, attn_ref, attn_pt)
    check_dropout(dropout_mask, dropout_p, mask, mask, causal)
    for name, got, ref, pt in zip(("dQ", "dK", "dV"), (dq, dk, dv), gr, gp):
        check_2x(name, got, ref, pt)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", HEADDIMS)
@pytest.mark.parametrize("seqlen", SEQLENS)
@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
def test_flash_attn_unpadded_kvpacked(seqlen, d, dropout_p, causal, dtype):
    """tests/test_flash_attn.py:421-511: independent q / k padding masks => seqlen_q != seqlen_k."""
    fi = _api()
    q, k, v = make_inputs(seqlen, d, dtype)
    qmask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    kmask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    P = U.generate_qkv(q, k, v, qmask, kmask, kvpacked=True)
    out_unpad, lse, S_dmask = fi.flash_attn_unpadded_kvpacked_func(P.q_unpad, P.kv_unpad, P.cu_q, P.cu_k, P.max_q,
                                                                   P.max_k, dropout_p, return_attn_probs=True,
                                                                   causal=causal)
    out = P.pad_q(out_unpad)
    dropout_mask, attn = U.decode_s_dmask(S_dmask, seqlen, seqlen)
    g = torch.randn_like(out)
    dq_unpad, dkv_unpad = torch.autograd.grad(out_unpad, (P.q_unpad, P.kv_unpad), g[qmask])
    dq, dk, dv = P.pad_q(dq_unpad), P.pad_k(dkv_unpad[:, 0]), P.pad_k(dkv_unpad[:, 1])
    out_ref, attn_ref, gr, out_pt, attn_pt, gp = reference_triplet(q, k, v, qmask, kmask, dropout_p, dropout_mask,
                                                                   causal, g)
    check_2x("O", out, out_ref, out_pt)
    check_2x("attn", attn, attn_ref, attn_pt)
    check_dropout(dropout_mask, dropout_p, qmask, kmask, causal)
    for name, got, ref, pt in zip(("dQ", "dK", "dV"), (dq, dk, dv), gr, gp):
        check_2x(name, got, ref, pt)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", HEADDIMS)
@pytest.mark.parametrize("seqlen", SEQLENS)
@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
def test_flash_attn_unpadded(seqlen, d, dropout_p, causal, dtype):
    """tests/test_flash_attn.py:514-607"""
    fi = _api()
    q, k, v = make_inputs(seqlen, d, dtype)
    qmask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    kmask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    P = U.generate_qkv(q, k, v, qmask, kmask)
    out_unpad, lse, S_dmask = fi.flash_attn_unpadded_func(P.q_unpad, P.k_unpad, P.v_unpad, P.cu_q, P.cu_k, P.max_q,
                                                          P.max_k, dropout_p, return_attn_probs=True, causal=causal)
    out = P.pad_q(out_unpad)
    dropout_mask, attn = U.decode_s_dmask(S_dmask, seqlen, seqlen)
    g = torch.randn_like(out)
    dq_unpad, dk_unpad, dv_unpad = torch.autograd.grad(out_unpad, (P.q_unpad, P.k_unpad, P.v_unpad), g[qmask])
    dq, dk, dv = P.pad_q(dq_unpad), P.pad_k(dk_unpad), P.pad_k(dv_unpad)
    out_ref, attn_ref, gr, out_pt, attn_pt, gp = reference_triplet(q, k, v, qmask, kmask, dropout_p, dropout_mask,
                                                                   causal, g)
    check_2x("O", out, out_ref, out_pt)
    check_2x("attn", attn, attn_ref, attn_pt)
    check_dropout(dropout_mask, dropout_p, qmask, kmask, causal)
    for name, got, ref, pt in zip(("dQ", "dK", "dV"), (dq, dk, dv), gr, gp):
        check_2x(name, got, ref, pt)
    # softmax_lse: natural-log LSE of the scaled masked scores (documented at flash_attn_interface.py:168-170)
    lse_ref = orc.softmax_lse_ref(q, k, qmask, kmask, causal)
    # lse has round16(max_seqlen_q) columns, which can be fewer than the padded test length (e.g. 96 < 97)
    n = min(seqlen, lse.shape[2])
    assert not bool(qmask[:, n:].any())
    ok = (qmask[:, None, :].expand_as(lse_ref) & torch.isfinite(lse_ref))[:, :, :n]
    assert (lse[:, :, :n] - lse_ref[:, :, :n])[ok].abs().max().item() < 2e-3


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [128, 64, 32, 16])
@pytest.mark.parametrize("seqlen", SEQLENS)
@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
@pytest.mark.parametrize("ordered_dq", [False, True])
def test_flash_attn_race_condition(seqlen, d, dropout_p, causal, dtype, ordered_dq):
    """tests/test_flash_attn.py:610-671: bitwise run-to-run determinism of O and S_dmask (the reference's asserts)
    and, here, of dK and dV.  dQ is accumulated across key blocks with fp32 bulk reductions in global memory
    (north_star: "dQ is accumulated by atomics"), so by default its summation order may differ run to run: it must
    agree to fp32 round-off, far below the 16-bit output resolution.  ordered_dq: with FA_FLAG_DETERMINISTIC
    (flash_attn_cuda.set_deterministic) the key blocks take turns and dQ is bitwise reproducible too, 10 runs."""
    import flash_attn_cuda
    fi = _api()
    q, k, v = make_inputs(seqlen, d, dtype)
    qmask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    kmask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
    P = U.generate_qkv(q, k, v, qmask, kmask)

    def run():
        torch.random.manual_seed(0)
        out, lse, S = fi.flash_attn_unpadded_func(P.q_unpad, P.k_unpad, P.v_unpad, P.cu_q, P.cu_k, P.max_q, P.max_k,
                                                  dropout_p, return_attn_probs=True, causal=causal)
        torch.random.manual_seed(1)
        g = torch.randn_like(out)
        grads = torch.autograd.grad(out, (P.q_unpad, P.k_unpad, P.v_unpad), g)
        return (out, S) + tuple(grads)

    prev = flash_attn_cuda.set_deterministic(True) if ordered_dq else None
    try:
        first = run()
        for _ in range(10):
            again = run()
            for i, (a, b) in enumerate(zip(first, again)):
                if i == 2 and not ordered_dq:   # dQ
                    assert torch.allclose(a.float(), b.float(), rtol=2e-2, atol=1e-3 * max(1.0, a.abs().max().item()))
                else:
                    assert torch.equal(a, b)
    finally:
        if ordered_dq:
            flash_attn_cuda.set_deterministic(prev)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [8, 16, 24, 40, 48, 72, 88, 104, 112, 120])
def test_every_multiple_of_8_head_dim(d, causal, dtype):
    """fmha_api.cpp:163-164 accepts every head dim that is a multiple of 8 up to 128.  The kernels run them in a 64- or
    128-column tile and issue only the MMA k-steps / output columns / dQ reduction chunks that hold data
    (ceil16(d)); this walks the head dims the reference grid does not (d % 16 == 8 included), p = 0.17, ragged lengths,
    and finishes with a d = 128 call on the same stream: the dQ workspace the trimmed call left behind must be clean."""
    fi = _api()
    seqlen = 257
    for dd in (d, 128):
        q, k, v = make_inputs(seqlen, dd, dtype)
        mask = orc.generate_random_padding_mask(seqlen, BATCH, "cuda", mode="random")
        P = U.generate_qkv(q, k, v, mask, mask, qkvpacked=True)
        out_unpad, lse, S_dmask = fi.flash_attn_unpadded_qkvpacked_func(P.qkv_unpad, P.cu_q, P.max_q, 0.17,
                                                                        return_attn_probs=True, causal=causal)
        out = P.pad_q(out_unpad)
        dropout_mask, attn = U.decode_s_dmask(S_dmask, seqlen, seqlen)
        g = torch.randn_like(out)
        dqkv_unpad, = torch.autograd.grad(out_unpad, P.qkv_unpad, g[mask])
        dq, dk, dv = [P.pad_q(dqkv_unpad[:, i]) for i in range(3)]
        out_ref, attn_ref, gr, out_pt, attn_pt, gp = reference_triplet(q, k, v, mask, mask, 0.17, dropout_mask, causal, g)
        check_2x(f"O d{dd}", out, out_ref, out_pt)
        check_2x(f"attn d{dd}", attn, attn_ref, attn_pt)
        for name, got, ref, pt in zip(("dQ", "dK", "dV"), (dq, dk, dv), gr, gp):
            check_2x(f"{name} d{dd}", got, ref, pt)
