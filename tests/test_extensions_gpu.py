"""GPU tests of what this build adds around the reference's hot path: isolation of packed sequences (NaN in a neighbour
must not leak), sub-range addressing (`seqused_*`, used by the K/V ring), the ALiBi attention bias, the old-order
`flash_attn_func` alias, the O(total_q) backward scratch kept between calls, and -- when the unmodified reference
Python package is installed under baseline/_ref -- the reference's OWN flash_attn_interface.py /
flash_blocksparse_attn_interface.py running on top of the `flash_attn_cuda` shim."""
import importlib.util
import os
import sys

import pytest
import torch

import fa_test_utils as U
from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _fi():
    from flash_attn import flash_attn_interface as fi
    return fi


@pytest.mark.parametrize("causal", [False, True])
def test_flash_attn_func_old_argument_order(causal):
    """flash_attn_func(qkv, cu_seqlens, dropout_p, max_s, ...) == the qkvpacked function (reference :246-252)."""
    fi = _fi()
    b, s, h, d = 3, 200, 2, 64
    torch.manual_seed(0)
    qkv = torch.randn(b * s, 3, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    cu = U.cu_seqlens_full(b, s, DEV)
    out_old = fi.flash_attn_func(qkv, cu, 0.0, s, causal=causal)
    out_new = fi.flash_attn_unpadded_qkvpacked_func(qkv, cu, s, 0.0, causal=causal)
    assert torch.equal(out_old, out_new)
    g = torch.randn_like(out_old)
    (g_old,) = torch.autograd.grad(out_old, qkv, g)
    x = qkv.detach().reshape(b, s, 3, h, d)
    xf = x.float().requires_grad_()
    ref, _ = orc.attention_qkvpacked_ref(xf, None, causal=causal)
    (gref,) = torch.autograd.grad(ref, xf, g.reshape(b, s, h, d).float())
    xp = x.clone().requires_grad_()
    pt, _ = orc.attention_qkvpacked_ref(xp, None, causal=causal, upcast=False, reorder_ops=True)
    (gpt,) = torch.autograd.grad(pt, xp, g.reshape(b, s, h, d))
    assert orc.within_2x(out_old.reshape(b, s, h, d), ref, pt)
    assert orc.max_abs_err(g_old.reshape(b, s, 3, h, d), gref) <= 2 * orc.max_abs_err(gpt, gref) + 1e-6
    # the S_dmask block-size helper answers like the reference on this device class (reference :8-14)
    assert fi._get_block_size(torch.device(DEV), 64, False) == 256 and fi._get_block_size(torch.device(DEV), 128, True) == 128


@pytest.mark.parametrize("d", [64, 128])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("poison", [float("nan"), float("inf")])
def test_sequences_are_isolated_from_their_neighbours(d, causal, poison):
    """The packed tiles of one sequence physically contain rows of the next sequence and of the buffer tail behind
    cu_seqlens[-1].  The reference predicates its loads and zero-fills; here the stray rows are zeroed in shared
    memory.  Poisoning sequence 1 and the tail must leave O / dQ / dK / dV of sequences 0 and 2 bit-identical."""
    fi = _fi()
    h = 2
    lens_q, lens_k = [100, 77, 150], [130, 60, 200]       # every boundary falls inside a 128-row tile
    tail = 40
    torch.manual_seed(1)
    cu_q = torch.tensor([0, 100, 177, 327], device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0, 130, 190, 390], device=DEV, dtype=torch.int32)

    def run(poisoned):
        g = torch.Generator(device=DEV).manual_seed(7)
        q = torch.randn(sum(lens_q) + tail, h, d, device=DEV, dtype=torch.bfloat16, generator=g)
        k = torch.randn(sum(lens_k) + tail, h, d, device=DEV, dtype=torch.bfloat16, generator=g)
        v = torch.randn(sum(lens_k) + tail, h, d, device=DEV, dtype=torch.bfloat16, generator=g)
        dout = torch.randn(sum(lens_q) + tail, h, d, device=DEV, dtype=torch.bfloat16, generator=g)
        if poisoned:
            q[100:177], k[130:190], v[130:190], dout[100:177] = poison, poison, poison, poison
            q[327:], k[390:], v[390:], dout[327:] = poison, poison, poison, poison
        q, k, v = [t.requires_grad_() for t in (q, k, v)]
        out = fi.flash_attn_unpadded_func(q, k, v, cu_q, cu_k, max(lens_q), max(lens_k), 0.0, causal=causal)
        dq, dk, dv = torch.autograd.grad(out, (q, k, v), dout)
        return out.detach(), dq, dk, dv

    clean, dirty = run(False), run(True)
    for name, a, b_, bounds in (("O", clean[0], dirty[0], cu_q), ("dQ", clean[1], dirty[1], cu_q),
                                ("dK", clean[2], dirty[2], cu_k), ("dV", clean[3], dirty[3], cu_k)):
        for i in (0, 2):
            sl = slice(bounds[i].item(), bounds[i + 1].item())
            assert torch.isfinite(b_[sl].float()).all(), f"{name} of sequence {i} is poisoned"
            if name == "dQ":    # fp32 reduction order may differ run to run
                assert torch.allclose(a[sl].float(), b_[sl].float(), rtol=2e-2, atol=1e-3 * max(1.0, a[sl].abs().max().item()))
            else:
                assert torch.equal(a[sl], b_[sl]), f"{name} of sequence {i} changed"


@pytest.mark.parametrize("d,dtype", [(128, torch.bfloat16), (64, torch.float16)])
@pytest.mark.parametrize("plan", [(0, 2, 1, False), (1, 1, 2, False), (0, 2, 2, True)])
def test_seqused_addresses_subranges_of_a_packed_batch(plan, d, dtype):
    """One launch over (q rows [q_off*c, (q_off+q_len)*c), key rows [0, k_len*c)) of every 2c-row batch element -- the
    three step kinds of the K/V ring -- equals the same attention on sliced, re-packed tensors."""
    import flash_attn_cuda
    q_off, q_len, k_len, causal = plan
    b, c, h = 3, 192, 2
    torch.manual_seed(2)
    q, k, v, dout = [torch.randn(b, 2 * c, h, d, device=DEV, dtype=dtype) for _ in range(4)]
    cu = torch.arange(0, (b + 1) * 2 * c, 2 * c, device=DEV, dtype=torch.int32)
    half = torch.full((b,), c, device=DEV, dtype=torch.int32)
    sq, sk = q_len * c, k_len * c
    su_q, su_k = (half if q_len == 1 else None), (half if k_len == 1 else None)
    scale = d ** -0.5
    flat = lambda t: t.reshape(b * 2 * c, h, d)
    qv = flat(q)[q_off * c:]
    (o, lse), _ = flash_attn_cuda.fwd_ex(qv, flat(k), flat(v), cu, cu, sq, sk, 0.0, scale, causal, seqused_q=su_q, seqused_k=su_k)
    dq, dk, dv = torch.full_like(qv, float("nan")), torch.full_like(flat(k), float("nan")), torch.full_like(flat(v), float("nan"))
    flash_attn_cuda.bwd(flat(dout)[q_off * c:], qv, flat(k), flat(v), o, lse, dq, dk, dv, cu, cu, sq, sk, 0.0, scale, False,
                        causal, None, seqused_q=su_q, seqused_k=su_k)
    # the same problem on re-packed slices
    qs, ks, vs, gs = q[:, q_off * c:q_off * c + sq], k[:, :sk], v[:, :sk], dout[:, q_off * c:q_off * c + sq]
    cq, ck = U.cu_seqlens_full(b, sq, DEV), U.cu_seqlens_full(b, sk, DEV)
    q2, k2, v2 = [t.reshape(-1, h, d).contiguous().requires_grad_() for t in (qs, ks, vs)]
    o2 = _fi().flash_attn_unpadded_func(q2, k2, v2, cq, ck, sq, sk, 0.0, causal=causal)
    dq2, dk2, dv2 = torch.autograd.grad(o2, (q2, k2, v2), gs.reshape(-1, h, d).contiguous())
    pad = lambda t, off: torch.cat([t.new_zeros(off * c, h, d), t]).reshape(b, 2 * c, h, d)
    assert torch.equal(pad(o, q_off)[:, q_off * c:q_off * c + sq], o2.reshape(b, sq, h, d))
    got_dq = pad(dq, q_off)[:, q_off * c:q_off * c + sq]
    assert torch.allclose(got_dq.float(), dq2.reshape(b, sq, h, d).float(), rtol=2e-2, atol=1e-3 * max(1.0, dq2.abs().max().item()))
    assert torch.equal(dk.reshape(b, 2 * c, h, d)[:, :sk], dk2.reshape(b, sk, h, d))
    assert torch.equal(dv.reshape(b, 2 * c, h, d)[:, :sk], dv2.reshape(b, sk, h, d))
    if k_len == 1:      # key rows outside the used range are never written
        assert torch.isnan(dk.reshape(b, 2 * c, h, d)[:, sk:]).all()


def _alibi_slopes(h):
    return torch.tensor([2.0 ** (-8.0 * (i + 1) / h) for i in range(h)], device=DEV, dtype=torch.float32)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [32, 64, 96, 128])
@pytest.mark.parametrize("seqlen", [113, 256, 600])
@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
def test_alibi_bias(seqlen, d, dropout_p, causal, dtype):
    """Additive bias -slope[h] * |i - j| (reference roadmap README.md:57) under the reference's 2x rule,
    against the oracle extended with the same bias; varlen batch with seqlen_q != seqlen_k."""
    fi = _fi()
    b, h = 4, 4
    torch.manual_seed(0)
    q, k, v = [torch.randn(b, seqlen, h, d, device=DEV, dtype=dtype) for _ in range(3)]
    qmask = orc.generate_random_padding_mask(seqlen, b, DEV, mode="random")
    kmask = orc.generate_random_padding_mask(seqlen, b, DEV, mode="random")
    slopes = _alibi_slopes(h)
    P = U.generate_qkv(q, k, v, qmask, kmask)
    out_u, lse, S = fi.flash_attn_unpadded_func(P.q_unpad, P.k_unpad, P.v_unpad, P.cu_q, P.cu_k, P.max_q, P.max_k,
                                                dropout_p, causal=causal, return_attn_probs=True, alibi_slopes=slopes)
    keep, probs = U.decode_s_dmask(S, seqlen, seqlen)
    g = torch.randn(b, seqlen, h, d, device=DEV, dtype=dtype)
    dq, dk, dv = torch.autograd.grad(out_u, (P.q_unpad, P.k_unpad, P.v_unpad), g[qmask])
    qf, kf, vf = [t.float().requires_grad_() for t in (q, k, v)]
    ref, attn_ref = orc.attention_ref(qf, kf, vf, qmask, kmask, dropout_p, keep, causal=causal, alibi_slopes=slopes)
    gr = torch.autograd.grad(ref, (qf, kf, vf), g.float())
    qp, kp, vp = [t.clone().requires_grad_() for t in (q, k, v)]
    pt, attn_pt = orc.attention_ref(qp, kp, vp, qmask, kmask, dropout_p, keep, causal=causal, upcast=False,
                                    reorder_ops=True, alibi_slopes=slopes)
    gp = torch.autograd.grad(pt, (qp, kp, vp), g)
    assert orc.within_2x(P.pad_q(out_u), ref, pt, floor=1e-6)
    assert orc.within_2x(probs, attn_ref, attn_pt, floor=1e-6)
    for name, got, r, p_ in (("dQ", P.pad_q(dq), gr[0], gp[0]), ("dK", P.pad_k(dk), gr[1], gp[1]), ("dV", P.pad_k(dv), gr[2], gp[2])):
        err, pt_err = orc.max_abs_err(got, r), orc.max_abs_err(p_, r)
        assert err <= 2 * pt_err + 1e-6, f"{name}: error {err:.3e} > 2 x PyTorch error {pt_err:.3e}"
    # and the bias really is in effect
    out_nobias = fi.flash_attn_unpadded_func(P.q_unpad, P.k_unpad, P.v_unpad, P.cu_q, P.cu_k, P.max_q, P.max_k, 0.0, causal=causal)
    assert dropout_p > 0 or (out_nobias.float() - out_u.float()).abs().max().item() > 1e-2


def test_alibi_matches_simt_cross_check_at_large_size():
    fi = _fi()
    b, s, h, d = 1, 4096, 4, 128
    torch.manual_seed(3)
    q, k, v = [torch.randn(b * s, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True) for _ in range(3)]
    cu = U.cu_seqlens_full(b, s, DEV)
    slopes = _alibi_slopes(h)
    g = torch.randn(b * s, h, d, device=DEV, dtype=torch.bfloat16)
    res = []
    for simt in (False, True):
        import contextlib
        with (U.simt_kernels() if simt else contextlib.nullcontext()):
            out, lse, _ = fi.flash_attn_unpadded_func(q, k, v, cu, cu, s, s, 0.0, causal=True, return_attn_probs=True,
                                                      alibi_slopes=slopes)
            res.append((out.detach(), lse) + torch.autograd.grad(out, (q, k, v), g))
    (o_a, l_a, *g_a), (o_b, l_b, *g_b) = res
    assert (l_a[:, :, :s] - l_b[:, :, :s]).abs().max().item() < 1e-3
    assert (o_a.float() - o_b.float()).abs().max().item() < 3e-2
    for x, y in zip(g_a, g_b):
        assert (x.float() - y.float()).abs().max().item() < 3e-2 * max(1.0, y.float().abs().max().item())


def test_backward_scratch_is_kept_zero_between_calls_of_different_shapes():
    """The fp32 dQ accumulator lives in a per-stream buffer that every backward leaves all-zero (FA_FLAG_WS_ZEROED).
    Alternating shapes moves the statistics area over the accumulator area: results must not depend on the history."""
    import flash_attn_cuda
    fi = _fi()
    shapes = [(2, 1024, 4, 128, True), (5, 160, 2, 64, False), (1, 2048, 3, 96, True), (3, 333, 4, 32, False), (2, 1024, 4, 128, True)]
    first = {}
    for rep in range(2):
        for (b, s, h, d, causal) in shapes:
            gcpu = torch.Generator().manual_seed(b * 1000 + s)
            qkv = torch.randn(b * s, 3, h, d, generator=gcpu).to(DEV, torch.bfloat16).requires_grad_()
            g = torch.randn(b * s, h, d, generator=gcpu).to(DEV, torch.bfloat16)
            cu = U.cu_seqlens_full(b, s, DEV)
            out = fi.flash_attn_unpadded_qkvpacked_func(qkv, cu, s, 0.0, causal=causal)
            (dqkv,) = torch.autograd.grad(out, qkv, g)
            key = (b, s, h, d, causal)
            if key in first:
                assert torch.equal(dqkv[:, 1:], first[key][:, 1:])
                assert torch.allclose(dqkv[:, 0].float(), first[key][:, 0].float(), rtol=2e-2, atol=1e-3 * dqkv[:, 0].abs().max().item())
            else:
                first[key] = dqkv
                x = qkv.detach().reshape(b, s, 3, h, d)
                xf = x.float().requires_grad_()
                ref, _ = orc.attention_qkvpacked_ref(xf, None, causal=causal)
                (gref,) = torch.autograd.grad(ref, xf, g.reshape(b, s, h, d).float())
                xp = x.clone().requires_grad_()
                pt, _ = orc.attention_qkvpacked_ref(xp, None, causal=causal, upcast=False, reorder_ops=True)
                (gpt,) = torch.autograd.grad(pt, xp, g.reshape(b, s, h, d))
                assert orc.max_abs_err(dqkv.reshape(b, s, 3, h, d), gref) <= 2 * orc.max_abs_err(gpt, gref) + 1e-6
    torch.cuda.synchronize()
    for ent in flash_attn_cuda._ws_cache.values():      # the zero range of the last layout really is zero again
        ws, lo, hi = ent
        assert int(ws[:lo].count_nonzero().item()) == 0


# ------------------------------------------------------------------------------------------------------------
# the reference's own Python layer on top of the shim
# ------------------------------------------------------------------------------------------------------------
REF_PKG = os.path.join(ROOT, "baseline", "_ref", "flash_attn")


def _load_reference_module(name):
    """The UNMODIFIED reference file (installed by __graft_entry__.build() into baseline/_ref, never part of this
    repository's history) imported under a private name; its `import flash_attn_cuda` resolves to the shim."""
    path = os.path.join(REF_PKG, name + ".py")
    if not os.path.exists(path):
        pytest.skip("baseline/_ref/flash_attn not installed (python __graft_entry__.py in a container that has /root/reference)")
    spec = importlib.util.spec_from_file_location("_reference_" + name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
@pytest.mark.parametrize("causal", [False, True])
def test_reference_flash_attn_interface_runs_unmodified_on_the_shim(causal, dropout_p):
    """INTEGRATION.md option A: the reference's flash_attn_interface.py (its autograd Functions, its RNG save / restore
    protocol, its positional calls into flash_attn_cuda.fwd / .bwd) works on top of this library as is."""
    ref_fi = _load_reference_module("flash_attn_interface")
    import flash_attn_cuda
    assert ref_fi.flash_attn_cuda is flash_attn_cuda
    b, s, h, d = 4, 300, 4, 64
    torch.manual_seed(0)
    q, k, v = [torch.randn(b, s, h, d, device=DEV, dtype=torch.float16) for _ in range(3)]
    qmask = orc.generate_random_padding_mask(s, b, DEV, mode="random")
    kmask = orc.generate_random_padding_mask(s, b, DEV, mode="random")
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.float16)
    for form in ("sep", "kv", "qkv"):
        km = qmask if form == "qkv" else kmask
        P = U.generate_qkv(q, k, v, qmask, km, kvpacked=form == "kv", qkvpacked=form == "qkv")
        if form == "sep":
            leaves = (P.q_unpad, P.k_unpad, P.v_unpad)
            out_u, lse, S = ref_fi.flash_attn_unpadded_func(*leaves, P.cu_q, P.cu_k, P.max_q, P.max_k, dropout_p,
                                                            return_attn_probs=True, causal=causal)
        elif form == "kv":
            leaves = (P.q_unpad, P.kv_unpad)
            out_u, lse, S = ref_fi.flash_attn_unpadded_kvpacked_func(*leaves, P.cu_q, P.cu_k, P.max_q, P.max_k, dropout_p,
                                                                     return_attn_probs=True, causal=causal)
        else:
            leaves = (P.qkv_unpad,)
            out_u, lse, S = ref_fi.flash_attn_unpadded_qkvpacked_func(*leaves, P.cu_q, P.max_q, dropout_p,
                                                                      return_attn_probs=True, causal=causal)
        keep, _ = U.decode_s_dmask(S, s, s)
        grads = torch.autograd.grad(out_u, leaves, g[qmask])
        if form == "sep":
            dq, dk, dv = P.pad_q(grads[0]), P.pad_k(grads[1]), P.pad_k(grads[2])
        elif form == "kv":
            dq, dk, dv = P.pad_q(grads[0]), P.pad_k(grads[1][:, 0]), P.pad_k(grads[1][:, 1])
        else:
            dq, dk, dv = [P.pad_q(grads[0][:, i]) for i in range(3)]
        qf, kf, vf = [t.float().requires_grad_() for t in (q, k, v)]
        ref, _ = orc.attention_ref(qf, kf, vf, qmask, km, dropout_p, keep, causal=causal)
        gr = torch.autograd.grad(ref, (qf, kf, vf), g.float())
        qp, kp, vp = [t.clone().requires_grad_() for t in (q, k, v)]
        pt, _ = orc.attention_ref(qp, kp, vp, qmask, km, dropout_p, keep, causal=causal, upcast=False, reorder_ops=True)
        gp = torch.autograd.grad(pt, (qp, kp, vp), g)
        assert orc.within_2x(P.pad_q(out_u), ref, pt)
        for got, r, p_ in zip((dq, dk, dv), gr, gp):     # the backward drew the forward's mask through the reference's RNG replay
            assert orc.max_abs_err(got, r) <= 2 * orc.max_abs_err(p_, r) + 1e-6


@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
def test_reference_blocksparse_interface_runs_unmodified_on_the_shim(dropout_p):
    ref_bs = _load_reference_module("flash_blocksparse_attn_interface")
    b, s, h, d = 4, 512, 4, 64
    torch.manual_seed(0)
    qkv = torch.randn(b, s, 3, h, d, device=DEV, dtype=torch.float16)
    blockmask = orc.generate_sparsity_mask(s, 0.4, device=DEV)
    blockmask[:, 0] = True
    amask = orc.generate_random_padding_mask(s, b, DEV, mode="random")
    from flash_attn.bert_padding import unpad_input
    xu, idx, cu, mx = unpad_input(qkv.reshape(b, s, -1), amask)
    xu = xu.reshape(-1, 3, h, d).detach().requires_grad_()
    out_u, S, lse = ref_bs.flash_blocksparse_attn_func(xu, cu, blockmask, dropout_p, mx, return_attn_probs=True)
    keep, _ = U.decode_s_dmask(S, s, s)
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.float16)
    (dqkv_u,) = torch.autograd.grad(out_u, xu, g[amask])
    qkv_f = qkv.float().requires_grad_()
    ref, _ = orc.attention_blocksparse_ref(qkv_f, blockmask, amask, dropout_p, keep)
    (gref,) = torch.autograd.grad(ref, qkv_f, g.float())
    qkv_p = qkv.clone().requires_grad_()
    pt, _ = orc.attention_blocksparse_ref(qkv_p, blockmask, amask, dropout_p, keep, upcast=False)
    (gpt,) = torch.autograd.grad(pt, qkv_p, g)
    assert orc.max_abs_err(U.unpack_rows(out_u, idx, b, s), ref) <= 2 * orc.max_abs_err(pt, ref) + 1e-6
    assert orc.max_abs_err(U.unpack_rows(dqkv_u, idx, b, s), gref) <= 2 * orc.max_abs_err(gpt, gref) + 1e-6


# ------------------------------------------------------------------------------------------------------------
# varlen bookkeeping on the device
# ------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("b,s", [(1, 7), (3, 257), (16, 512), (33, 1000), (1500, 40), (4, 4096)])
def test_unpad_input_nosync_matches_unpad_input(b, s):
    """indices / cu_seqlens / packed rows equal the reference-style unpad_input (integer work: bit-exact), including
    empty and full rows and more sequences than one scan chunk; entries past the total are -1 / zero rows."""
    from flash_attn.bert_padding import pad_input, unpad_input, unpad_input_nosync
    g = torch.Generator(device=DEV).manual_seed(b * 7919 + s)
    lens = torch.randint(0, s + 1, (b,), device=DEV, generator=g)
    lens[0] = s                      # a full row
    if b > 2:
        lens[1] = 0                  # an empty row
    mask = torch.arange(s, device=DEV)[None] < lens[:, None]
    if b > 3:                        # and one row with holes (a mask need not be a prefix)
        mask[3] = torch.rand(s, device=DEV, generator=g) < 0.5
    x = torch.randn(b, s, 24, device=DEV, dtype=torch.float16, generator=g)
    xu, idx, cu, mx = unpad_input(x, mask)
    xn, idxn, cun, mxn = unpad_input_nosync(x, mask)
    nnz = idx.numel()
    assert idxn.shape == (b * s,) and torch.equal(idxn[:nnz], idx) and bool((idxn[nnz:] == -1).all())
    assert torch.equal(cun, cu) and mxn == s and mx <= mxn
    assert torch.equal(xn[:nnz], xu) and xn[nnz:].abs().max().item() == 0 if nnz < b * s else torch.equal(xn, xu)
    assert torch.equal(pad_input(xn, idxn, b, s), pad_input(xu, idx, b, s))
    import flash_attn_cuda
    _, _, stats = flash_attn_cuda.varlen_index(mask)
    assert stats.tolist() == [int(mask.sum(-1).max()), nnz]


def test_flash_attention_module_makes_no_host_synchronisation():
    """FlashAttention.forward with a key_padding_mask, forward and backward, under torch's sync debug mode: the
    reference synchronises twice per call in unpad_input (bert_padding.py:109-112); this path must not at all."""
    from flash_attn.flash_attention import FlashAttention
    b, s, h, d = 4, 300, 4, 64
    torch.manual_seed(0)
    qkv = torch.randn(b, s, 3, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    mask = orc.generate_random_padding_mask(s, b, DEV, mode="third")
    mod = FlashAttention().eval()
    g = torch.randn(b, s, h, d, device=DEV, dtype=torch.float16)
    out, _ = mod(qkv, key_padding_mask=mask, causal=True)          # warm-up: allocator, one-time kernel attributes
    torch.autograd.grad(out, qkv, g)
    torch.cuda.synchronize()
    torch.cuda.set_sync_debug_mode("error")
    try:
        out, _ = mod(qkv, key_padding_mask=mask, causal=True)
        (dqkv,) = torch.autograd.grad(out, qkv, g)
    finally:
        torch.cuda.set_sync_debug_mode("default")
    xf = qkv.detach().float().requires_grad_()
    ref, _ = orc.attention_qkvpacked_ref(xf, mask, causal=True)
    (gref,) = torch.autograd.grad(ref, xf, g.float())
    xp = qkv.detach().clone().requires_grad_()
    pt, _ = orc.attention_qkvpacked_ref(xp, mask, causal=True, upcast=False, reorder_ops=True)
    (gpt,) = torch.autograd.grad(pt, xp, g)
    assert orc.within_2x(out, ref, pt)
    valid = mask[:, :, None, None, None]
    assert orc.max_abs_err(dqkv * valid, gref * valid) <= 2 * orc.max_abs_err(gpt * valid, gref * valid) + 1e-6
    assert dqkv[~mask].abs().max().item() == 0           # padded tokens get exactly zero gradient


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16, torch.float32])
def test_index_first_axis_residual_backward_kernel(dtype):
    """forward (x[indices], x) / backward grad_residual + scatter(grad_output) (reference bert_padding.py:67-96),
    through fa_scatter_add_rows; indices unique, so the result equals index_add bit for bit."""
    from flash_attn.bert_padding import index_first_axis_residual
    n, w = 3000, 96
    g = torch.Generator(device=DEV).manual_seed(1)
    x = torch.randn(n, 4, w // 4, device=DEV, dtype=dtype, generator=g, requires_grad=True)
    idx = torch.randperm(n, device=DEV, generator=g)[:1234].sort().values
    y, res = index_first_axis_residual(x, idx)
    assert torch.equal(y, x[idx]) and torch.equal(res, x.detach())
    gy, gres = torch.randn_like(y), torch.randn_like(res)
    from flash_attn.bert_padding import IndexFirstAxisResidual

    class _Ctx:
        saved_tensors = (idx,)
        first_axis_dim = n
    got, _ = IndexFirstAxisResidual.backward(_Ctx, gy, gres)
    want = gres.reshape(n, -1).index_add(0, idx, gy.reshape(idx.numel(), -1)).reshape(x.shape)
    assert torch.equal(got, want)


@pytest.mark.parametrize("d,dtype", [(128, torch.bfloat16), (64, torch.float16)])
def test_dq_accumulates_across_backward_calls_in_the_workspace(d, dtype):
    """FA_FLAG_KEEP_DQ_ACCUM + fa_bwd_dq_finalize (what the K/V ring does per step): the keys are split in two halves, each
    differentiated by its own backward call with the FINAL lse, dQ kept in the caller's fp32 workspace; one finalise.  Must
    equal the ordinary backward over all keys: dK / dV bitwise (same per-key-block work), dQ within its usual tolerance.
    Query sub-ranges (second call: only the later half of the queries, through cu_seqlens_q + offset) hit the same slots."""
    import flash_attn_cuda as fc
    b, c, h = 3, 256, 2
    torch.manual_seed(4)
    q, k, v, dout = [torch.randn(b * 2 * c, h, d, device=DEV, dtype=dtype) for _ in range(4)]
    cu = torch.arange(0, (b + 1) * 2 * c, 2 * c, device=DEV, dtype=torch.int32)
    half = torch.full((b,), c, device=DEV, dtype=torch.int32)
    scale = d ** -0.5
    # reference: rows [0, c) of every sequence see keys [0, c); rows [c, 2c) see all keys = two "ring steps":
    #   step A: all queries x first key half;   step B: later query half x second key half
    (o_a, lse_a), _ = fc.fwd_ex(q, k, v, cu, cu, 2 * c, c, 0.0, scale, False, seqused_k=half)
    (o_b, lse_b), _ = fc.fwd_ex(q, k, v, cu + c, cu + c, c, c, 0.0, scale, False, seqused_q=half, seqused_k=half)
    o_acc = torch.empty(b * 2 * c, h, d, device=DEV)
    lse = torch.empty(b, h, 2 * c, device=DEV)
    fc.merge_partials(o_acc, lse, o_a, lse_a, b, 2 * c, True)
    fc.merge_partials(o_acc[c:], lse[:, :, c:], o_b[c:], lse_b, b, c, False, batch_rows=2 * c)
    out = o_acc.to(dtype)
    # (1) the two-call backward with dQ kept in the workspace
    ws = torch.zeros(fc.bwd_workspace_bytes(q, cu, 2 * c), dtype=torch.uint8, device=DEV)
    dk1, dv1, dk2, dv2 = [torch.zeros_like(k) for _ in range(4)]
    fc.bwd(dout, q, k, v, out, lse, None, dk1, dv1, cu, cu, 2 * c, c, 0.0, scale, False, False, None, seqused_k=half, dq_workspace=ws)
    fc.bwd(dout, q, k, v, out, lse[:, :, c:].contiguous(), None, dk2, dv2, cu + c, cu + c, c, c, 0.0, scale, False, False, None,
           seqused_q=half, seqused_k=half, dq_workspace=ws)
    dq = fc.bwd_dq_finalize(torch.full_like(q, float("nan")), cu, 2 * c, scale, ws)
    torch.cuda.synchronize()
    assert int(ws[: fc._lib.fa_workspace_zero_range(__import__("ctypes").byref(fc._dq_layout_params(q, cu, 2 * c, b)), fc.FA_OP_BWD, None)].count_nonzero()) == 0
    # (2) the same gradient from separate ordinary calls, dq blocks added in fp32
    dq_a, dq_b = torch.zeros_like(q), torch.zeros_like(q)
    dk1r, dv1r, dk2r, dv2r = [torch.zeros_like(k) for _ in range(4)]
    fc.bwd(dout, q, k, v, out, lse, dq_a, dk1r, dv1r, cu, cu, 2 * c, c, 0.0, scale, False, False, None, seqused_k=half)
    fc.bwd(dout, q, k, v, out, lse[:, :, c:].contiguous(), dq_b, dk2r, dv2r, cu + c, cu + c, c, c, 0.0, scale, False, False, None,
           seqused_q=half, seqused_k=half)
    want = dq_a.float()
    want.view(b, 2 * c, h, d)[:, c:] += dq_b.view(b, 2 * c, h, d)[:, c:].float()
    assert torch.isfinite(dq.float()).all()
    assert torch.allclose(dq.float(), want, rtol=2e-2, atol=2e-3 * max(1.0, want.abs().max().item()))
    for a, r in ((dk1, dk1r), (dv1, dv1r), (dk2, dk2r), (dv2, dv2r)):
        assert torch.equal(a, r)
