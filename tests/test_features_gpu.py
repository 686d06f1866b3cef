"""GPU tests beyond the reference's grid: committed golden fixtures, bit-exact dropout stream, edge cases,
blocksparse, modules, varlen pack/unpack kernels, error behaviour, and size-independent properties at
BASELINE.json's full sizes."""
import math

import pytest
import torch

import fa_test_utils as U
from oracle import attention_ref as orc

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _fi():
    from flash_attn import flash_attn_interface as fi
    return fi


@pytest.mark.parametrize("idx", range(8))
def test_golden_fixtures(idx):
    """Kernel vs outputs of the reference's own attention_ref (tests/golden/make_golden.py)."""
    fi = _fi()
    c = U.load_golden("attention_ref_cases.pt")[idx]
    q, k, v, dout = [c[n].to(DEV) for n in ("q", "k", "v", "dout")]
    qmask, kmask = c["qmask"].to(DEV), c["kmask"].to(DEV)
    P = U.generate_qkv(q, k, v, qmask, kmask)
    # p_dropout = 0 here: the fixture's dropout mask came from torch.rand, not from the kernel's Philox stream
    out_u = fi.flash_attn_unpadded_func(P.q_unpad, P.k_unpad, P.v_unpad, P.cu_q, P.cu_k, P.max_q, P.max_k, 0.0,
                                        causal=c["causal"])
    if c["p"] == 0.0:
        out = P.pad_q(out_u)
        assert orc.max_abs_err(out.cpu(), c["out_ref"]) <= 2 * c["out_pt_maxerr"] + 1e-6
        dq, dk, dv = torch.autograd.grad(out_u, (P.q_unpad, P.k_unpad, P.v_unpad), dout[qmask])
        # the reference's 2x rule, with the 16-bit PyTorch gradient error the generator measured on the same inputs
        for got, ref, m, padf, pt_err in ((dq, c["dq"], qmask, P.pad_q, c["grad_pt_maxerr"][0]),
                                          (dk, c["dk"], kmask, P.pad_k, c["grad_pt_maxerr"][1]),
                                          (dv, c["dv"], kmask, P.pad_k, c["grad_pt_maxerr"][2])):
            ref = ref.to(DEV)
            assert orc.max_abs_err(padf(got), ref * m[:, :, None, None]) <= 2 * pt_err + 1e-6


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("d", [64, 128])
def test_dropout_mask_is_the_documented_philox_stream(d, causal):
    """Integer work is bit-exact: the sign pattern of S_dmask equals the host evaluation of the RNG."""
    fi = _fi()
    b, h, sq, sk, p = 2, 2, 150, 150, 0.3
    torch.manual_seed(3)
    q = torch.randn(b * sq, h, d, device=DEV, dtype=torch.float16)
    k = torch.randn(b * sk, h, d, device=DEV, dtype=torch.float16)
    v = torch.randn(b * sk, h, d, device=DEV, dtype=torch.float16)
    cu = U.cu_seqlens_full(b, sq, DEV)
    gen = torch.cuda.default_generators[torch.cuda.current_device()]
    seed, offset = gen.initial_seed(), gen.get_offset()
    out, lse, S = fi.flash_attn_unpadded_func(q, k, v, cu, cu, sq, sk, p, causal=causal, return_attn_probs=True)
    assert gen.get_offset() == offset + 4
    keep, probs = U.decode_s_dmask(S, sq, sk)
    expect = U.expected_dropout_mask(seed, offset, b, h, sq, sk, p).to(DEV)
    visible = probs > 0
    assert torch.equal(keep & visible, expect & visible)
    # and the main (tcgen05) kernel used exactly that mask: O matches the oracle evaluated with it
    qp, kp, vp = [t.reshape(b, -1, h, d) for t in (q, k, v)]
    out_ref, _ = orc.attention_ref(qp, kp, vp, None, None, p, expect, causal=causal)
    out_pt, _ = orc.attention_ref(qp, kp, vp, None, None, p, expect, causal=causal, upcast=False, reorder_ops=True)
    assert orc.within_2x(out.reshape(b, sq, h, d), out_ref, out_pt)


def test_p_zero_consumes_no_rng():
    fi = _fi()
    gen = torch.cuda.default_generators[torch.cuda.current_device()]
    q = torch.randn(64, 1, 64, device=DEV, dtype=torch.float16)
    cu = torch.tensor([0, 64], device=DEV, dtype=torch.int32)
    before = gen.get_offset()
    fi.flash_attn_unpadded_func(q, q, q, cu, cu, 64, 64, 0.0)
    assert gen.get_offset() == before   # fmha_api.cpp:231


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_ragged_and_empty_sequences(dtype):
    """Empty sequences, length-1 sequences, lengths straddling tile boundaries, seqlen_q != seqlen_k."""
    fi = _fi()
    h, d = 3, 64
    ql = [0, 1, 127, 128, 129, 255, 256, 257, 300]
    kl = [5, 0, 128, 1, 257, 3, 129, 511, 300]
    torch.manual_seed(0)
    cu_q = torch.tensor([0] + list(torch.tensor(ql).cumsum(0)), device=DEV, dtype=torch.int32)
    cu_k = torch.tensor([0] + list(torch.tensor(kl).cumsum(0)), device=DEV, dtype=torch.int32)
    q = torch.randn(sum(ql), h, d, device=DEV, dtype=dtype, requires_grad=True)
    k = torch.randn(sum(kl), h, d, device=DEV, dtype=dtype, requires_grad=True)
    v = torch.randn(sum(kl), h, d, device=DEV, dtype=dtype, requires_grad=True)
    for causal in (False, True):
        out, lse, _ = fi.flash_attn_unpadded_func(q, k, v, cu_q, cu_k, max(ql), max(kl), 0.0, causal=causal,
                                                  return_attn_probs=True)
        g = torch.randn_like(out)
        dq, dk, dv = torch.autograd.grad(out, (q, k, v), g)
        assert not torch.isnan(out).any() and not torch.isnan(dq).any() and not torch.isnan(dk).any()
        for i, (a, b) in enumerate(zip(ql, kl)):
            qs, ks = slice(cu_q[i].item(), cu_q[i + 1].item()), slice(cu_k[i].item(), cu_k[i + 1].item())
            if a == 0:
                continue
            if b == 0:   # no keys: zero output, lse = -inf, zero gradients (appendix A item 5)
                assert out[qs].abs().max().item() == 0 and dq[qs].abs().max().item() == 0
                assert bool(torch.isinf(lse[i, :, :a]).all())
                continue
            qi, ki, vi = [t.detach()[None].clone().requires_grad_() for t in (q[qs], k[ks], v[ks])]
            ref, _ = orc.attention_ref(qi, ki, vi, causal=causal)
            pt, _ = orc.attention_ref(qi, ki, vi, causal=causal, upcast=False, reorder_ops=True)
            assert orc.within_2x(out[qs][None], ref, pt, floor=1e-5), (i, a, b, causal)
            gr = torch.autograd.grad(ref, (qi, ki, vi), g[qs][None])
            gp = torch.autograd.grad(pt, (qi, ki, vi), g[qs][None])
            for got, r, p_ in zip((dq[qs], dk[ks], dv[ks]), gr, gp):
                assert orc.within_2x(got[None], r, p_, floor=1e-3 * max(1.0, r.abs().max().item())), (i, a, b, causal)


def test_strided_views_and_inplace_gradients():
    """q/k/v given as strided views of one packed tensor; bwd writes into strided views of dqkv."""
    import flash_attn_cuda
    b, s, h, d = 4, 200, 2, 64
    torch.manual_seed(1)
    qkv = torch.randn(b * s, 3, h, d, device=DEV, dtype=torch.bfloat16)
    cu = U.cu_seqlens_full(b, s, DEV)
    scale = d ** -0.5
    o1, lse1 = flash_attn_cuda.fwd(qkv[:, 0], qkv[:, 1], qkv[:, 2], cu, cu, s, s, 0.0, scale, False, True, False, None)
    o2, lse2 = flash_attn_cuda.fwd(qkv[:, 0].contiguous(), qkv[:, 1].contiguous(), qkv[:, 2].contiguous(), cu, cu, s, s,
                                   0.0, scale, False, True, False, None)
    assert torch.equal(o1, o2) and torch.equal(lse1[:, :, :s], lse2[:, :, :s])   # the lse tail beyond s is uninitialised
    dout = torch.randn_like(o1)
    dqkv = torch.full_like(qkv, float("nan"))
    flash_attn_cuda.bwd(dout, qkv[:, 0], qkv[:, 1], qkv[:, 2], o1, lse1, dqkv[:, 0], dqkv[:, 1], dqkv[:, 2], cu, cu,
                        s, s, 0.0, scale, False, True, None)
    assert not torch.isnan(dqkv).any()   # fully overwritten


def test_tcgen05_matches_simt_cross_check_kernels():
    """Two independent implementations inside the library agree (size beyond what the CPU oracle can do)."""
    import flash_attn_cuda
    fi = _fi()
    b, s, h, d = 2, 4096, 4, 128
    torch.manual_seed(2)
    q, k, v = [torch.randn(b * s, h, d, device=DEV, dtype=torch.bfloat16) for _ in range(3)]
    cu = U.cu_seqlens_full(b, s, DEV)
    res = {}
    gen = torch.cuda.default_generators[torch.cuda.current_device()]
    import contextlib
    for impl in (flash_attn_cuda.FA_IMPL_AUTO, flash_attn_cuda.FA_IMPL_SIMT):
        gen.set_offset(0)   # same Philox (seed, offset) for both implementations
        with (U.simt_kernels() if impl == flash_attn_cuda.FA_IMPL_SIMT else contextlib.nullcontext()):
            res[impl] = fi.flash_attn_unpadded_func(q, k, v, cu, cu, s, s, 0.1, causal=True, return_attn_probs=True)
    (o_a, lse_a, _), (o_b, lse_b, _) = res[0], res[2]
    assert (lse_a[:, :, :s] - lse_b[:, :, :s]).abs().max().item() < 1e-3
    assert (o_a.float() - o_b.float()).abs().max().item() < 3e-2


def test_linearity_in_v_at_full_size():
    """Size-independent property at BASELINE's d=128 config (b16 h12 S4096): O is linear in V."""
    fi = _fi()
    b, s, h, d = 16, 4096, 12, 128
    torch.manual_seed(5)
    q, k, v1, v2 = [torch.randn(b * s, h, d, device=DEV, dtype=torch.bfloat16) for _ in range(4)]
    cu = U.cu_seqlens_full(b, s, DEV)
    o1 = fi.flash_attn_unpadded_func(q, k, v1, cu, cu, s, s, 0.0, causal=True).float()
    o2 = fi.flash_attn_unpadded_func(q, k, v2, cu, cu, s, s, 0.0, causal=True).float()
    o12 = fi.flash_attn_unpadded_func(q, k, (v1.float() + v2.float()).to(torch.bfloat16), cu, cu, s, s, 0.0, causal=True).float()
    assert (o12 - (o1 + o2)).abs().max().item() < 6e-2
    # first row of every sequence attends only to key 0 under the causal mask: O[0] == V[0]
    first = torch.arange(b, device=DEV) * s
    assert torch.equal(fi.flash_attn_unpadded_func(q, k, v1, cu, cu, s, s, 0.0, causal=True)[first], v1[first])


def test_error_behaviour():
    """RuntimeError like TORCH_CHECK (fmha_api.cpp:131-170), never exit()."""
    import flash_attn_cuda
    cu = torch.tensor([0, 16], device=DEV, dtype=torch.int32)
    q = torch.randn(16, 2, 64, device=DEV, dtype=torch.float16)
    args = (cu, cu, 16, 16, 0.0, 0.125, False, False, False, None)
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(q.float(), q.float(), q.float(), *args)          # dtype
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(q[..., :12], q[..., :12], q[..., :12], *args)    # d % 8
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(q, q.bfloat16(), q, *args)                       # mismatched dtypes
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(q, q, q, cu.long(), cu, *args[2:])               # cu_seqlens dtype
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(q.transpose(1, 2), q, q, *args)                  # last stride != 1
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(q, q, q, cu, cu, 16, 16, 1.0, 0.125, False, False, False, None)   # p_dropout >= 1
    big = torch.randn(16, 2, 136, device=DEV, dtype=torch.float16)
    with pytest.raises(RuntimeError):
        flash_attn_cuda.fwd(big, big, big, *args)                            # d > 128
    out, lse = flash_attn_cuda.fwd(q, q, q, *args)                           # context still healthy afterwards
    assert torch.isfinite(out).all()


# ------------------------------------------------------------------------------------------------------------
# blocksparse
# ------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("idx", range(3))
def test_blocksparse_golden(idx):
    """vs outputs of the reference's attention_blocksparse_ref (tests/test_flash_attn.py:189-215)."""
    from flash_attn.flash_blocksparse_attn_interface import flash_blocksparse_attn_func
    c = U.load_golden("blocksparse_cases.pt")["attn"][idx]
    qkv, amask, bmask = c["qkv"].to(DEV), c["amask"].to(DEV), c["blockmask"].to(DEV)
    b, s = c["b"], c["s"]
    xu, idx_, cu, mx = __import__("flash_attn.bert_padding", fromlist=["unpad_input"]).unpad_input(
        qkv.reshape(b, s, -1), amask)
    out_u = flash_blocksparse_attn_func(xu.reshape(-1, 3, c["h"], c["d"]), cu, bmask, 0.0, mx)
    out = U.unpack_rows(out_u, idx_, b, s)
    # the kernel runs without dropout (the fixture's mask came from torch.rand, not from the kernel's Philox stream): the
    # p = 0 fixtures are compared with the reference's own output, the p > 0 one with the oracle (pinned to the reference
    # on this very fixture by tests/test_oracle_cpu.py) evaluated at p = 0 -- both under the 2x rule
    ref32 = c["out32"].to(DEV) if c["p"] == 0.0 else orc.attention_blocksparse_ref(qkv.float(), bmask, amask, 0.0, None)[0]
    pt16, _ = orc.attention_blocksparse_ref(qkv, bmask, amask, 0.0, None, upcast=False)
    assert orc.max_abs_err(out, ref32) <= 2 * orc.max_abs_err(pt16, ref32) + 1e-6


@pytest.mark.parametrize("dropout_p", [0.0, 0.17])
@pytest.mark.parametrize("d", [32, 64, 128])
@pytest.mark.parametrize("seqlen", [256, 512, 1000, 2048])
def test_blocksparse_fwd_bwd(seqlen, d, dropout_p):
    from flash_attn.bert_padding import unpad_input
    from flash_attn.flash_blocksparse_attn_interface import flash_blocksparse_attn_func
    b, h = 8, 4
    dtype = torch.float16
    torch.manual_seed(0)
    s_r = (seqlen + 255) // 256 * 256
    qkv = torch.randn(b, seqlen, 3, h, d, device=DEV, dtype=dtype)
    blockmask = orc.generate_sparsity_mask(s_r, 0.4, device=DEV)
    blockmask[:, 0] = True   # every query row sees at least the first key block (keeps the fp reference finite)
    amask = orc.generate_random_padding_mask(seqlen, b, DEV, mode="random")
    xu, idx, cu, mx = unpad_input(qkv.reshape(b, seqlen, -1), amask)
    xu = xu.reshape(-1, 3, h, d).detach().requires_grad_()
    out_u, S, lse = flash_blocksparse_attn_func(xu, cu, blockmask, dropout_p, mx, return_attn_probs=True)
    out = U.unpack_rows(out_u, idx, b, seqlen)
    keep, probs = U.decode_s_dmask(S, seqlen, seqlen)
    qkv_f = qkv.detach().float().requires_grad_()
    # the reference's 2x rule (tests/test_flash_attn.py:407-418) with a 16-bit PyTorch baseline of the same math
    out_ref32, attn_ref32 = orc.attention_blocksparse_ref(qkv_f.float(), blockmask, amask, dropout_p, keep)
    qkv_p = qkv.detach().clone().requires_grad_()
    out_pt, attn_pt = orc.attention_blocksparse_ref(qkv_p, blockmask, amask, dropout_p, keep, upcast=False)
    assert orc.max_abs_err(out, out_ref32) <= 2 * orc.max_abs_err(out_pt, out_ref32) + 1e-6
    assert orc.max_abs_err(probs, attn_ref32) <= 2 * orc.max_abs_err(attn_pt, attn_ref32) + 1e-6
    if dropout_p > 0:
        dense = orc.expand_blockmask(blockmask, seqlen)
        visible = amask[:, None, :, None] & amask[:, None, None, :] & dense[None, None]
        frac = ((~keep) & visible).sum().item() / (visible.sum().item() * h)
        assert 0.98 <= frac / dropout_p <= 1.02
    g = torch.randn_like(out)
    dqkv_u, = torch.autograd.grad(out_u, xu, g[amask])
    dqkv = U.unpack_rows(dqkv_u, idx, b, seqlen)
    dqkv_ref, = torch.autograd.grad(out_ref32, qkv_f, g.float())
    dqkv_pt, = torch.autograd.grad(out_pt, qkv_p, g)
    for i, name in enumerate(("dQ", "dK", "dV")):
        err, pt_err = orc.max_abs_err(dqkv[:, :, i], dqkv_ref[:, :, i]), orc.max_abs_err(dqkv_pt[:, :, i], dqkv_ref[:, :, i])
        assert err <= 2 * pt_err + 1e-6, f"{name}: error {err:.3e} > 2 x PyTorch error {pt_err:.3e}"


def test_blocksparse_skips_masked_blocks_butterfly():
    """Butterfly layout at S=4096: result equals dense attention restricted to the layout."""
    from flash_attn.flash_blocksparse_attention import ButterflySparsityConfig, FlashBlocksparseAttention
    b, s, h, d = 2, 4096, 4, 64
    torch.manual_seed(0)
    mod = FlashBlocksparseAttention(ButterflySparsityConfig(), max_seq_length=s).to(DEV)
    qkv = torch.randn(b, s, 3, h, d, device=DEV, dtype=torch.float16)
    out, _ = mod(qkv)
    amask = torch.ones(b, s, dtype=torch.bool, device=DEV)
    ref, _ = orc.attention_blocksparse_ref(qkv.float(), mod.layout.bool(), amask, 0.0, None)
    assert orc.max_abs_err(out, ref) < 3e-3


# ------------------------------------------------------------------------------------------------------------
# modules and varlen pack / unpack kernels
# ------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("causal", [False, True])
def test_flash_attention_module_three_input_modes(causal, dtype):
    from flash_attn.bert_padding import unpad_input
    from flash_attn.flash_attention import FlashAttention
    b, s, h, d = 4, 300, 4, 64
    torch.manual_seed(0)
    qkv = torch.randn(b, s, 3, h, d, device=DEV, dtype=dtype)
    mod = FlashAttention().eval()
    out_dense, none = mod(qkv, causal=causal)
    assert none is None
    ref, _ = orc.attention_qkvpacked_ref(qkv, None, causal=causal)
    pt, _ = orc.attention_qkvpacked_ref(qkv, None, causal=causal, upcast=False, reorder_ops=True)
    assert orc.within_2x(out_dense, ref, pt)
    mask = orc.generate_random_padding_mask(s, b, DEV, mode="third")
    out_masked, _ = mod(qkv, key_padding_mask=mask, causal=causal)
    ref, _ = orc.attention_qkvpacked_ref(qkv, mask, causal=causal)
    pt, _ = orc.attention_qkvpacked_ref(qkv, mask, causal=causal, upcast=False, reorder_ops=True)
    assert orc.within_2x(out_masked, ref, pt)
    xu, idx, cu, mx = unpad_input(qkv.reshape(b, s, -1), mask)
    out_unp, _ = mod(xu.reshape(-1, 3, h, d), cu_seqlens=cu, max_s=mx, causal=causal)
    assert torch.equal(U.unpack_rows(out_unp, idx, b, s), out_masked)


@pytest.mark.parametrize("rotary", [None, "1d", "2d"])
def test_flash_mha_module(rotary):
    from flash_attn.flash_attention import FlashMHA
    b, s, e, h = 2, 256, 256, 4
    torch.manual_seed(0)
    mha = FlashMHA(e, h, causal=True, use_rotary_emb=rotary, device=DEV, dtype=torch.float16).eval()
    x = torch.randn(b, s, e, device=DEV, dtype=torch.float16, requires_grad=True)
    y, w = mha(x)
    assert w is None and y.shape == x.shape
    qkv = mha.Wqkv(x).reshape(b, s, 3, h, e // h)
    if rotary:
        q, k, v = qkv.unbind(2)
        q, k = mha.rotary_emb(q, k, seq_dimension=-3)
        qkv = torch.stack([q, k, v], 2)
    ctx, _ = orc.attention_qkvpacked_ref(qkv, None, causal=True)
    y_ref = mha.out_proj(ctx.reshape(b, s, e))
    assert orc.max_abs_err(y, y_ref) < 2e-2
    # backward through Wqkv -> (rotary) -> attention -> out_proj against the fp32 oracle, 2x rule on dL/dx
    gy = torch.randn_like(y)
    (gx,) = torch.autograd.grad(y, x, gy)

    def torch_path(dtype):
        xx = x.detach().to(dtype).requires_grad_()
        w = {n: p_.detach().to(dtype) for n, p_ in mha.named_parameters()}
        qkv_ = torch.nn.functional.linear(xx, w["Wqkv.weight"], w["Wqkv.bias"]).reshape(b, s, 3, h, e // h)
        if rotary:
            q_, k_, v_ = qkv_.unbind(2)
            q_, k_ = mha.rotary_emb(q_, k_, seq_dimension=-3)
            qkv_ = torch.stack([q_, k_, v_], 2)
        c_, _ = orc.attention_qkvpacked_ref(qkv_, None, causal=True, upcast=(dtype == torch.float32),
                                            reorder_ops=(dtype != torch.float32))
        y_ = torch.nn.functional.linear(c_.reshape(b, s, e), w["out_proj.weight"], w["out_proj.bias"])
        (g_,) = torch.autograd.grad(y_, xx, gy.to(dtype))
        return g_
    g32, g16 = torch_path(torch.float32), torch_path(torch.float16)
    assert orc.max_abs_err(gx, g32) <= 2 * orc.max_abs_err(g16, g32) + 1e-6


@pytest.mark.parametrize("p", [0.0, 0.17])
def test_cuda_graph_capture_and_replay(p):
    """Forward + backward are stream-ordered with no host synchronisation (fmha_api.cpp:132), so one training step can be
    captured in a CUDA graph and replayed on new data in the same buffers -- with dropout too: a captured call reads its
    Philox offset from a device counter (fa_params.philox_offset_ptr, the reference's PhiloxCudaState, fmha.h:39), so
    every replay draws a fresh mask and the captured backward regenerates the mask of its own forward."""
    from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
    b, s, h, d = 3, 384, 4, 64
    torch.manual_seed(0)
    qkv = torch.randn(b * s, 3, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    dout = torch.randn(b * s, h, d, device=DEV, dtype=torch.float16)
    cu = torch.arange(0, (b + 1) * s, s, device=DEV, dtype=torch.int32)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2):   # warm-up (also creates the device-side Philox counter)
            out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=True)
            torch.autograd.grad(out, qkv, dout)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        g_out, g_lse, g_S = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, p, causal=True, return_attn_probs=True)
        (g_dqkv,) = torch.autograd.grad(g_out, qkv, dout)
    with torch.no_grad():
        qkv.copy_(torch.randn_like(qkv))
        dout.copy_(torch.randn_like(dout))
    graph.replay()
    torch.cuda.synchronize()
    keep1, probs = U.decode_s_dmask(g_S, s, s)
    out1, dqkv1 = g_out.clone(), g_dqkv.clone()
    # the replayed step equals the oracle evaluated with the mask the replay drew (forward AND backward: same mask)
    x = qkv.detach().reshape(b, s, 3, h, d)
    xf = x.float().requires_grad_()
    ref, _ = orc.attention_qkvpacked_ref(xf, None, p, keep1, causal=True)
    (gref,) = torch.autograd.grad(ref, xf, dout.reshape(b, s, h, d).float())
    xp = x.clone().requires_grad_()
    pt, _ = orc.attention_qkvpacked_ref(xp, None, p, keep1, causal=True, upcast=False, reorder_ops=True)
    (gpt,) = torch.autograd.grad(pt, xp, dout.reshape(b, s, h, d))
    assert orc.within_2x(out1.reshape(b, s, h, d), ref, pt)
    assert orc.max_abs_err(dqkv1.reshape(b, s, 3, h, d), gref) <= 2 * orc.max_abs_err(gpt, gref) + 1e-6
    graph.replay()
    torch.cuda.synchronize()
    keep2, _ = U.decode_s_dmask(g_S, s, s)
    if p == 0.0:
        assert torch.equal(g_out, out1) and torch.equal(g_dqkv[:, 1:], dqkv1[:, 1:])      # O, dK, dV: bitwise deterministic
        assert orc.max_abs_err(g_dqkv[:, 0], dqkv1[:, 0]) < 1e-2                          # dQ: fp32 reduction order may differ
    else:
        visible = probs > 0
        frac_same = ((keep1 == keep2) & visible).sum().item() / visible.sum().item()
        assert 0.6 < frac_same < 0.85      # fresh mask: two independent masks at p = 0.17 agree on 1 - 2p(1-p) = 72 % of the entries
        assert not torch.equal(g_out, out1)


@pytest.mark.parametrize("mode", ["1d", "2d"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("d", [32, 64, 128])
def test_fused_rotary_kernel_matches_reference_modules(mode, dtype, d):
    """fa_rotary_qk (in place, and its transposed backward) against the reference's elementwise formulation
    (flash_attn/rotary.py) evaluated in fp32; tolerance: the reference's own rule, error <= 2x the error of the
    16-bit PyTorch path."""
    from flash_attn.rotary import RotaryEmbedding, RotaryEmbedding2D, apply_rotary_qkvpacked_, rotary_tables
    b, s, h = 3, 144, 5
    torch.manual_seed(0)
    qkv = torch.randn(b, s, 3, h, d, device=DEV, dtype=dtype)
    mod = (RotaryEmbedding(d) if mode == "1d" else RotaryEmbedding2D(d)).to(DEV)
    q32, k32 = mod(qkv[:, :, 0].float(), qkv[:, :, 1].float(), seq_dimension=-3)
    q16, k16 = mod(qkv[:, :, 0], qkv[:, :, 1], seq_dimension=-3)
    cos, sin = rotary_tables(mode, s, d, DEV)
    x = qkv.reshape(b, s, -1).clone().requires_grad_()
    y = apply_rotary_qkvpacked_(x * 1.0, cos, sin, h)       # x * 1.0: a fresh non-leaf tensor, as in FlashMHA
    y5 = y.reshape(b, s, 3, h, d)
    assert torch.equal(y5[:, :, 2], qkv[:, :, 2])                               # v untouched
    for got, ref, pt in ((y5[:, :, 0], q32, q16), (y5[:, :, 1], k32, k16)):
        assert orc.max_abs_err(got, ref) <= 2 * orc.max_abs_err(pt, ref) + 1e-6
    # backward = transposed rotation of the q, k parts, identity on v
    g = torch.randn_like(y)
    (gx,) = torch.autograd.grad(y, x, g)
    xr = qkv.float().clone().requires_grad_()
    qr, kr = mod(xr[:, :, 0], xr[:, :, 1], seq_dimension=-3)
    yr = torch.stack([qr, kr, xr[:, :, 2]], dim=2)
    (gr,) = torch.autograd.grad(yr, xr, g.reshape(b, s, 3, h, d).float())
    x16 = qkv.clone().requires_grad_()
    q_, k_ = mod(x16[:, :, 0], x16[:, :, 1], seq_dimension=-3)
    (g16,) = torch.autograd.grad(torch.stack([q_, k_, x16[:, :, 2]], dim=2), x16, g.reshape(b, s, 3, h, d))
    assert orc.max_abs_err(gx.reshape(b, s, 3, h, d), gr) <= 2 * orc.max_abs_err(g16, gr) + 1e-6
    assert torch.equal(gx.reshape(b, s, 3, h, d)[:, :, 2], g.reshape(b, s, 3, h, d)[:, :, 2])


def test_fused_rotary_rejects_bad_arguments():
    import flash_attn_cuda
    x = torch.randn(8, 3, 2, 32, device=DEV, dtype=torch.float16)
    cos = torch.zeros(8, 16, device=DEV)
    with pytest.raises(RuntimeError):
        flash_attn_cuda.rotary_qk(x, x, cos[:, :8].contiguous(), cos[:, :8].contiguous(), 8)
    with pytest.raises(RuntimeError):
        flash_attn_cuda.rotary_qk(x.float(), x.float(), cos, cos, 8)
    with pytest.raises(RuntimeError):
        flash_attn_cuda.rotary_qk(x.transpose(1, 2), x.transpose(1, 2), cos, cos, 8)


@pytest.mark.parametrize("width", [2, 8, 64, 768, 1000])
@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
def test_bert_padding_kernels_match_torch(width, dtype):
    from flash_attn.bert_padding import index_first_axis, pad_input, unpad_input
    b, s = 16, 512
    torch.manual_seed(0)
    x = torch.randn(b, s, width, device=DEV, dtype=dtype, requires_grad=True)
    mask = orc.generate_random_padding_mask(s, b, DEV, mode="third")
    xu, idx, cu, mx = unpad_input(x, mask)
    assert torch.equal(xu, x.reshape(b * s, width)[idx])
    assert torch.equal(cu[1:].long(), mask.sum(-1).cumsum(0)) and mx == mask.sum(-1).max().item()
    back = pad_input(xu, idx, b, s)
    assert torch.equal(back, x * mask[..., None])
    back.backward(torch.ones_like(back))
    assert torch.equal(x.grad, mask[..., None].expand_as(x).to(dtype))
    x3 = torch.randn(b * s, 3, 4, 8, device=DEV, dtype=dtype)
    assert torch.equal(index_first_axis(x3, idx), x3[idx])


def test_blocksparse_actually_skips_masked_blocks():
    """Not just the right result: the butterfly layout must cost a fraction of the all-ones layout through the very same
    entry point (forward + backward), i.e. masked (16 x 256) blocks are skipped, not computed and thrown away."""
    from flash_attn.flash_blocksparse_attention import ButterflySparsityConfig
    from flash_attn.flash_blocksparse_attn_interface import convert_blockmask, flash_blocksparse_attn_func
    b, s, h, d = 4, 8192, 8, 64
    torch.manual_seed(0)
    qkv = torch.randn(b * s, 3, h, d, device=DEV, dtype=torch.float16, requires_grad=True)
    dout = torch.randn(b * s, h, d, device=DEV, dtype=torch.float16)
    cu = torch.arange(0, (b + 1) * s, s, device=DEV, dtype=torch.int32)
    layout = ButterflySparsityConfig().make_layout(s).to(DEV)
    density = float(layout.float().mean())
    masks = {"butterfly": convert_blockmask(layout, causal=False), "dense": convert_blockmask(torch.ones_like(layout), causal=False)}

    def step(mask):
        out = flash_blocksparse_attn_func(qkv, cu, mask, 0.0, s, convert_mask=False)
        torch.autograd.grad(out, qkv, dout)

    ms = {}
    for name, mask in masks.items():
        for _ in range(2):
            step(mask)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            step(mask)
        e1.record()
        torch.cuda.synchronize()
        ms[name] = e0.elapsed_time(e1) / 5
    ratio = ms["butterfly"] / ms["dense"]
    assert density < 0.5, density
    assert ratio < min(0.8, density + 0.35), f"butterfly (density {density:.2f}) takes {ratio:.2f} of the dense time: nothing is skipped"
