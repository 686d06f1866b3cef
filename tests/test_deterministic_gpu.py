"""FA_FLAG_DETERMINISTIC (include/fa_b200.h): dQ accumulated in key-block order.

By default the key blocks of a sequence add their parts of dQ with fp32 reductions in arrival order, so dQ moves in the
last bits from run to run; dK, dV and the forward never do (the reference's only determinism test is on the forward,
tests/test_flash_attn.py:652-664).  With the flag -- `flash_attn_cuda.set_deterministic(True)` or torch's global
`use_deterministic_algorithms` -- dQ has to be bit-identical across runs, equal to the default mode's dQ up to fp32
summation order, and dK / dV bit-identical to the default mode's."""
import pytest
import torch

import fa_test_utils as U

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _grads(fn, leaves, dout, reps):
    out = fn()
    return [torch.autograd.grad(out, leaves, dout, retain_graph=True) for _ in range(reps)]


def _ragged_cu(lens):
    return torch.nn.functional.pad(torch.tensor(lens, dtype=torch.int64).cumsum(0), (1, 0)).to(device=DEV, dtype=torch.int32)


CASES = [
    # d, dtype, causal, p, lengths
    (128, torch.bfloat16, True, 0.0, [2048] * 4),
    (128, torch.bfloat16, False, 0.0, [1024, 777, 1500, 128, 1]),
    (64, torch.float16, True, 0.17, [512, 300, 1025, 64]),
    (64, torch.float16, False, 0.0, [2048] * 3),
    (96, torch.bfloat16, True, 0.0, [1300, 257]),
    (40, torch.float16, False, 0.17, [600, 900]),
]


@pytest.mark.parametrize("d,dtype,causal,p,lens", CASES)
def test_dq_is_bitwise_reproducible_and_matches_the_default_mode(d, dtype, causal, p, lens):
    import flash_attn_cuda
    from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
    h = 4
    torch.manual_seed(1)
    cu, total, mx = _ragged_cu(lens), sum(lens), max(lens)
    qkv = torch.randn(total, 3, h, d, device=DEV, dtype=dtype, requires_grad=True)
    dout = torch.randn(total, h, d, device=DEV, dtype=dtype)
    fn = lambda: flash_attn_unpadded_qkvpacked_func(qkv, cu, mx, p, causal=causal)
    torch.manual_seed(7)                       # same dropout stream for both modes
    (ref,) = _grads(fn, qkv, dout, 1)[0]
    prev = flash_attn_cuda.set_deterministic(True)
    try:
        torch.manual_seed(7)
        runs = [g[0] for g in _grads(fn, qkv, dout, 4)]
    finally:
        flash_attn_cuda.set_deterministic(prev)
    for r in runs[1:]:
        assert torch.equal(r, runs[0]), "dQ / dK / dV differ between two deterministic runs"
    assert torch.equal(runs[0][:, 1:], ref[:, 1:]), "dK / dV must not depend on the mode"
    err = (runs[0][:, 0].float() - ref[:, 0].float()).abs().max().item()
    scale = ref[:, 0].float().abs().max().item()
    assert err <= 2e-2 * scale, f"dQ differs from the default mode by {err} (max |dQ| {scale})"   # 16-bit rounding of two fp32 sums


def test_cross_attention_lengths_and_the_torch_switch():
    """seqlen_q != seqlen_k (more key blocks than query tiles and the reverse), switched on through torch's global flag"""
    import flash_attn_cuda
    from flash_attn.flash_attn_interface import flash_attn_unpadded_kvpacked_func
    h, d = 3, 128
    torch.manual_seed(2)
    lq, lk = [700, 130, 2048], [2048, 1500, 96]
    cq, ck = _ragged_cu(lq), _ragged_cu(lk)
    q = torch.randn(sum(lq), h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    kv = torch.randn(sum(lk), 2, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    dout = torch.randn_like(q)
    assert flash_attn_cuda._deterministic is None and not flash_attn_cuda._want_deterministic()
    for causal in (False, True):
        fn = lambda: flash_attn_unpadded_kvpacked_func(q, kv, cq, ck, max(lq), max(lk), 0.0, causal=causal)
        ref = _grads(fn, (q, kv), dout, 1)[0]
        torch.use_deterministic_algorithms(True, warn_only=True)
        try:
            assert flash_attn_cuda._want_deterministic()
            runs = _grads(fn, (q, kv), dout, 3)
        finally:
            torch.use_deterministic_algorithms(False)
        for r in runs[1:]:
            assert torch.equal(r[0], runs[0][0]) and torch.equal(r[1], runs[0][1])
        assert torch.equal(runs[0][1], ref[1])
        assert (runs[0][0].float() - ref[0].float()).abs().max().item() <= 2e-2 * ref[0].float().abs().max().item()


def test_full_size_and_graph_replay():
    """the headline shape (every SM busy, 32 key blocks taking turns on every query tile), eagerly and from a CUDA graph"""
    import flash_attn_cuda
    from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
    b, s, h, d = 16, 4096, 12, 128
    torch.manual_seed(3)
    qkv = torch.randn(b * s, 3, h, d, device=DEV, dtype=torch.bfloat16, requires_grad=True)
    dout = torch.randn(b * s, h, d, device=DEV, dtype=torch.bfloat16)
    cu = U.cu_seqlens_full(b, s, DEV)
    prev = flash_attn_cuda.set_deterministic(True)
    try:
        for causal in (True, False):
            out = flash_attn_unpadded_qkvpacked_func(qkv, cu, s, 0.0, causal=causal)
            a, = torch.autograd.grad(out, qkv, dout, retain_graph=True)
            c, = torch.autograd.grad(out, qkv, dout, retain_graph=True)
            assert torch.equal(a, c)
        # forward + backward captured together (autograd runs a backward on its forward's stream), non-causal as the last eager step
        # (a fresh leaf made on the side stream: the eager steps above tied qkv's gradient sink to the default stream, and
        # autograd would synchronise with that stream in the middle of the capture)
        del out, a
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            leaf = qkv.detach().clone().requires_grad_()

        def step():
            o = flash_attn_unpadded_qkvpacked_func(leaf, cu, s, 0.0, causal=False)
            return torch.autograd.grad(o, leaf, dout)[0]
        with torch.cuda.stream(side):
            step()
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            g = step()
        for _ in range(2):
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(g, c)
    finally:
        flash_attn_cuda.set_deterministic(prev)
