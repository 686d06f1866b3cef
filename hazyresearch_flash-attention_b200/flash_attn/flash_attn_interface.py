"""Functional attention API: thin torch.autograd Functions over the C-ABI library (via `flash_attn_cuda`).

Drop-in for the reference's flash_attn/flash_attn_interface.py: same function names, argument order, defaults
and return values (reference lines cited per function).  The native calls replace csrc/flash_attn's mha_fwd
and the (absent-from-snapshot) mha_bwd.
"""
import torch

import flash_attn_cuda


def _get_block_size(device, head_dim, is_dropout):
    """K-block size the REFERENCE kernel would use on `device` (reference :8-14).  Kept importable for callers
    that decode the reference's S_dmask layout; this build's S_dmask is already in natural layout, so nothing
    in this package depends on it."""
    assert head_dim % 8 == 0 and head_dim <= 128
    cc = torch.cuda.get_device_capability(device)
    if head_dim <= 64:
        return 128 if (cc == (7, 5) and is_dropout) else 256
    return 256 if (cc == (8, 0) and not is_dropout) else 128


def _flash_attn_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                        softmax_scale, causal, return_softmax, alibi_slopes=None):
    """reference :17-26.  -> out, softmax_lse, S_dmask or None, and the dropout stream the call used (handed to the
    backward instead of the reference's global RNG-state round trip, reference :44, :61-63)."""
    res, rng = flash_attn_cuda.fwd_ex(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                                      softmax_scale, causal, return_softmax, None, alibi_slopes=alibi_slopes)
    return res[0], res[1], (res[2] if return_softmax else None), rng


def _flash_attn_backward(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k,
                         max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale, causal, rng=None, alibi_slopes=None):
    """reference :29-36"""
    softmax_d = flash_attn_cuda.bwd(dout, q, k, v, out, softmax_lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k,
                                    max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale, False, causal, rng,
                                    alibi_slopes=alibi_slopes)
    return dq, dk, dv, softmax_d


def _default_scale(softmax_scale, head_dim):
    return head_dim ** (-0.5) if softmax_scale is None else softmax_scale


def _save_hparams(ctx, dropout_p, max_seqlen_q, max_seqlen_k, softmax_scale, causal, rng, alibi_slopes):
    ctx.dropout_p, ctx.softmax_scale, ctx.causal = dropout_p, softmax_scale, causal
    ctx.max_seqlen_q, ctx.max_seqlen_k = max_seqlen_q, max_seqlen_k
    ctx.rng, ctx.alibi_slopes = rng, alibi_slopes


class FlashAttnQKVPackedFunc(torch.autograd.Function):
    """qkv: (total, 3, h, d).  The kernel reads the three strided views directly and the backward writes
    straight into views of one packed dqkv (reference :39-72)."""

    @staticmethod
    def forward(ctx, qkv, cu_seqlens, max_seqlen, dropout_p, softmax_scale, causal, return_softmax, alibi_slopes=None):
        softmax_scale = _default_scale(softmax_scale, qkv.shape[-1])
        q, k, v = qkv.unbind(1)      # three strided views in one call
        out, lse, s_dmask, rng = _flash_attn_forward(q, k, v, cu_seqlens, cu_seqlens, max_seqlen, max_seqlen, dropout_p,
                                                     softmax_scale, causal, return_softmax, alibi_slopes)
        ctx.save_for_backward(qkv, out, lse, cu_seqlens)
        _save_hparams(ctx, dropout_p, max_seqlen, max_seqlen, softmax_scale, causal, rng, alibi_slopes)
        return (out, lse, s_dmask) if return_softmax else out

    @staticmethod
    def backward(ctx, dout, *unused):
        qkv, out, lse, cu_seqlens = ctx.saved_tensors
        dqkv = torch.empty_like(qkv)
        (q, k, v), (dq, dk, dv) = qkv.unbind(1), dqkv.unbind(1)
        _flash_attn_backward(dout, q, k, v, out, lse, dq, dk, dv, cu_seqlens, cu_seqlens, ctx.max_seqlen_q, ctx.max_seqlen_k,
                             ctx.dropout_p, ctx.softmax_scale, ctx.causal, ctx.rng, ctx.alibi_slopes)
        return dqkv, None, None, None, None, None, None, None


class FlashAttnKVPackedFunc(torch.autograd.Function):
    """q: (total_q, h, d), kv: (total_k, 2, h, d) -- cross attention with seqlen_q != seqlen_k (reference :75-111)."""

    @staticmethod
    def forward(ctx, q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale,
                causal, return_softmax, alibi_slopes=None):
        softmax_scale = _default_scale(softmax_scale, q.shape[-1])
        out, lse, s_dmask, rng = _flash_attn_forward(q, kv[:, 0], kv[:, 1], cu_seqlens_q, cu_seqlens_k, max_seqlen_q,
                                                     max_seqlen_k, dropout_p, softmax_scale, causal, return_softmax,
                                                     alibi_slopes)
        ctx.save_for_backward(q, kv, out, lse, cu_seqlens_q, cu_seqlens_k)
        _save_hparams(ctx, dropout_p, max_seqlen_q, max_seqlen_k, softmax_scale, causal, rng, alibi_slopes)
        return (out, lse, s_dmask) if return_softmax else out

    @staticmethod
    def backward(ctx, dout, *unused):
        q, kv, out, lse, cu_seqlens_q, cu_seqlens_k = ctx.saved_tensors
        dq, dkv = torch.empty_like(q), torch.empty_like(kv)
        _flash_attn_backward(dout, q, kv[:, 0], kv[:, 1], out, lse, dq, dkv[:, 0], dkv[:, 1], cu_seqlens_q,
                             cu_seqlens_k, ctx.max_seqlen_q, ctx.max_seqlen_k, ctx.dropout_p,
                             ctx.softmax_scale, ctx.causal, ctx.rng, ctx.alibi_slopes)
        return dq, dkv, None, None, None, None, None, None, None, None, None


class FlashAttnFunc(torch.autograd.Function):
    """Separate q, k, v (reference :114-148)."""

    @staticmethod
    def forward(ctx, q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p, softmax_scale,
                causal, return_softmax, alibi_slopes=None):
        softmax_scale = _default_scale(softmax_scale, q.shape[-1])
        out, lse, s_dmask, rng = _flash_attn_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k,
                                                     dropout_p, softmax_scale, causal, return_softmax, alibi_slopes)
        ctx.save_for_backward(q, k, v, out, lse, cu_seqlens_q, cu_seqlens_k)
        _save_hparams(ctx, dropout_p, max_seqlen_q, max_seqlen_k, softmax_scale, causal, rng, alibi_slopes)
        return (out, lse, s_dmask) if return_softmax else out

    @staticmethod
    def backward(ctx, dout, *unused):
        q, k, v, out, lse, cu_seqlens_q, cu_seqlens_k = ctx.saved_tensors
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        _flash_attn_backward(dout, q, k, v, out, lse, dq, dk, dv, cu_seqlens_q, cu_seqlens_k,
                             ctx.max_seqlen_q, ctx.max_seqlen_k, ctx.dropout_p, ctx.softmax_scale, ctx.causal,
                             ctx.rng, ctx.alibi_slopes)
        return dq, dk, dv, None, None, None, None, None, None, None, None, None


def flash_attn_unpadded_qkvpacked_func(qkv, cu_seqlens, max_seqlen, dropout_p, softmax_scale=None, causal=False,
                                       return_attn_probs=False, alibi_slopes=None):
    """Self-attention on a packed varlen batch (reference :151-176).  dropout_p must be 0.0 at evaluation.

    qkv: (total, 3, nheads, headdim); cu_seqlens: (batch + 1,) int32 cumulative lengths; max_seqlen: int.
    softmax_scale defaults to 1/sqrt(headdim); causal masks key j > query i (both relative to the sequence start).
    Returns out (total, nheads, headdim); with return_attn_probs=True (testing only) also
    softmax_lse (batch, nheads, round16(max_seqlen)) fp32 and S_dmask (batch, nheads, sq_rounded, sk_rounded):
    softmax probabilities in natural layout, NEGATIVE where the position was dropped.
    alibi_slopes (extension, the reference's roadmap item README.md:57): fp32 (nheads,) tensor; adds the bias
    -slope[h] * |i - j| to the scaled scores (no gradient flows to the slopes).
    """
    return FlashAttnQKVPackedFunc.apply(qkv, cu_seqlens, max_seqlen, dropout_p, softmax_scale, causal,
                                        return_attn_probs, alibi_slopes)


def flash_attn_unpadded_kvpacked_func(q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                                      softmax_scale=None, causal=False, return_attn_probs=False, alibi_slopes=None):
    """q: (total_q, nheads, headdim), kv: (total_k, 2, nheads, headdim) (reference :179-210).  See
    flash_attn_unpadded_qkvpacked_func for the remaining arguments and the return value."""
    return FlashAttnKVPackedFunc.apply(q, kv, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                                       softmax_scale, causal, return_attn_probs, alibi_slopes)


def flash_attn_unpadded_func(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                             softmax_scale=None, causal=False, return_attn_probs=False, alibi_slopes=None):
    """q: (total_q, nheads, headdim), k and v: (total_k, nheads, headdim) (reference :213-243).  See
    flash_attn_unpadded_qkvpacked_func for the remaining arguments and the return value."""
    return FlashAttnFunc.apply(q, k, v, cu_seqlens_q, cu_seqlens_k, max_seqlen_q, max_seqlen_k, dropout_p,
                               softmax_scale, causal, return_attn_probs, alibi_slopes)


def flash_attn_func(qkv, cu_seqlens, dropout_p, max_s, softmax_scale=None, causal=False, return_attn_probs=False):
    """Old argument order, kept for backward compatibility (reference :246-252)."""
    return flash_attn_unpadded_qkvpacked_func(qkv, cu_seqlens, max_s, dropout_p, softmax_scale, causal,
                                              return_attn_probs)
