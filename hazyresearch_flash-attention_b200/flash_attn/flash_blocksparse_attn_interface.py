"""Blocksparse attention functional API: drop-in for the reference's flash_attn/flash_blocksparse_attn_interface.py.
Mask granularity is 16 query rows x 256 key columns; masked blocks are skipped by the kernels.  The reference's
native fwd_block / bwd_block sources are absent from its snapshot; semantics follow the Python call sites and the
test oracle attention_blocksparse_ref (tests/test_flash_attn.py:189-215)."""
import torch

import flash_attn_cuda


def convert_blockmask(blockmask, causal):
    """0/1 mask (nrow, ncol) -> the kernels' format (reference :8-40): int32 (ncol, nrow); column j lists the row
    indices of its nonzero blocks in increasing order, times 4, +1 when the block is the first nonzero of its row
    and +2 when it is the last nonzero of its row; the tail of each column is padded with -1."""
    assert not causal
    nrow, ncol = blockmask.shape
    on = blockmask != 0
    rows = torch.arange(nrow, device=blockmask.device)
    cols = torch.arange(ncol, device=blockmask.device)
    # first / last nonzero column of every row (rows without nonzeros get flags nowhere)
    big = ncol + 1
    first_col = torch.where(on, cols[None, :], big).min(dim=1).values
    last_col = torch.where(on, cols[None, :], -1).max(dim=1).values
    code = rows[:, None] * 4 + (cols[None, :] == first_col[:, None]).int() + 2 * (cols[None, :] == last_col[:, None]).int()
    # stable sort puts each column's nonzero rows first, in increasing row order
    order = torch.sort((~on).to(torch.uint8), dim=0, stable=True).indices
    packed = torch.gather(code, 0, order)
    packed = torch.where(torch.gather(on, 0, order), packed, -1)
    return packed.T.contiguous().to(torch.int32)


def _flash_blocksparse_attn_forward(qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal,
                                     return_softmax):
    """reference :43-51"""
    res, rng = flash_attn_cuda.fwd_block_ex(qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal,
                                            return_softmax, None)
    return res[0], res[1], (res[2] if return_softmax else None), rng


def _flash_blocksparse_attn_backward(dout, qkv, out, S_dmask, softmax_lse, cu_seqlens, blockmask, dropout_p, max_s,
                                      softmax_scale, causal, rng=None):
    """reference :54-61 (bwd_block takes softmax_scale BEFORE max_s)"""
    dqkv, dp, softmax_d = flash_attn_cuda.bwd_block(dout, qkv, out, S_dmask, softmax_lse, cu_seqlens, blockmask,
                                                    dropout_p, softmax_scale, max_s, causal, rng)
    return dqkv


class _BlocksparseBase(torch.autograd.Function):

    @staticmethod
    def _fwd(ctx, qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal, return_softmax):
        if softmax_scale is None:
            softmax_scale = qkv.shape[-1] ** (-0.5)
        # rng: the dropout stream of this call; the backward regenerates the mask from it (the reference saves and
        # restores the global CUDA RNG state instead, reference :72, :85-90)
        context, softmax_lse, S_dmask, rng = _flash_blocksparse_attn_forward(
            qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal=causal, return_softmax=return_softmax)
        ctx.save_for_backward(qkv, context, S_dmask, softmax_lse, cu_seqlens, blockmask)
        ctx.dropout_p, ctx.max_s, ctx.softmax_scale, ctx.causal, ctx.rng = dropout_p, max_s, softmax_scale, causal, rng
        return context, S_dmask, softmax_lse

    @staticmethod
    def _bwd(ctx, dout):
        qkv, context, S_dmask, softmax_lse, cu_seqlens, blockmask = ctx.saved_tensors
        return _flash_blocksparse_attn_backward(dout, qkv, context, S_dmask if S_dmask is not None else context,
                                                softmax_lse, cu_seqlens, blockmask, ctx.dropout_p, ctx.max_s,
                                                ctx.softmax_scale, ctx.causal, ctx.rng)


class FlashBlocksparseAttnFun(torch.autograd.Function):
    """returns context only (reference :64-96)"""

    @staticmethod
    def forward(ctx, qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal):
        return _BlocksparseBase._fwd(ctx, qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal, False)[0]

    @staticmethod
    def backward(ctx, dout):
        return _BlocksparseBase._bwd(ctx, dout), None, None, None, None, None, None


class FlashBlocksparseAttnFunWithS(torch.autograd.Function):
    """returns (context, S_dmask, softmax_lse) for testing (reference :101-132)"""

    @staticmethod
    def forward(ctx, qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal):
        return _BlocksparseBase._fwd(ctx, qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal, True)

    @staticmethod
    def backward(ctx, dout, _dS_dmask_ignored, _dsoftmax_sum_ignored):
        return _BlocksparseBase._bwd(ctx, dout), None, None, None, None, None, None


def flash_blocksparse_attn_func(qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale=None, causal=False,
                                 return_attn_probs=False, convert_mask=True):
    """dropout_p should be 0.0 during evaluation (reference :135-142).  blockmask: (max_s/16, max_s/256) 0/1 mask,
    or the converted format when convert_mask=False."""
    func = FlashBlocksparseAttnFunWithS if return_attn_probs else FlashBlocksparseAttnFun
    if convert_mask:
        blockmask = convert_blockmask(blockmask, causal=causal)
    return func.apply(qkv, cu_seqlens, blockmask, dropout_p, max_s, softmax_scale, causal)
