"""Varlen pack / unpack: drop-in for the reference's flash_attn/bert_padding.py (lines cited per function).

On CUDA tensors the row gather / scatter runs in this project's own kernels (fa_gather_rows / fa_scatter_rows of
libfa_b200.so) instead of torch.gather / scatter_ / index_put; on CPU tensors -- the reference's functions are
device agnostic -- plain torch indexing is used.
"""
import torch
import torch.nn.functional as F


def _flatten_rows(x):
    return x.reshape(x.shape[0], -1)


def _use_kernels(x2d, indices):
    return x2d.is_cuda and indices.is_cuda and (x2d.shape[1] * x2d.element_size()) % 4 == 0


def _gather(x, indices):
    """x[indices] along axis 0."""
    other = x.shape[1:]
    x2d = _flatten_rows(x)
    if _use_kernels(x2d, indices):
        import flash_attn_cuda
        out = flash_attn_cuda.gather_rows(x2d.contiguous(), indices.contiguous())
    else:
        out = x2d[indices]
    return out.reshape(-1, *other)


def _scatter_into_zeros(values, indices, first_axis_dim):
    """zeros(first_axis_dim, ...)[indices] = values."""
    other = values.shape[1:]
    v2d = _flatten_rows(values)
    if _use_kernels(v2d, indices):
        import flash_attn_cuda
        out = flash_attn_cuda.scatter_rows(v2d.contiguous(), indices.contiguous(), first_axis_dim)
    else:
        out = torch.zeros(first_axis_dim, v2d.shape[1], device=values.device, dtype=values.dtype)
        out[indices] = v2d
    return out.reshape(first_axis_dim, *other)


class IndexFirstAxis(torch.autograd.Function):
    """forward: input[indices]; backward: scatter into zeros (reference :11-38)."""

    @staticmethod
    def forward(ctx, input, indices):
        ctx.save_for_backward(indices)
        assert input.ndim >= 2
        ctx.first_axis_dim = input.shape[0]
        return _gather(input, indices)

    @staticmethod
    def backward(ctx, grad_output):
        indices, = ctx.saved_tensors
        assert grad_output.ndim >= 2
        return _scatter_into_zeros(grad_output, indices, ctx.first_axis_dim), None


index_first_axis = IndexFirstAxis.apply


class IndexPutFirstAxis(torch.autograd.Function):
    """forward: zeros(first_axis_dim, ...)[indices] = values; backward: grad[indices] (reference :41-64)."""

    @staticmethod
    def forward(ctx, values, indices, first_axis_dim):
        ctx.save_for_backward(indices)
        assert indices.ndim == 1
        assert values.ndim >= 2
        return _scatter_into_zeros(values, indices, first_axis_dim)

    @staticmethod
    def backward(ctx, grad_output):
        indices, = ctx.saved_tensors
        return _gather(grad_output, indices), None, None


index_put_first_axis = IndexPutFirstAxis.apply


class IndexFirstAxisResidual(torch.autograd.Function):
    """forward: (input[indices], input.detach()); backward: grad_residual with grad_output added at indices
    (reference :67-96)."""

    @staticmethod
    def forward(ctx, input, indices):
        ctx.save_for_backward(indices)
        assert input.ndim >= 2
        ctx.first_axis_dim = input.shape[0]
        return _gather(input, indices), input.detach()

    @staticmethod
    def backward(ctx, grad_output, grad_residual):
        indices, = ctx.saved_tensors
        assert grad_output.ndim >= 2
        assert grad_residual.shape[1:] == grad_output.shape[1:]
        g2d, r2d = _flatten_rows(grad_output), _flatten_rows(grad_residual)
        if (_use_kernels(g2d, indices) and g2d.dtype == r2d.dtype and g2d.dtype in (torch.float16, torch.bfloat16, torch.float32)
                and (g2d.shape[1] * g2d.element_size()) % 16 == 0):
            import flash_attn_cuda
            grad_input = flash_attn_cuda.scatter_add_rows(r2d.clone(), g2d.contiguous(), indices.contiguous())
        else:
            grad_input = r2d.index_add(0, indices, g2d)
        return grad_input.reshape(ctx.first_axis_dim, *grad_output.shape[1:]), None


index_first_axis_residual = IndexFirstAxisResidual.apply


def unpad_input(hidden_states, attention_mask):
    """(batch, seqlen, ...) + (batch, seqlen) mask (1 = valid) -> packed tokens (reference :99-119).

    Returns (hidden_states (total_nnz, ...), indices (total_nnz,) int64, cu_seqlens (batch + 1,) int32,
    max_seqlen_in_batch: python int)."""
    seqlens_in_batch = attention_mask.sum(dim=-1, dtype=torch.int32)
    indices = torch.nonzero(attention_mask.flatten(), as_tuple=False).flatten()
    max_seqlen_in_batch = seqlens_in_batch.max().item()
    cu_seqlens = F.pad(torch.cumsum(seqlens_in_batch, dim=0, dtype=torch.int32), (1, 0))
    flat = hidden_states.reshape(hidden_states.shape[0] * hidden_states.shape[1], *hidden_states.shape[2:])
    return index_first_axis(flat, indices), indices, cu_seqlens, max_seqlen_in_batch


def unpad_input_nosync(hidden_states, attention_mask):
    """unpad_input without the host round trip (the reference synchronises twice per call: nonzero() and .item(),
    reference :109-112).  CUDA only.  Same 4-tuple, with upper bounds where the reference has exact values:

      hidden_states  (batch * seqlen, ...): the valid tokens first, in order; the remaining rows are zero
      indices        (batch * seqlen,) int64: positions of the valid tokens, then -1 (pad_input and the gradients skip them)
      cu_seqlens     (batch + 1,) int32, exact, on the device
      max_seqlen     `seqlen` (python int): an upper bound, which is all the attention kernels need

    Everything downstream (flash_attn_unpadded_*_func, pad_input, autograd) gives the same result for the valid tokens."""
    import flash_attn_cuda
    assert hidden_states.is_cuda and attention_mask.is_cuda, "unpad_input_nosync is the CUDA path; use unpad_input on CPU"
    batch, seqlen = attention_mask.shape
    indices, cu_seqlens, _ = flash_attn_cuda.varlen_index(attention_mask)
    flat = hidden_states.reshape(batch * seqlen, *hidden_states.shape[2:])
    return index_first_axis(flat, indices), indices, cu_seqlens, seqlen


def pad_input(hidden_states, indices, batch, seqlen):
    """packed (total_nnz, ...) -> (batch, seqlen, ...) with zeros at padded positions (reference :122-134)."""
    output = index_put_first_axis(hidden_states, indices, batch * seqlen)
    return output.reshape(batch, seqlen, *hidden_states.shape[1:])
