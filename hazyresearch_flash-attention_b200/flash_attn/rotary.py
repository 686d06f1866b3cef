"""Rotary position embeddings used by FlashMHA(use_rotary_emb=...).  Behavioural equivalent of the reference's
pure-PyTorch flash_attn/rotary.py (interleaved pairs (x0, x1), (x2, x3), ...; tables cached per sequence length).
The modules keep the reference's elementwise formulation (any device); FlashMHA goes through `apply_rotary_qkvpacked_`,
which rotates q and k of the packed qkv tensor (and copies v) in one CUDA kernel pass (fa_rotary_qk) instead of the reference's
elementwise passes + torch.stack copy (flash_attention.py:105-110)."""
import math
from typing import Tuple

import torch


def rotate_half(x):
    """(x0, x1, x2, x3, ...) -> (-x1, x0, -x3, x2, ...)   (reference rotary.py:21-28)"""
    pairs = x.unflatten(-1, (-1, 2))
    return torch.stack((-pairs[..., 1], pairs[..., 0]), dim=-1).flatten(-2)


def apply_rotary_pos_emb(x, cos, sin, seq_dimension: int = -2):
    """x * cos + rotate_half(x) * sin with the tables cropped to x's sequence length (reference rotary.py:31-42)."""
    n = x.shape[seq_dimension]
    cos, sin = cos[:n], sin[:n]
    if seq_dimension == -3:
        cos, sin = cos[:, None, :], sin[:, None, :]
    return x * cos + rotate_half(x) * sin


class RotaryEmbedding(torch.nn.Module):
    """RoFormer rotary embedding (reference rotary.py:45-100)."""

    def __init__(self, dim_model: int, *_, **__):
        super().__init__()
        inv_freq = 1.0 / (10000 ** (torch.arange(0, dim_model, 2).float() / dim_model))
        self.register_buffer("inv_freq", inv_freq)
        self._seq_len_cached = None
        self._cos_cached = None
        self._sin_cached = None

    def _update_cos_sin_tables(self, x, seq_dimension=-2):
        seq_len = x.shape[seq_dimension]
        stale = (seq_len != self._seq_len_cached or self._cos_cached is None
                 or self._cos_cached.device != x.device or self._cos_cached.dtype != x.dtype)
        if stale:
            self._seq_len_cached = seq_len
            t = torch.arange(seq_len, device=x.device, dtype=self.inv_freq.dtype)
            freqs = torch.outer(t, self.inv_freq.to(x.device))
            self._cos_cached = torch.cos(freqs).to(x.dtype).repeat_interleave(2, dim=-1)
            self._sin_cached = torch.sin(freqs).to(x.dtype).repeat_interleave(2, dim=-1)
        return self._cos_cached, self._sin_cached

    def forward(self, q: torch.Tensor, k: torch.Tensor, seq_dimension=-2) -> Tuple[torch.Tensor, torch.Tensor]:
        assert seq_dimension in [-2, -3]   # (b, h, s, d) or (b, s, h, d)
        cos, sin = self._update_cos_sin_tables(k, seq_dimension=seq_dimension)
        return (apply_rotary_pos_emb(q, cos, sin, seq_dimension), apply_rotary_pos_emb(k, cos, sin, seq_dimension))


class RotaryEmbedding2D(torch.nn.Module):
    """2-D rotary for square token grids: first half of the head dim rotates with the row index, second half
    with the column index (reference rotary.py:103-135)."""

    def __init__(self, dim: int):
        super().__init__()
        assert dim % 4 == 0
        self.rotary_emb1d = RotaryEmbedding(dim // 2)

    def forward(self, q: torch.Tensor, k: torch.Tensor, seq_dimension=-2):
        assert seq_dimension in [-2, -3]
        seqlen = q.shape[seq_dimension]
        side = int(math.sqrt(seqlen))
        assert seqlen == side ** 2
        if seq_dimension == -3:   # (b, s, h, d) -> (b, h, s, d)
            q, k = q.transpose(1, 2), k.transpose(1, 2)

        def grid(x):   # (b, h, s, d) -> (b, h, side, side, d)
            return x.unflatten(2, (side, side))

        q0, q1 = q.chunk(2, dim=-1)
        k0, k1 = k.chunk(2, dim=-1)
        q0e, k0e = self.rotary_emb1d(grid(q0), grid(k0), seq_dimension=-2)   # varies along grid columns (w)
        q1e, k1e = self.rotary_emb1d(grid(q1), grid(k1), seq_dimension=-3)   # varies along grid rows (h)
        q_emb = torch.cat([q0e.flatten(2, 3), q1e.flatten(2, 3)], dim=-1)
        k_emb = torch.cat([k0e.flatten(2, 3), k1e.flatten(2, 3)], dim=-1)
        if seq_dimension == -3:
            q_emb, k_emb = q_emb.transpose(1, 2), k_emb.transpose(1, 2)
        return q_emb, k_emb


# ---------------------------------------------------------------------------------------------------
# fused path used by FlashMHA
# ---------------------------------------------------------------------------------------------------
def rotary_tables(mode, seqlen, head_dim, device):
    """fp32 (seqlen, head_dim/2) cos / sin tables: entry [s, j] is the angle applied to the pair (2j, 2j+1) of a head at
    sequence position s.  '1d': s * 10000^(-2j/d) (reference rotary.py:68, 84-88).  '2d' (square grids, rotary.py:103-135):
    pairs of the first half of the head rotate with the grid column s % side, pairs of the second half with the grid row
    s // side, both with the frequencies of a (d/2)-dimensional 1-D embedding."""
    pos = torch.arange(seqlen, device=device, dtype=torch.float32)
    if mode == '1d':
        inv_freq = 1.0 / (10000 ** (torch.arange(0, head_dim, 2, device=device).float() / head_dim))
        freqs = torch.outer(pos, inv_freq)
    else:
        assert mode == '2d' and head_dim % 4 == 0
        side = int(math.sqrt(seqlen))
        assert seqlen == side ** 2
        half = head_dim // 2
        inv_freq = 1.0 / (10000 ** (torch.arange(0, half, 2, device=device).float() / half))
        idx = torch.arange(seqlen, device=device)
        freqs = torch.cat([torch.outer((idx % side).float(), inv_freq),
                           torch.outer(torch.div(idx, side, rounding_mode='floor').float(), inv_freq)], dim=-1)
    return torch.cos(freqs).contiguous(), torch.sin(freqs).contiguous()


class _RotaryQKVPacked(torch.autograd.Function):
    """qkv (b, s, 3*h*d): q and k rotated, v copied, in ONE kernel pass into a new tensor; the backward applies the
    transposed rotation to the q and k parts of the incoming gradient.
    Not in place on purpose: an in-place autograd Function (mark_dirty) on the Wqkv output makes autograd differentiate
    the later reshape views through its generic as_strided fallback, which zero-fills a full qkv-sized gradient tensor
    in every backward (measured: a 14 us FillFunctor kernel per step at b8 s1024 e1024, against 4 us for copying v here)."""

    @staticmethod
    def forward(ctx, qkv, cos, sin, nheads):
        import flash_attn_cuda
        b, s, w = qkv.shape
        shape = (b * s, 3, nheads, w // (3 * nheads))
        out = torch.empty_like(qkv)
        flash_attn_cuda.rotary_qk(qkv.view(shape), out.view(shape), cos, sin, s, n_rot=2)
        ctx.save_for_backward(cos, sin)
        ctx.nheads = nheads
        return out

    @staticmethod
    def backward(ctx, dqkv):
        import flash_attn_cuda
        cos, sin = ctx.saved_tensors
        dqkv = dqkv.contiguous()
        b, s, w = dqkv.shape
        shape = (b * s, 3, ctx.nheads, w // (3 * ctx.nheads))
        out = torch.empty_like(dqkv)
        flash_attn_cuda.rotary_qk(dqkv.view(shape), out.view(shape), cos, sin, s, n_rot=2, conjugate=True)
        return out, None, None, None


def apply_rotary_qkvpacked_(qkv, cos, sin, nheads):
    """Rotary of q and k inside the contiguous (b, s, 3*h*d) output of a QKV projection (columns ordered (three, h, d) as
    in flash_attention.py:105-106), one kernel pass; returns a new tensor (the name keeps its historical underscore)."""
    assert qkv.dim() == 3 and qkv.is_contiguous() and qkv.shape[-1] % (3 * nheads) == 0
    return _RotaryQKVPacked.apply(qkv, cos, sin, nheads)
