"""Rotary position embeddings used by FlashMHA(use_rotary_emb=...).  Behavioural equivalent of the reference's
pure-PyTorch flash_attn/rotary.py (interleaved pairs (x0, x1), (x2, x3), ...; tables cached per sequence length).
Outside the attention kernel's hot path -- plain torch elementwise code, no kernels here."""
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
