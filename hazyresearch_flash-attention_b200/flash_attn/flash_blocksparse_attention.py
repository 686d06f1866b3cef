"""Blocksparse nn.Modules: drop-in for the reference's flash_attn/flash_blocksparse_attention.py.

The reference builds its layout with `hydra.utils.instantiate(sparsity_config).make_layout(max_seq_length)`
(flash_blocksparse_attention.py:27,33); hydra and that config class are external to the reference repo.  Here
`sparsity_config` may be (a) any object with a `make_layout(seqlen)` method returning a (seqlen/16, seqlen/256)
0/1 tensor, (b) a ready-made layout tensor, or (c) a hydra-style dict -- instantiated through hydra when it is
installed.  `ButterflySparsityConfig` is the layout this build benchmarks.
"""
import math

import torch
import torch.nn as nn

from flash_attn.bert_padding import pad_input, unpad_input
from flash_attn.flash_blocksparse_attn_interface import convert_blockmask, flash_blocksparse_attn_func


class ButterflySparsityConfig:
    """256 x 256 key/query blocks (I, J) are on iff I == J or I xor J is a power of two; emitted at the kernels'
    (16 x 256) granularity."""

    def make_layout(self, seqlen):
        n = (seqlen + 255) // 256
        i = torch.arange(n)[:, None]
        j = torch.arange(n)[None, :]
        x = i ^ j
        on = (x == 0) | ((x & (x - 1)) == 0)
        return on.repeat_interleave(16, dim=0)[: (seqlen + 15) // 16].to(torch.int32)


def _resolve_layout(sparsity_config, max_seq_length):
    if torch.is_tensor(sparsity_config):
        return sparsity_config
    if hasattr(sparsity_config, "make_layout"):
        return sparsity_config.make_layout(max_seq_length)
    try:
        import hydra
    except ImportError as e:
        raise ImportError("sparsity_config is a hydra config but hydra is not installed; pass an object with "
                          "make_layout(seqlen) or a layout tensor instead") from e
    return hydra.utils.instantiate(sparsity_config).make_layout(max_seq_length)


class FlashBlocksparseAttention(nn.Module):
    """Blocksparse scaled dot product attention (reference :14-104)."""

    def __init__(self, sparsity_config, softmax_temp=None, attention_dropout=0.0, max_seq_length=2048,
                 device=None, dtype=None):
        super().__init__()
        self.softmax_temp = softmax_temp
        self.dropout_p = attention_dropout
        max_seq_length = ((max_seq_length + 256 - 1) // 256) * 256
        self.max_seq_length = max_seq_length
        layout = _resolve_layout(sparsity_config, max_seq_length)
        self.register_buffer("layout", layout)
        self.register_buffer("blockmask_converted", convert_blockmask(layout, causal=False))

    def forward(self, qkv, attn_mask=None, key_padding_mask=None, causal=False, cu_seqlens=None, max_s=None,
                need_weights=False, convert_mask=True):
        """qkv: (B, S, 3, H, D) (or unpadded (nnz, 3, H, D) with cu_seqlens); key_padding_mask: bool (B, S)."""
        assert not need_weights
        assert attn_mask is None
        assert qkv.dtype in (torch.float16, torch.bfloat16)
        assert qkv.is_cuda
        p = self.dropout_p if self.training else 0.0

        def mask_for(seqlen):
            seqlen_rounded = ((seqlen + 256 - 1) // 256) * 256
            assert seqlen_rounded // 16 <= self.layout.shape[0] and seqlen_rounded // 256 <= self.layout.shape[1]
            return self.layout[:seqlen_rounded // 16, :seqlen_rounded // 256]

        if cu_seqlens is not None:
            assert max_s is not None
            if convert_mask:
                out = flash_blocksparse_attn_func(qkv, cu_seqlens, mask_for(max_s), p, max_s,
                                                  softmax_scale=self.softmax_temp, causal=causal)
            else:   # pre-converted full-length mask (reference :97-101)
                out = flash_blocksparse_attn_func(qkv, cu_seqlens, self.blockmask_converted, p, max_s,
                                                  softmax_scale=self.softmax_temp, causal=causal, convert_mask=False)
            return out, None
        batch, seqlen = qkv.shape[0], qkv.shape[1]
        blockmask = mask_for(seqlen)
        if key_padding_mask is None:
            flat = qkv.reshape(batch * seqlen, *qkv.shape[2:])
            cu = torch.arange(0, (batch + 1) * seqlen, step=seqlen, dtype=torch.int32, device=qkv.device)
            out = flash_blocksparse_attn_func(flat, cu, blockmask, p, seqlen, softmax_scale=self.softmax_temp,
                                              causal=causal)
            return out.reshape(batch, seqlen, *out.shape[1:]), None
        key_padding_mask_bool = key_padding_mask.bool_matrix if hasattr(key_padding_mask, "bool_matrix") \
            else key_padding_mask
        nheads, headdim = qkv.shape[-2], qkv.shape[-1]
        x_unpad, indices, cu, max_s = unpad_input(qkv.reshape(batch, seqlen, 3 * nheads * headdim),
                                                  key_padding_mask_bool)
        out_unpad = flash_blocksparse_attn_func(x_unpad.reshape(-1, 3, nheads, headdim), cu, blockmask, p, max_s,
                                                softmax_scale=self.softmax_temp, causal=causal)
        out = pad_input(out_unpad.reshape(-1, nheads * headdim), indices, batch, seqlen)
        return out.reshape(batch, seqlen, nheads, headdim), None


class FlashBlocksparseMHA(nn.Module):
    """Wqkv -> FlashBlocksparseAttention -> out_proj (reference :107-136)."""

    def __init__(self, embed_dim, num_heads, sparsity_config, bias=True, batch_first=True, attention_dropout=0.0,
                 causal=False, max_seq_length=2048, device=None, dtype=None, **kwargs) -> None:
        assert batch_first
        factory_kwargs = {'device': device, 'dtype': dtype}
        super().__init__()
        self.embed_dim = embed_dim
        self.causal = causal
        self.num_heads = num_heads
        assert self.embed_dim % num_heads == 0, "self.kdim must be divisible by num_heads"
        self.head_dim = self.embed_dim // num_heads
        assert self.head_dim in [16, 32, 64], "Only support head_dim == 16, 32, or 64"
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim, bias=bias, **factory_kwargs)
        self.inner_attn = FlashBlocksparseAttention(sparsity_config, attention_dropout=attention_dropout,
                                                    max_seq_length=max_seq_length, **factory_kwargs)
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=bias, **factory_kwargs)

    def forward(self, x, x_ignored_, x_ignored_1_, attn_mask=None, key_padding_mask=None, need_weights=False):
        b, s, _ = x.shape
        qkv = self.Wqkv(x).reshape(b, s, 3, self.num_heads, self.head_dim)
        context, attn_weights = self.inner_attn(qkv, key_padding_mask=key_padding_mask, need_weights=need_weights,
                                                causal=self.causal)
        return self.out_proj(context.reshape(b, s, self.embed_dim)), attn_weights
