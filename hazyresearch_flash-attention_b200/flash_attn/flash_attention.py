"""nn.Modules over the functional API: drop-in for the reference's flash_attn/flash_attention.py.

FlashAttention accepts the reference's three input modes (padded dense, padded + key_padding_mask, already
unpadded with cu_seqlens); FlashMHA is Wqkv -> optional rotary -> FlashAttention -> out_proj.
Deviation: bf16 inputs are accepted as well as fp16 (the reference asserts fp16 only, flash_attention.py:37),
because the kernels support both and the headline d=128 configuration is bf16.
"""
import torch
import torch.nn as nn

from flash_attn.bert_padding import pad_input, unpad_input_nosync
from flash_attn.flash_attn_interface import flash_attn_unpadded_qkvpacked_func
from flash_attn.rotary import RotaryEmbedding, RotaryEmbedding2D, apply_rotary_qkvpacked_, rotary_tables


class FlashAttention(nn.Module):
    """Scaled dot product attention with softmax (reference :12-72).

    softmax_scale: temperature (default 1/sqrt(d) at runtime); attention_dropout: dropout rate, applied only in
    training mode.
    """

    def __init__(self, softmax_scale=None, attention_dropout=0.0, device=None, dtype=None):
        super().__init__()
        self.softmax_scale = softmax_scale
        self.dropout_p = attention_dropout
        self._cu_cache = (None, None)     # ((batch, seqlen, device), cu_seqlens of a dense batch)

    def forward(self, qkv, key_padding_mask=None, causal=False, cu_seqlens=None, max_s=None, need_weights=False):
        """qkv: (B, S, 3, H, D) when cu_seqlens is None, else already unpadded (nnz, 3, H, D);
        key_padding_mask: bool (B, S), True = valid token.  Returns (output, None)."""
        assert not need_weights
        assert qkv.dtype in (torch.float16, torch.bfloat16)
        assert qkv.is_cuda
        p = self.dropout_p if self.training else 0.0
        if cu_seqlens is not None:
            assert max_s is not None
            out = flash_attn_unpadded_qkvpacked_func(qkv, cu_seqlens, max_s, p, softmax_scale=self.softmax_scale,
                                                     causal=causal)
            return out, None
        batch, seqlen = qkv.shape[0], qkv.shape[1]
        if key_padding_mask is None:
            flat = qkv.reshape(batch * seqlen, *qkv.shape[2:])
            key = (batch, seqlen, qkv.device)
            if self._cu_cache[0] != key:      # one arange launch per shape, not per call
                self._cu_cache = (key, torch.arange(0, (batch + 1) * seqlen, step=seqlen, dtype=torch.int32, device=qkv.device))
            cu = self._cu_cache[1]
            out = flash_attn_unpadded_qkvpacked_func(flat, cu, seqlen, p, softmax_scale=self.softmax_scale,
                                                     causal=causal)
            return out.reshape(batch, seqlen, *out.shape[1:]), None
        nheads, headdim = qkv.shape[-2], qkv.shape[-1]
        # indices / cu_seqlens are built on the device: no host synchronisation anywhere in this module (the reference
        # synchronises twice here, bert_padding.py:109-112); max_s is the padded length, an upper bound
        x_unpad, indices, cu, max_s = unpad_input_nosync(qkv.reshape(batch, seqlen, 3 * nheads * headdim), key_padding_mask)
        out_unpad = flash_attn_unpadded_qkvpacked_func(x_unpad.reshape(-1, 3, nheads, headdim), cu, max_s, p,
                                                       softmax_scale=self.softmax_scale, causal=causal)
        out = pad_input(out_unpad.reshape(-1, nheads * headdim), indices, batch, seqlen)
        return out.reshape(batch, seqlen, nheads, headdim), None


class FlashMHA(nn.Module):
    """Multi-head self attention block (reference :75-115)."""

    def __init__(self, embed_dim, num_heads, bias=True, batch_first=True, attention_dropout=0.0, causal=False,
                 use_rotary_emb=None, device=None, dtype=None, **kwargs) -> None:
        assert batch_first
        factory_kwargs = {'device': device, 'dtype': dtype}
        super().__init__()
        self.embed_dim = embed_dim
        self.causal = causal
        self.num_heads = num_heads
        assert self.embed_dim % num_heads == 0, "self.kdim must be divisible by num_heads"
        self.head_dim = self.embed_dim // num_heads
        assert self.head_dim in [16, 32, 64, 128], "Only support head_dim == 16, 32, 64, or 128"
        assert use_rotary_emb in [None, '1d', '2d']
        self.use_rotary_emb = use_rotary_emb
        if self.use_rotary_emb == '1d':
            self.rotary_emb = RotaryEmbedding(self.head_dim)
        elif self.use_rotary_emb == '2d':
            self.rotary_emb = RotaryEmbedding2D(self.head_dim)
        self._rotary_cache = (None, None)   # ((seqlen, device), (cos, sin)) for the fused rotary kernel
        self.Wqkv = nn.Linear(embed_dim, 3 * embed_dim, bias=bias, **factory_kwargs)
        self.inner_attn = FlashAttention(attention_dropout=attention_dropout, **factory_kwargs)
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=bias, **factory_kwargs)

    def forward(self, x, key_padding_mask=None, need_weights=False):
        """x: (batch, seqlen, embed_dim); key_padding_mask: bool (batch, seqlen).  Returns (out, None)."""
        b, s, _ = x.shape
        qkv = self.Wqkv(x)
        if self.use_rotary_emb:
            # one kernel pass (the reference: elementwise torch ops + torch.stack, :105-110)
            key = (s, qkv.device)
            if self._rotary_cache[0] != key:
                self._rotary_cache = (key, rotary_tables(self.use_rotary_emb, s, self.head_dim, qkv.device))
            cos, sin = self._rotary_cache[1]
            qkv = apply_rotary_qkvpacked_(qkv.contiguous(), cos, sin, self.num_heads)
        qkv = qkv.reshape(b, s, 3, self.num_heads, self.head_dim)
        context, attn_weights = self.inner_attn(qkv, key_padding_mask=key_padding_mask, need_weights=need_weights,
                                                causal=self.causal)
        return self.out_proj(context.reshape(b, s, self.embed_dim)), attn_weights
