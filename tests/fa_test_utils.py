"""Shared helpers for the parity tests: input generators modelled on the reference's test-suite
(tests/test_flash_attn.py:29-112 generate_qkv) built on this repo's own unpad_input / pad_input, plus the
S_dmask decoding used by every test.  Device-agnostic."""
import os

import torch

from flash_attn.bert_padding import pad_input, unpad_input

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    return torch.load(os.path.join(GOLDEN_DIR, name), weights_only=False)


def unpack_bool(packed):
    import numpy as np
    shape = packed["shape"]
    n = 1
    for s in shape:
        n *= s
    bits = np.unpackbits(packed["bits"].numpy())[:n]
    return torch.from_numpy(bits.astype(bool)).reshape(shape)


def dtype_from_str(s):
    return {"torch.float16": torch.float16, "torch.bfloat16": torch.bfloat16}[s]


class Packed:
    """Unpadded tensors + the closures that re-pad outputs / gradients."""
    pass


def cu_seqlens_full(batch, seqlen, device):
    return torch.arange(0, (batch + 1) * seqlen, step=seqlen, dtype=torch.int32, device=device)


def pack_rows(x, mask):
    """x: (b, s, ...) -> (unpadded (nnz, ...), indices, cu_seqlens, max_seqlen); mask None = all valid."""
    b, s = x.shape[:2]
    if mask is None:
        flat = x.reshape(b * s, *x.shape[2:])
        return flat, None, cu_seqlens_full(b, s, x.device), s
    tail = x.shape[2:]
    xu, idx, cu, mx = unpad_input(x.reshape(b, s, -1), mask)
    return xu.reshape(-1, *tail), idx, cu, mx


def unpack_rows(xu, idx, batch, seqlen):
    """inverse of pack_rows: (nnz, ...) -> (b, s, ...) with zeros at padded positions."""
    tail = xu.shape[1:]
    if idx is None:
        return xu.reshape(batch, seqlen, *tail)
    return pad_input(xu.reshape(xu.shape[0], -1), idx, batch, seqlen).reshape(batch, seqlen, *tail)


def generate_qkv(q, k, v, query_padding_mask=None, key_padding_mask=None, kvpacked=False, qkvpacked=False):
    """q: (b, sq, h, d), k/v: (b, sk, h, d) padded tensors -> Packed with the unpadded views for each API form."""
    assert not (kvpacked and qkvpacked)
    P = Packed()
    b, sq = q.shape[:2]
    sk = k.shape[1]
    P.q_unpad, P.idx_q, P.cu_q, P.max_q = pack_rows(q, query_padding_mask)
    P.k_unpad, P.idx_k, P.cu_k, P.max_k = pack_rows(k, key_padding_mask)
    P.v_unpad, _, _, _ = pack_rows(v, key_padding_mask)
    P.pad_q = lambda t: unpack_rows(t, P.idx_q, b, sq)
    P.pad_k = lambda t: unpack_rows(t, P.idx_k, b, sk)
    if qkvpacked:
        P.qkv_unpad = torch.stack([P.q_unpad, P.k_unpad, P.v_unpad], dim=1).detach().requires_grad_()
    elif kvpacked:
        P.q_unpad = P.q_unpad.detach().requires_grad_()
        P.kv_unpad = torch.stack([P.k_unpad, P.v_unpad], dim=1).detach().requires_grad_()
    else:
        P.q_unpad = P.q_unpad.detach().requires_grad_()
        P.k_unpad = P.k_unpad.detach().requires_grad_()
        P.v_unpad = P.v_unpad.detach().requires_grad_()
    return P


def decode_s_dmask(S_dmask, seqlen_q, seqlen_k):
    """This build returns S_dmask in natural layout (b, h, sq_rounded, sk_rounded): |S| = softmax probability,
    sign bit set = dropped.  (The reference needs convert_flash_attn_S_to_softmax + normalize_flash_attn_S,
    tests/test_flash_attn.py:218-297, to undo its MMA-fragment layout and per-block scaling.)"""
    S = S_dmask[:, :, :seqlen_q, :seqlen_k]
    # the kernel's extents are round-ups of max_seqlen, which can be smaller than the padded test length:
    # pad with +0 (probability 0, "kept") like the reference's converter does (tests/test_flash_attn.py:253-260)
    pad_q, pad_k = seqlen_q - S.shape[2], seqlen_k - S.shape[3]
    if pad_q > 0 or pad_k > 0:
        S = torch.nn.functional.pad(S, (0, max(pad_k, 0), 0, max(pad_q, 0)))
    dropout_mask = ~torch.signbit(S.float())
    return dropout_mask, S.abs()


def expected_dropout_mask(seed, offset, batch, heads, seqlen_q, seqlen_k, p_dropout):
    """CPU evaluation of the kernels' position-keyed Philox stream (bit-exact integer work)."""
    import flash_attn_cuda
    m = torch.zeros(batch, heads, seqlen_q, (seqlen_k + 7) // 8 * 8, dtype=torch.bool)
    for b in range(batch):
        for h in range(heads):
            for i in range(seqlen_q):
                for g in range((seqlen_k + 7) // 8):
                    keep = flash_attn_cuda.dropout_keep8_host(seed, offset, b * heads + h, i, g, p_dropout)
                    m[b, h, i, g * 8:(g + 1) * 8] = torch.tensor(keep)
    return m[..., :seqlen_k]


class simt_kernels:
    """Context manager: route the calls inside through the library's CUDA-core cross-check kernels (fa_params.impl =
    FA_IMPL_SIMT).  Test-only: the product shim has no switch for this, the tests patch its module attribute."""

    def __enter__(self):
        import flash_attn_cuda
        self.mod, self.old = flash_attn_cuda, flash_attn_cuda._impl
        flash_attn_cuda._impl = flash_attn_cuda.FA_IMPL_SIMT
        return self

    def __exit__(self, *exc):
        self.mod._impl = self.old
        return False
