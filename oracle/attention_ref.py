"""ORACLE (test infrastructure only -- never imported by the product path).

CPU/GPU-agnostic PyTorch restatement of the reference's own test oracle for the
attention hot path.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this module.

Parity status: PINNED for the forward math against outputs of the reference's own
``attention_ref`` / ``attention_blocksparse_ref`` / ``convert_blockmask`` /
``get_dropout_fraction`` (imported from /root/reference by
``tests/golden/make_golden.py``; fixtures committed under ``tests/golden/``).
The reference ships no golden vectors of its own and its
backward / blocksparse CUDA sources are absent, so gradients are pinned only to
autograd of this same math ("parity unpinned" for bwd/blocksparse beyond that).

Each function cites the reference lines it restates (paths relative to
/root/reference).
"""
import math

import torch


def _expand_key_mask(mask):
    # (b, s) -> (b, 1, 1, s)
    return mask[:, None, None, :]


def _expand_query_mask(mask):
    # (b, t) -> (b, 1, t, 1)
    return mask[:, None, :, None]


def causal_mask_topleft(seqlen_q, seqlen_k, device):
    """True where key j is NOT visible to query i (j > i), top-left aligned.
    tests/test_flash_attn.py:147 (torch.triu(ones(sq, sk), 1))."""
    i = torch.arange(seqlen_q, device=device)[:, None]
    j = torch.arange(seqlen_k, device=device)[None, :]
    return j > i


def alibi_bias(alibi_slopes, seqlen_q, seqlen_k, device):
    """EXTENSION (not in the reference's test oracle; its README.md:57 roadmap names "attention bias (e.g. ALiBi)"):
    (h, sq, sk) additive bias -slope[h] * |i - j| on the scaled scores, i / j = query / key index inside the sequence
    (top-left aligned like the causal mask above)."""
    i = torch.arange(seqlen_q, device=device, dtype=torch.float32)[:, None]
    j = torch.arange(seqlen_k, device=device, dtype=torch.float32)[None, :]
    return -alibi_slopes.float().to(device)[:, None, None] * (i - j).abs()[None]


def attention_ref(q, k, v, query_padding_mask=None, key_padding_mask=None, dropout_p=0.0,
                  dropout_mask=None, causal=False, upcast=True, reorder_ops=False, alibi_slopes=None):
    """Standard attention.  Restates tests/test_flash_attn.py:115-159.

    q: (b, sq, h, d); k, v: (b, sk, h, d); masks are bool (True = valid token);
    dropout_mask: (b, h, sq, sk) bool, True = keep.  ``None`` means keep-all (the
    reference raises UnboundLocalError for None -- SURVEY.md appendix B).
    Returns (output (b, sq, h, d), attention (b, h, sq, sk) = softmax BEFORE dropout).
    upcast=True -> fp32 math ("reference"); upcast=False, reorder_ops=True -> the
    low-precision "PyTorch baseline" whose error defines the 2x tolerance.
    """
    out_dtype = q.dtype
    if upcast:
        q, k, v = q.float(), k.float(), v.float()
    sq, sk, d = q.shape[1], k.shape[1], q.shape[-1]
    rsqrt_d = math.sqrt(d)
    if reorder_ops:
        scores = torch.einsum('bthd,bshd->bhts', q, k / rsqrt_d)        # :143
    else:
        scores = torch.einsum('bthd,bshd->bhts', q / rsqrt_d, k)        # :141
    if alibi_slopes is not None:                                         # extension, see alibi_bias
        scores = scores + alibi_bias(alibi_slopes, sq, sk, q.device)[None].to(scores.dtype)
    if key_padding_mask is not None:
        scores = scores.masked_fill(~_expand_key_mask(key_padding_mask), float('-inf'))   # :145
    if causal:
        scores = scores.masked_fill(causal_mask_topleft(sq, sk, q.device), float('-inf'))  # :147-148
    attention = torch.softmax(scores, dim=-1)                            # :149
    keep_scale = 1.0 / (1.0 - dropout_p)                                 # :150
    attention_drop = attention if dropout_mask is None else attention.masked_fill(~dropout_mask, 0.0)
    # the reference folds the dropout rescale into V, not P (:155)
    output = torch.einsum('bhts,bshd->bthd', attention_drop, v * keep_scale)
    if query_padding_mask is not None:
        output = output.masked_fill(~query_padding_mask[:, :, None, None], 0.0)            # :157
        attention = attention.masked_fill(~_expand_query_mask(query_padding_mask), 0.0)    # :158
    return output.to(out_dtype), attention.to(out_dtype)


def attention_kvpacked_ref(q, kv, query_padding_mask=None, key_padding_mask=None, dropout_p=0.0,
                           dropout_mask=None, causal=False, upcast=True, reorder_ops=False):
    """tests/test_flash_attn.py:162-166. kv: (b, sk, 2, h, d)."""
    return attention_ref(q, kv[:, :, 0], kv[:, :, 1], query_padding_mask, key_padding_mask,
                         dropout_p, dropout_mask, causal=causal, upcast=upcast,
                         reorder_ops=reorder_ops)


def attention_qkvpacked_ref(qkv, key_padding_mask=None, dropout_p=0.0, dropout_mask=None,
                            causal=False, upcast=True, reorder_ops=False):
    """tests/test_flash_attn.py:169-173. qkv: (b, s, 3, h, d)."""
    return attention_ref(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], key_padding_mask,
                         key_padding_mask, dropout_p, dropout_mask, causal=causal, upcast=upcast,
                         reorder_ops=reorder_ops)


def softmax_lse_ref(q, k, query_padding_mask=None, key_padding_mask=None, causal=False,
                    softmax_scale=None):
    """Natural-log logsumexp of the scaled, masked scores in fp32: (b, h, sq).
    Semantics of the kernel's softmax_lse output (csrc/flash_attn/src/fmha_fprop_kernel_1xN.h:593,617;
    doc flash_attn/flash_attn_interface.py:168-170).  Fully masked rows give -inf."""
    q, k = q.float(), k.float()
    sq, sk, d = q.shape[1], k.shape[1], q.shape[-1]
    scale = (1.0 / math.sqrt(d)) if softmax_scale is None else softmax_scale
    scores = torch.einsum('bthd,bshd->bhts', q * scale, k)
    if key_padding_mask is not None:
        scores = scores.masked_fill(~_expand_key_mask(key_padding_mask), float('-inf'))
    if causal:
        scores = scores.masked_fill(causal_mask_topleft(sq, sk, q.device), float('-inf'))
    return torch.logsumexp(scores, dim=-1)


def expand_blockmask(blockmask, seqlen):
    """(S/16, S/256) 0/1 mask -> (seqlen, seqlen) bool. tests/test_flash_attn.py:206-207."""
    m = blockmask.bool().repeat_interleave(16, dim=0).repeat_interleave(256, dim=1)
    return m[:seqlen, :seqlen]


def attention_blocksparse_ref(qkv, blockmask, attn_mask, dropout_p, dropout_mask, upcast=True):
    """Blocksparse attention oracle.  Restates tests/test_flash_attn.py:189-215 (upcast=False, an extension, runs the
    same math in the input precision: the "PyTorch baseline" of the 2x rule, which the reference only has for the dense
    oracle).
    qkv: (b, s, 3, h, d); blockmask (s/16, s/256); attn_mask (b, s) bool.
    Rows whose every block is masked produce NaN in the reference math (softmax of all -inf);
    here they are defined as zero output / zero attention, which is what the kernel emits
    (fully-masked-row rule)."""
    q, k, v = (qkv.float() if upcast else qkv).unbind(dim=2)
    d, seqlen = qkv.shape[-1], qkv.shape[1]
    scores = torch.einsum('bthd,bshd->bhts', q / math.sqrt(d), k)
    scores = scores.masked_fill(~_expand_key_mask(attn_mask), float('-inf'))
    dense = expand_blockmask(blockmask, seqlen).to(qkv.device)
    scores = scores.masked_fill(~dense[None, None], float('-inf'))
    attention = torch.softmax(scores, dim=-1)
    attention = torch.nan_to_num(attention, nan=0.0)
    attention = attention.masked_fill(~_expand_query_mask(attn_mask), 0.0)
    attention = attention.masked_fill(~dense[None, None], 0.0)
    drop = attention if dropout_mask is None else attention.masked_fill(~dropout_mask, 0.0)
    drop = drop / (1.0 - dropout_p)
    output = torch.einsum('bhts,bshd->bthd', drop, v)
    output = output.masked_fill(~attn_mask[:, :, None, None], 0.0)
    return output.to(qkv.dtype), attention.to(qkv.dtype)


def generate_random_padding_mask(max_seqlen, batch_size, device, mode='random'):
    """tests/test_flash_attn.py:17-26. 'random': lengths in [max(1, S-20), S) (upper bound exclusive)."""
    assert mode in ('full', 'random', 'third')
    if mode == 'full':
        lengths = torch.full((batch_size, 1), max_seqlen, device=device, dtype=torch.int32)
    elif mode == 'random':
        lengths = torch.randint(max(1, max_seqlen - 20), max_seqlen, (batch_size, 1), device=device)
    else:
        lengths = torch.randint(max_seqlen // 3, max_seqlen, (batch_size, 1), device=device)
    return torch.arange(max_seqlen, device=device)[None, :].expand(batch_size, -1) < lengths


def generate_sparsity_mask(seqlen, sparsity=0.3, device='cpu'):
    """tests/test_flash_attn.py:176-186 (device made a parameter)."""
    return torch.rand(seqlen // 16, seqlen // 256, device=device) < sparsity


def butterfly_blockmask(seqlen, device='cpu'):
    """Butterfly layout at the kernel's (16 x 256) granularity.  The reference takes its
    layout from an external hydra sparsity_config (flash_attn/flash_blocksparse_attention.py:27,33)
    that is not in the repo; this build defines: 256-block (I, J) is on iff I == J or I xor J is a
    power of two; every 16-row block inside 256-row block I shares that row."""
    n = (seqlen + 255) // 256
    I = torch.arange(n, device=device)[:, None]
    J = torch.arange(n, device=device)[None, :]
    x = I ^ J
    on = (x == 0) | ((x & (x - 1)) == 0)
    return on.repeat_interleave(16, dim=0)[: (seqlen + 15) // 16]


def get_dropout_fraction(dropout_mask, query_padding_mask=None, key_padding_mask=None, causal=False):
    """Fraction of VALID positions that were dropped.  tests/test_flash_attn.py:300-329."""
    b, h, sq, sk = dropout_mask.shape
    dropped = ~dropout_mask
    if query_padding_mask is not None:
        dropped = dropped & _expand_query_mask(query_padding_mask)
    if key_padding_mask is not None:
        dropped = dropped & _expand_key_mask(key_padding_mask)
    if causal:
        dropped = dropped & ~causal_mask_topleft(sq, sk, dropout_mask.device)
    dev = dropout_mask.device
    ql = query_padding_mask.sum(-1) if query_padding_mask is not None else torch.full((b,), sq, device=dev)
    kl = key_padding_mask.sum(-1) if key_padding_mask is not None else torch.full((b,), sk, device=dev)
    if not causal:
        per_batch = ql * kl
    else:
        per_batch = torch.where(ql <= kl, ql * (ql + 1) / 2, ql * kl - (kl * (kl - 1) / 2))   # :321-328
    return dropped.sum() / (per_batch.sum() * h)


def convert_blockmask_ref(blockmask):
    """Plain-loop restatement of flash_attn/flash_blocksparse_attn_interface.py:8-40 for checking the
    product's vectorised convert_blockmask.  Returns (ncol, nrow) int32: column j lists the row indices
    of its nonzero blocks x4, +1 if that block is the first nonzero of its row, +2 if the last,
    padded with -1."""
    bm = blockmask.to(torch.int64).cpu()
    nrow, ncol = bm.shape
    out = torch.full((ncol, nrow), -1, dtype=torch.int32)
    first = [None] * nrow
    last = [None] * nrow
    for r in range(nrow):
        nz = [c for c in range(ncol) if bm[r, c] != 0]
        if nz:
            first[r], last[r] = nz[0], nz[-1]
        else:
            # reference: sort-based argmax on an all-zero row picks column index from a stable sort:
            # descending stable sort -> first index 0; ascending stable sort -> last index ncol-1.
            first[r], last[r] = 0, ncol - 1
    for c in range(ncol):
        slot = 0
        for r in range(nrow):
            if bm[r, c] != 0:
                val = 4 * r
                if first[r] == c:
                    val += 1
                if last[r] == c:
                    val += 2
                out[c, slot] = val
                slot += 1
    return out


def max_abs_err(a, b):
    return (a.float() - b.float()).abs().max().item()


def within_2x(x, x_ref, x_pt, floor=0.0):
    """The reference's acceptance rule (tests/test_flash_attn.py:407,409): error of the kernel against the
    fp32 reference is at most twice the error of the same-precision PyTorch implementation.
    ``floor`` guards the degenerate case where the PyTorch error is exactly zero."""
    return max_abs_err(x, x_ref) <= 2.0 * max_abs_err(x_pt, x_ref) + floor
