"""Causal K/V ring attention across the GPUs of one NVLink box (context parallelism for long sequences).

New functionality with respect to the reference (which is single-GPU; its README claims sequence lengths "up to 64K",
README.md:75, only through the in-kernel K-block loop).  The merge rule is the reference's own inter-K-block rule
(csrc/flash_attn/src/fmha_fprop_kernel_1xN.h:606-650: lse' = log(e^lse + e^lse_blk), O rescaled by exp(lse - lse'))
lifted from K blocks inside one kernel to K/V chunks living on different GPUs.

Layout ("zig-zag", balances causal work): the sequence is cut into 2N chunks of c tokens; rank r owns chunks r and
2N-1-r.  A local tensor is CHUNK-MAJOR: (2, b, c, h, d), so each chunk is a packed (b*c, h, d) batch of b sequences the
varlen kernels consume directly.  K/V blocks travel rank -> rank+1 with torch.distributed P2P (NCCL over NVLink on the
GPU box; gloo in the CPU tests) while the attention kernels work on the block already here.  At step t a rank holds the
block of rank (r - t) mod N and computes exactly two c x c block products (three at t = 0, two of them causal):

    t = 0      (qa, ka) causal, (qb, kb) causal  -> one causal call over 2b sequences;   (qb, ka) full
    src < r    (qa, ka) full, (qb, ka) full
    src > r    (qb, ka) full, (qb, kb) full

Backward: a second ring in which every K/V block travels together with its fp32 (dK, dV) accumulator and is home again
after N hops; dQ stays local.  Each block product is differentiated with the FINAL softmax_lse and D = rowsum(dO o O),
so the per-block gradients simply add up.

The per-block attention and the merge are injectable (`ops`) so that the host logic can be tested on CPU
(tests/test_ring_cpu.py uses the oracle there); the default `CudaBlockOps` calls the tcgen05 kernels through
`flash_attn_cuda` and has no fallback.
"""
import torch
import torch.distributed as dist


# ---------------------------------------------------------------------------------------------------------
# sharding helpers
# ---------------------------------------------------------------------------------------------------------
def zigzag_shard(x, world_size, rank):
    """x: (b, S, ...) full sequence -> local (2, b, c, ...) with chunks `rank` and `2*world_size-1-rank`."""
    b, S = x.shape[:2]
    assert S % (2 * world_size) == 0, "sequence length must be divisible by 2 * world_size"
    chunks = x.reshape(b, 2 * world_size, S // (2 * world_size), *x.shape[2:])
    return torch.stack([chunks[:, rank], chunks[:, 2 * world_size - 1 - rank]], dim=0).contiguous()


def zigzag_unshard(parts):
    """inverse of zigzag_shard given the list of every rank's local (2, b, c, ...) tensor -> (b, S, ...)."""
    n = len(parts)
    chunks = [None] * (2 * n)
    for r, p in enumerate(parts):
        chunks[r], chunks[2 * n - 1 - r] = p[0], p[1]
    return torch.cat(chunks, dim=1)


# ---------------------------------------------------------------------------------------------------------
# block operators (CUDA)
# ---------------------------------------------------------------------------------------------------------
class CudaBlockOps:
    """Per-block forward / backward / merge on the tcgen05 kernels (libfa_b200.so).  q: (nq*c, h, d) packed, nq sequences."""

    def __init__(self):
        import flash_attn_cuda
        self.lib = flash_attn_cuda
        self._cu = {}

    def _cu_seqlens(self, n, c, device):
        key = (n, c, device)
        if key not in self._cu:
            self._cu[key] = torch.arange(0, (n + 1) * c, c, dtype=torch.int32, device=device)
        return self._cu[key]

    def fwd(self, q, k, v, n_seq, c, scale, causal):
        cu = self._cu_seqlens(n_seq, c, q.device)
        o, lse = self.lib.fwd(q, k, v, cu, cu, c, c, 0.0, scale, False, causal, False, None)
        return o, lse                                     # (n_seq*c, h, d) 16 bit, (n_seq, h, round16(c)) fp32

    def bwd(self, dout, q, k, v, out, lse, n_seq, c, scale, causal):
        cu = self._cu_seqlens(n_seq, c, q.device)
        dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
        self.lib.bwd(dout, q, k, v, out, lse, dq, dk, dv, cu, cu, c, c, 0.0, scale, False, causal, None)
        return dq, dk, dv

    def merge(self, o_acc, lse_acc, lse_out, o_blk, lse_blk, n_seq, c, first):
        self.lib.merge_partials(o_acc, lse_acc, lse_out, o_blk, lse_blk, n_seq, c, first)


# ---------------------------------------------------------------------------------------------------------
# communication
# ---------------------------------------------------------------------------------------------------------
class _Ring:
    """send to rank+1 / receive from rank-1, as one batched P2P so that NCCL pairs them without deadlock."""

    def __init__(self, group):
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.next = dist.get_global_rank(group, (self.rank + 1) % self.world) if group is not None else (self.rank + 1) % self.world
        self.prev = dist.get_global_rank(group, (self.rank - 1) % self.world) if group is not None else (self.rank - 1) % self.world

    def exchange(self, send_tensors, recv_tensors):
        if self.world == 1:                      # a ring of one: the block comes straight back
            for s, r in zip(send_tensors, recv_tensors):
                r.copy_(s)
            return []
        ops = []
        for s, r in zip(send_tensors, recv_tensors):
            ops.append(dist.P2POp(dist.isend, s, self.next, group=self.group))
            ops.append(dist.P2POp(dist.irecv, r, self.prev, group=self.group))
        return dist.batch_isend_irecv(ops)

    @staticmethod
    def wait(reqs):
        for r in reqs:
            r.wait()


def _block_plan(step, rank, world):
    """-> list of (q_half, k_half, causal): which c x c products rank `rank` computes at ring step `step`.
    q_half / k_half: 0 = early chunk, 1 = late chunk, None = both halves as one call over 2b sequences."""
    if step == 0:
        return [(None, None, True), (1, 0, False)]
    src = (rank - step) % world
    if src < rank:
        return [(0, 0, False), (1, 0, False)]
    return [(1, 0, False), (1, 1, False)]


# ---------------------------------------------------------------------------------------------------------
# forward / backward
# ---------------------------------------------------------------------------------------------------------
def ring_forward(q, k, v, softmax_scale, group=None, ops=None):
    """q, k, v: local (2, b, c, h, d) zig-zag shards.  -> out (2, b, c, h, d) (dtype of q), lse (2b, h, c) fp32."""
    ops = ops or CudaBlockOps()
    ring = _Ring(group)
    two, b, c, h, d = q.shape
    assert two == 2 and k.shape == q.shape and v.shape == q.shape
    assert c % 16 == 0, "chunk length must be a multiple of 16"
    qf = q.reshape(2 * b * c, h, d)
    o_acc = torch.empty(2 * b * c, h, d, dtype=torch.float32, device=q.device)
    lse_a = torch.empty(2 * b, h, c, dtype=torch.float32, device=q.device)
    lse_b = torch.empty_like(lse_a)
    started = [False, False]                       # per q half: accumulator initialised?
    kv = torch.stack([k.reshape(2 * b * c, h, d), v.reshape(2 * b * c, h, d)])       # (2, 2bc, h, d): one message
    kv_next = torch.empty_like(kv)
    half = b * c
    for step in range(ring.world):
        reqs = ring.exchange([kv], [kv_next]) if step + 1 < ring.world else []
        kf, vf = kv[0], kv[1]
        for qh, kh, causal in _block_plan(step, ring.rank, ring.world):
            if qh is None:                         # both diagonal blocks in one causal call over 2b sequences
                o_blk, lse_blk = ops.fwd(qf, kf, vf, 2 * b, c, softmax_scale, True)
                ops.merge(o_acc, lse_a, lse_b, o_blk, lse_blk, 2 * b, c, True)
                lse_a, lse_b = lse_b, lse_a
                started = [True, True]
                continue
            qs, ks = slice(qh * half, (qh + 1) * half), slice(kh * half, (kh + 1) * half)
            o_blk, lse_blk = ops.fwd(qf[qs], kf[ks], vf[ks], b, c, softmax_scale, causal)
            ls = slice(qh * b, (qh + 1) * b)
            # the merge reads lse_a and writes lse_b for this half only; copy the other half across first
            other = slice((1 - qh) * b, (2 - qh) * b)
            lse_b[other].copy_(lse_a[other])
            ops.merge(o_acc[qs], lse_a[ls], lse_b[ls], o_blk, lse_blk, b, c, not started[qh])
            started[qh] = True
            lse_a, lse_b = lse_b, lse_a
        if reqs:
            ring.wait(reqs)
            kv, kv_next = kv_next, kv
    return o_acc.to(q.dtype).reshape(2, b, c, h, d), lse_a


def ring_backward(dout, q, k, v, out, lse, softmax_scale, group=None, ops=None):
    """-> dq, dk, dv, each (2, b, c, h, d) in the dtype of q.  `lse` is ring_forward's (2b, h, c) final log-sum-exp."""
    ops = ops or CudaBlockOps()
    ring = _Ring(group)
    two, b, c, h, d = q.shape
    half = b * c
    qf, of, gf = (t.reshape(2 * b * c, h, d) for t in (q, out, dout.contiguous()))
    dq_acc = torch.zeros(2 * b * c, h, d, dtype=torch.float32, device=q.device)
    kv = torch.stack([k.reshape(2 * b * c, h, d), v.reshape(2 * b * c, h, d)])
    dkv = torch.zeros(2, 2 * b * c, h, d, dtype=torch.float32, device=q.device)       # travels with its K/V block
    kv_next, dkv_next = torch.empty_like(kv), torch.empty_like(dkv)
    for step in range(ring.world):
        # the NEXT block's K/V can move while this block is being differentiated
        kv_reqs = ring.exchange([kv], [kv_next]) if step + 1 < ring.world else []
        kf, vf = kv[0], kv[1]
        for qh, kh, causal in _block_plan(step, ring.rank, ring.world):
            if qh is None:
                dq, dk, dv = ops.bwd(gf, qf, kf, vf, of, lse, 2 * b, c, softmax_scale, True)
                dq_acc += dq
                dkv[0] += dk
                dkv[1] += dv
                continue
            qs, ks = slice(qh * half, (qh + 1) * half), slice(kh * half, (kh + 1) * half)
            ls = slice(qh * b, (qh + 1) * b)
            dq, dk, dv = ops.bwd(gf[qs], qf[qs], kf[ks], vf[ks], of[qs], lse[ls].contiguous(), b, c, softmax_scale, causal)
            dq_acc[qs] += dq
            dkv[0, ks] += dk
            dkv[1, ks] += dv
        # pass the block's gradient accumulator on (after the last step it arrives back at the block's owner)
        g_reqs = ring.exchange([dkv], [dkv_next])
        ring.wait(kv_reqs)
        ring.wait(g_reqs)
        if step + 1 < ring.world:
            kv, kv_next = kv_next, kv
        dkv, dkv_next = dkv_next, dkv
    shape = (2, b, c, h, d)
    return dq_acc.to(q.dtype).reshape(shape), dkv[0].to(q.dtype).reshape(shape), dkv[1].to(q.dtype).reshape(shape)


class RingFlashAttnFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, softmax_scale, group, ops):
        softmax_scale = q.shape[-1] ** (-0.5) if softmax_scale is None else softmax_scale
        out, lse = ring_forward(q, k, v, softmax_scale, group, ops)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.softmax_scale, ctx.group, ctx.ops = softmax_scale, group, ops
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        dq, dk, dv = ring_backward(dout, q, k, v, out, lse, ctx.softmax_scale, ctx.group, ctx.ops)
        return dq, dk, dv, None, None, None


def ring_flash_attn_func(q, k, v, softmax_scale=None, group=None, ops=None):
    """Causal attention over a sequence sharded zig-zag across the ranks of `group`.
    q, k, v: local (2, b, c, h, d) shards (see zigzag_shard).  Returns the local (2, b, c, h, d) output shard."""
    return RingFlashAttnFunc.apply(q, k, v, softmax_scale, group, ops)
