"""Causal K/V ring attention across the GPUs of one NVLink box (context parallelism for long sequences).

New functionality with respect to the reference (which is single-GPU; its README claims sequence lengths "up to 64K",
README.md:75, only through the in-kernel K-block loop).  The merge rule is the reference's own inter-K-block rule
(csrc/flash_attn/src/fmha_fprop_kernel_1xN.h:606-650: lse' = log(e^lse + e^lse_blk), O rescaled by exp(lse - lse'))
lifted from K blocks inside one kernel to K/V chunks living on different GPUs.

Layout ("zig-zag", balances causal work): the sequence is cut into 2N chunks of c tokens; rank r owns chunks r and
2N-1-r.  A local tensor is (b, 2, c, h, d): the two chunks of a batch element are adjacent, so a batch element is one
packed "sequence" of 2c rows whose halves a kernel launch can address through `seqused_q` / `seqused_k` (fa_params).
K/V blocks travel rank -> rank+1 with torch.distributed P2P (NCCL over NVLink on the GPU box; gloo in the CPU tests)
while the attention kernel works on the block already here.  At step t a rank holds the block of rank src = (r - t)
mod N and computes exactly two c x c block products' worth of work in ONE kernel launch:

    t = 0      [qa; qb] x [ka; kb] causal     (a precedes b in the global order: plain causal attention over 2c rows)
    src < r    [qa; qb] x ka       full       (both local query chunks come after the early chunk of an earlier rank)
    src > r    qb       x [ka; kb] full       (only the late query chunk sees a later rank's chunks -- both of them)

Backward: a second ring in which every K/V block is followed, one step behind, by its fp32 (dK, dV) accumulator, which
is home again after N hops; dQ stays local.  Each block product is differentiated with the FINAL softmax_lse and
D = rowsum(dO o O), so the per-block gradients simply add up.  The accumulator exchange of step t overlaps the
attention kernel of step t+1 (it is only needed when that kernel's dK / dV are added), the K/V exchange overlaps the
kernel of the same step.

The per-block attention, merge and accumulate operators are injectable (`ops`) so that the host logic can be tested
on CPU (tests/test_ring_cpu.py uses an fp32 PyTorch restatement there); the default `CudaBlockOps` calls the tcgen05
kernels through `flash_attn_cuda` and has no fallback.
"""
import torch
import torch.distributed as dist


# ---------------------------------------------------------------------------------------------------------
# sharding helpers
# ---------------------------------------------------------------------------------------------------------
def zigzag_shard(x, world_size, rank):
    """x: (b, S, ...) full sequence -> local (b, 2, c, ...) with chunks `rank` and `2*world_size-1-rank`."""
    b, S = x.shape[:2]
    assert S % (2 * world_size) == 0, "sequence length must be divisible by 2 * world_size"
    chunks = x.reshape(b, 2 * world_size, S // (2 * world_size), *x.shape[2:])
    return torch.stack([chunks[:, rank], chunks[:, 2 * world_size - 1 - rank]], dim=1).contiguous()


def zigzag_unshard(parts):
    """inverse of zigzag_shard given the list of every rank's local (b, 2, c, ...) tensor -> (b, S, ...)."""
    n = len(parts)
    chunks = [None] * (2 * n)
    for r, p in enumerate(parts):
        chunks[r], chunks[2 * n - 1 - r] = p[:, 0], p[:, 1]
    return torch.cat(chunks, dim=1)


def _step_plan(step, rank, world):
    """-> (q_off, q_len, k_len, causal) in chunks: rank `rank` attends, at ring step `step`, the query rows
    [q_off*c, (q_off+q_len)*c) of every local 2c-row batch element to the first k_len*c key rows of the block it holds."""
    if step == 0:
        return 0, 2, 2, True
    src = (rank - step) % world
    return (0, 2, 1, False) if src < rank else (1, 1, 2, False)


# ---------------------------------------------------------------------------------------------------------
# block operators (CUDA)
# ---------------------------------------------------------------------------------------------------------
class CudaBlockOps:
    """One launch per ring step on the tcgen05 kernels (libfa_b200.so).  q, k, v: (b*2c, h, d) local rows."""

    def __init__(self):
        import flash_attn_cuda
        self.lib = flash_attn_cuda
        self._cache = {}

    def _index(self, b, c, device):
        key = (b, c, device)
        if key not in self._cache:
            self._cache[key] = (torch.arange(0, (b + 1) * 2 * c, 2 * c, dtype=torch.int32, device=device),
                                torch.full((b,), c, dtype=torch.int32, device=device))
        return self._cache[key]

    def fwd(self, q, k, v, b, c, plan, scale):
        """-> o_blk (rows from q_off*c on, 16 bit), lse_blk (b, h, q_len*c) fp32"""
        q_off, q_len, k_len, causal = plan
        cu, half = self._index(b, c, q.device)
        (o, lse), _ = self.lib.fwd_ex(q[q_off * c:], k, v, cu, cu, q_len * c, k_len * c, 0.0, scale, causal,
                                      seqused_q=half if q_len == 1 else None, seqused_k=half if k_len == 1 else None)
        return o, lse

    def bwd(self, dout, q, k, v, out, lse, b, c, plan, scale):
        """lse: the FINAL (b, h, 2c) log-sum-exp.  -> dq_blk (rows from q_off*c on), dk_blk, dv_blk ((b*2c) rows, the
        first k_len*c of every batch element written)"""
        q_off, q_len, k_len, causal = plan
        cu, half = self._index(b, c, q.device)
        qv = q[q_off * c:]
        dq, dk, dv = torch.empty_like(qv), torch.empty_like(k), torch.empty_like(v)
        lse_v = lse if q_len == 2 else lse[:, :, q_off * c:(q_off + q_len) * c].contiguous()
        self.lib.bwd(dout[q_off * c:], qv, k, v, out[q_off * c:], lse_v, dq, dk, dv, cu, cu, q_len * c, k_len * c, 0.0,
                     scale, False, causal, None, seqused_q=half if q_len == 1 else None,
                     seqused_k=half if k_len == 1 else None)
        return dq, dk, dv

    def merge(self, o_acc, lse_acc, o_blk, lse_blk, b, c, plan, first):
        q_off, q_len = plan[0], plan[1]
        self.lib.merge_partials(o_acc[q_off * c:], lse_acc[:, :, q_off * c:], o_blk, lse_blk, b, q_len * c, first,
                                batch_rows=2 * c)

    def accumulate(self, dst, src, b, c, off, length):
        """dst fp32 (b*2c, h, d) rows [off*c, (off+length)*c) of every batch element += src (rows from off*c on)"""
        row = dst.shape[1] * dst.shape[2]
        self.lib.accumulate_f32(dst[off * c:], src, b, length * c * row, 2 * c * row, 2 * c * row)


_default_ops = None


def _ops_or_default(ops):
    global _default_ops
    if ops is not None:
        return ops
    if _default_ops is None:
        _default_ops = CudaBlockOps()       # kept: it caches the small index tensors of every (b, c) it has seen
    return _default_ops


# ---------------------------------------------------------------------------------------------------------
# communication
# ---------------------------------------------------------------------------------------------------------
class _Ring:
    """send to rank+1 / receive from rank-1, as one batched P2P so that NCCL pairs them without deadlock."""

    def __init__(self, group):
        self.group = group
        if not dist.is_initialized():          # a single process without a process group: a ring of one
            self.world, self.rank, self.next, self.prev = 1, 0, 0, 0
            return
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


# ---------------------------------------------------------------------------------------------------------
# forward / backward
# ---------------------------------------------------------------------------------------------------------
def ring_forward(q, k, v, softmax_scale, group=None, ops=None):
    """q, k, v: local (b, 2, c, h, d) zig-zag shards.  -> out (b, 2, c, h, d) (dtype of q), lse (b, h, 2c) fp32."""
    ops = _ops_or_default(ops)
    ring = _Ring(group)
    b, two, c, h, d = q.shape
    assert two == 2 and k.shape == q.shape and v.shape == q.shape
    assert c % 16 == 0, "chunk length must be a multiple of 16"
    rows = b * 2 * c
    qf = q.reshape(rows, h, d)
    o_acc = torch.empty(rows, h, d, dtype=torch.float32, device=q.device)
    lse_acc = torch.empty(b, h, 2 * c, dtype=torch.float32, device=q.device)
    kv = torch.stack([k.reshape(rows, h, d), v.reshape(rows, h, d)])       # (2, b*2c, h, d): one message
    kv_next = torch.empty_like(kv)
    for step in range(ring.world):
        reqs = ring.exchange([kv], [kv_next]) if step + 1 < ring.world else []   # the next block moves under this step's kernel
        plan = _step_plan(step, ring.rank, ring.world)
        o_blk, lse_blk = ops.fwd(qf, kv[0], kv[1], b, c, plan, softmax_scale)
        ops.merge(o_acc, lse_acc, o_blk, lse_blk, b, c, plan, step == 0)
        if reqs:
            ring.wait(reqs)
            kv, kv_next = kv_next, kv
    return o_acc.to(q.dtype).reshape(b, 2, c, h, d), lse_acc


def ring_backward(dout, q, k, v, out, lse, softmax_scale, group=None, ops=None):
    """-> dq, dk, dv, each (b, 2, c, h, d) in the dtype of q.  `lse` is ring_forward's (b, h, 2c) final log-sum-exp."""
    ops = _ops_or_default(ops)
    ring = _Ring(group)
    b, two, c, h, d = q.shape
    rows = b * 2 * c
    qf, of, gf = (t.reshape(rows, h, d) for t in (q, out, dout.contiguous()))
    dq_acc = torch.zeros(rows, h, d, dtype=torch.float32, device=q.device)
    kv = torch.stack([k.reshape(rows, h, d), v.reshape(rows, h, d)])
    dkv = torch.zeros(2, rows, h, d, dtype=torch.float32, device=q.device)       # follows its K/V block, one step behind
    kv_next, dkv_next = torch.empty_like(kv), torch.empty_like(dkv)
    g_reqs = None
    for step in range(ring.world):
        kv_reqs = ring.exchange([kv], [kv_next]) if step + 1 < ring.world else []
        plan = _step_plan(step, ring.rank, ring.world)
        q_off, q_len, k_len, _ = plan
        dq_blk, dk_blk, dv_blk = ops.bwd(gf, qf, kv[0], kv[1], of, lse, b, c, plan, softmax_scale)
        ops.accumulate(dq_acc, dq_blk, b, c, q_off, q_len)
        if g_reqs is not None:          # the accumulator of the block held now was sent by the previous rank a step ago
            ring.wait(g_reqs)
            dkv, dkv_next = dkv_next, dkv
        ops.accumulate(dkv[0], dk_blk, b, c, 0, k_len)
        ops.accumulate(dkv[1], dv_blk, b, c, 0, k_len)
        g_reqs = ring.exchange([dkv], [dkv_next])   # travels under the next step's kernel; after the last step it goes home
        if kv_reqs:
            ring.wait(kv_reqs)
            kv, kv_next = kv_next, kv
    ring.wait(g_reqs)
    dkv = dkv_next
    shape = (b, 2, c, h, d)
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
    q, k, v: local (b, 2, c, h, d) shards (see zigzag_shard).  Returns the local (b, 2, c, h, d) output shard."""
    return RingFlashAttnFunc.apply(q, k, v, softmax_scale, group, ops)
