"""Causal K/V ring attention across the GPUs of one NVLink box (context parallelism for long sequences).

New functionality with respect to the reference (which is single-GPU; its README claims sequence lengths "up to 64K",
README.md:75, only through the in-kernel K-block loop).  The merge rule is the reference's own inter-K-block rule
(csrc/flash_attn/src/fmha_fprop_kernel_1xN.h:606-650: lse' = log(e^lse + e^lse_blk), O rescaled by exp(lse - lse'))
lifted from K blocks inside one kernel to K/V chunks living on different GPUs.

Layout ("zig-zag", balances causal work): the sequence is cut into 2N chunks of c tokens; rank r owns chunks r and
2N-1-r.  A local tensor is (b, 2, c, h, d): the two chunks of a batch element are adjacent, so a batch element is one
packed "sequence" of 2c rows whose halves a kernel launch can address through `seqused_q` / `seqused_k` (fa_params).
K/V blocks travel rank -> rank+1 while the attention kernel works on the block already here: on GPUs as copy-engine pulls
out of the predecessor's CUDA-IPC-mapped buffer over NVLink, ordered by flag kernels (`PeerTransport`, the default; NCCL
isend / irecv cost a constant ~0.6 ms per exchange on an 8 x B200 box and left the ring transfer-bound), or through
torch.distributed P2P (`DistTransport`: NCCL, and gloo in the CPU tests).  At step t a rank holds the block of rank
src = (r - t) mod N and computes exactly two c x c block products' worth of work in ONE kernel launch:

    t = 0      [qa; qb] x [ka; kb] causal     (a precedes b in the global order: plain causal attention over 2c rows)
    src < r    [qa; qb] x ka       full       (both local query chunks come after the early chunk of an earlier rank)
    src > r    qb       x [ka; kb] full       (only the late query chunk sees a later rank's chunks -- both of them)

Backward: a second ring in which every K/V block is followed, one step behind, by its fp32 (dK, dV) accumulator, which
is home again after N hops.  dQ stays local and never leaves the backward kernel's own fp32 workspace accumulator between
steps (FA_FLAG_KEEP_DQ_ACCUM): every step adds its partial sums to the same tile slots and one conversion pass runs at the
end (fa_bwd_dq_finalize).  Each block product is differentiated with the FINAL softmax_lse and D = rowsum(dO o O), so the
per-block gradients simply add up.  The accumulator exchange of step t overlaps the attention kernel of step t+1 (it is only
needed when that kernel's dK / dV are added), the K/V exchange overlaps the kernel of the same step.  Not capturable in a CUDA
graph with the peer transport (the flag values of a pass are host-side counters baked into the launches).

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

    def dq_begin(self, q, b, c):
        """dQ of a whole ring pass accumulates in fp32 INSIDE the backward's own workspace (FA_FLAG_KEEP_DQ_ACCUM): every step
        adds its partial sums to the same tile slots, one conversion pass at the end -- no per-step 16-bit dQ block, no
        separate accumulate pass.  -> state for bwd() / dq_end()."""
        cu, _ = self._index(b, c, q.device)
        ws = torch.zeros(self.lib.bwd_workspace_bytes(q, cu, 2 * c), dtype=torch.uint8, device=q.device)
        return ws

    def bwd(self, dout, q, k, v, out, lse, b, c, plan, scale, dq_state):
        """lse: the FINAL (b, h, 2c) log-sum-exp.  dQ is added into dq_state; -> dkv_blk (2, b*2c, h, d) 16 bit, the first
        k_len*c rows of every batch element written.  The query sub-range is addressed through cu_seqlens_q (+ q_off*c) on
        the UNSHIFTED tensors, so that every step maps its query tiles to the same accumulator slots."""
        q_off, q_len, k_len, causal = plan
        cu, half = self._index(b, c, q.device)
        cu_q = cu if q_off == 0 else self._shifted(cu, q_off * c)
        dkv = torch.empty((2,) + tuple(k.shape), dtype=k.dtype, device=k.device)
        lse_v = lse if q_len == 2 else lse[:, :, q_off * c:(q_off + q_len) * c].contiguous()
        self.lib.bwd(dout, q, k, v, out, lse_v, None, dkv[0], dkv[1], cu_q, cu, q_len * c, k_len * c, 0.0, scale, False, causal,
                     None, seqused_q=half if q_len == 1 else None, seqused_k=half if k_len == 1 else None, dq_workspace=dq_state)
        return dkv

    def dq_end(self, dq_state, q, b, c, scale):
        cu, _ = self._index(b, c, q.device)
        return self.lib.bwd_dq_finalize(torch.empty_like(q), cu, 2 * c, scale, dq_state)

    def _shifted(self, cu, off):
        key = ("shift", cu.data_ptr(), off)
        if key not in self._cache:
            self._cache[key] = cu + off
        return self._cache[key]

    def merge(self, o_acc, lse_acc, o_blk, lse_blk, b, c, plan, first):
        q_off, q_len = plan[0], plan[1]
        self.lib.merge_partials(o_acc[q_off * c:], lse_acc[:, :, q_off * c:], o_blk, lse_blk, b, q_len * c, first,
                                batch_rows=2 * c)

    def dkv_end(self, dkv, dtype):
        """the own block's fp32 (dK, dV) sums, home after the last hop -> 16 bit"""
        return dkv[0].to(dtype), dkv[1].to(dtype)

    def accumulate_dkv(self, acc, dkv_blk, b, c, k_len):
        """acc fp32 (2, b*2c, h, d) += dkv_blk (same shape, 16 bit) on the first k_len*c rows of every batch element of both
        halves: one launch ((dK, dV) x batch = 2b runs of k_len*c rows)"""
        row = acc.shape[2] * acc.shape[3]
        self.lib.accumulate_f32(acc, dkv_blk, 2 * b, k_len * c * row, 2 * c * row, 2 * c * row)


_default_ops = None

# Optional timeline of one call (tools/ring_probe.py): when set to a list, every phase appends (tag, step, start event,
# end event), recorded on the compute stream -- a `wait_*` span is the time the compute stream stalled on a transfer.
_trace = None


class _span:
    __slots__ = ("tag", "step", "e0")

    def __init__(self, tag, step):
        self.tag, self.step, self.e0 = tag, step, None

    def __enter__(self):
        if _trace is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *exc):
        if self.e0 is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            _trace.append((self.tag, self.step, self.e0, e1))
        return False


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
    """rank / neighbours of this process in `group`, and the batched send-to-next / receive-from-previous of the
    torch.distributed transport (one batch so that NCCL pairs the two without deadlock)."""

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
# transports.  Both expose the same two objects to ring_forward / ring_backward:
#   K/V stream        .k, .v (the block held now) | .prefetch(step): start moving the next block | .advance(): it is
#                     current now (the compute stream is ordered after its arrival)
#   dK/dV accumulator .cur fp32 (2, rows, h, d) of the block held now | .arrive(): the accumulator that follows this
#                     block, one step behind, has come in | .depart(step): hand it on | .home(): the own block's sum
# ---------------------------------------------------------------------------------------------------------
class _DistKV:
    """torch.distributed P2P (NCCL on GPUs, gloo in the CPU tests)."""

    def __init__(self, ring, k_rows, v_rows):
        self.ring = ring
        self.kv = torch.stack([k_rows, v_rows])       # (2, rows, h, d): one message
        self.nxt = torch.empty_like(self.kv)
        self.reqs = []

    k = property(lambda self: self.kv[0])
    v = property(lambda self: self.kv[1])

    def prefetch(self, step):
        self.reqs = self.ring.exchange([self.kv], [self.nxt])

    def advance(self):
        self.ring.wait(self.reqs)
        self.kv, self.nxt = self.nxt, self.kv


class _DistDKV:
    def __init__(self, ring, rows, h, d, device):
        self.ring = ring
        self.cur = torch.zeros(2, rows, h, d, dtype=torch.float32, device=device)
        self.nxt = torch.empty_like(self.cur)
        self.reqs = None

    def arrive(self):
        self.ring.wait(self.reqs)
        self.cur, self.nxt = self.nxt, self.cur

    def depart(self, step):
        self.reqs = self.ring.exchange([self.cur], [self.nxt])

    def home(self):
        self.arrive()
        return self.cur


class DistTransport:
    name = "dist"

    def __init__(self, ring):
        self.ring = ring

    def kv_stream(self, k_rows, v_rows):
        return _DistKV(self.ring, k_rows, v_rows)

    def dkv_stream(self, rows, h, d, device):
        return _DistDKV(self.ring, rows, h, d, device)


_KV_READY, _KV_CONSUMED, _DKV_READY, _DKV_CONSUMED = 0, 1, 2, 3     # flag words (64 bytes apart)
_FLAG_BYTES = 256


class _CudaRt:
    """Everything the peer transport asks of the device: streams, events, the flag / copy primitives of libfa_b200.so and
    the mapping of the neighbours' memory.  The protocol classes below talk to the device only through an object like this
    one, so that tests/test_ring_protocol_cpu.py can run the very same protocol code on a simulator that executes the queued
    stream operations of N ranks in random interleavings."""

    def __init__(self, ring, device):
        import flash_attn_cuda
        self.lib, self.ring, self.device = flash_attn_cuda, ring, device

    # --- memory -------------------------------------------------------------------------------------------
    def alloc(self, nbytes, zero_head):
        mem = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        mem[:zero_head].zero_()
        return mem

    def view(self, mem, off, shape, dtype):
        n = 1
        for x in shape:
            n *= x
        return mem[off:off + n * torch.empty((), dtype=dtype).element_size()].view(dtype).view(shape)

    def addr(self, t):
        return t.data_ptr()

    def neighbours(self, mem):
        """-> (address of the predecessor's `mem`, address of the successor's): CUDA IPC handles exchanged over the group"""
        ring = self.ring
        handles = [None] * ring.world
        dist.all_gather_object(handles, self.lib.ipc_export(mem), group=ring.group)
        grank = {(dist.get_global_rank(ring.group, i) if ring.group is not None else i): i for i in range(ring.world)}
        opened = {}                       # a handle may be mapped once per process: at world 2 both neighbours are the same rank
        for peer in (ring.prev, ring.next):
            if peer not in opened:
                handle, offset = handles[grank[peer]]
                opened[peer] = self.lib.ipc_open(self.device, handle) + offset
        torch.cuda.synchronize(self.device)
        dist.barrier(group=ring.group)                      # every flag word is zero before anybody writes one
        return opened[ring.prev], opened[ring.next]

    def copy_(self, dst, src):            # on the current stream
        dst.copy_(src)

    def zero_(self, t):
        t.zero_()

    # --- streams and events ---------------------------------------------------------------------------------
    def new_stream(self):
        return torch.cuda.Stream(self.device)

    def current_stream(self):
        return torch.cuda.current_stream(self.device)

    def wait_stream(self, waiter, awaited):
        waiter.wait_stream(awaited)

    def record(self, stream):
        ev = torch.cuda.Event()
        ev.record(stream)
        return ev

    def wait_event(self, stream, ev):
        stream.wait_event(ev)

    # --- the primitives of include/fa_b200.h ----------------------------------------------------------------
    def flag_write(self, addr, value, stream):
        self.lib.flag_write(addr, value, stream.cuda_stream)

    def flag_wait_geq(self, addr, value, stream):
        self.lib.flag_wait_geq(addr, value, stream.cuda_stream)

    def peer_copy(self, dst_addr, src_addr, nbytes, stream):
        self.lib.peer_copy(dst_addr, src_addr, nbytes, stream.cuda_stream)


class _PeerLink:
    """Memory this rank shares with its ring neighbours through CUDA IPC, for one message shape:
    [4 flag words | K/V buffer 0 | K/V buffer 1 | dK/dV buffer 0 | dK/dV buffer 1], the same layout on every rank.
    Blocks are PULLED: a rank copies the block its predecessor holds out of the predecessor's buffer with a copy-engine
    transfer over NVLink (measured 631 GB/s for 96 MiB on an 8 x B200 NVSwitch box; NCCL isend / irecv took a constant
    ~0.6 ms per exchange, 83 - 163 GB/s, profiles/r02_ring_probe_n8_nccl.log).  Ordering across processes: monotone 64-bit
    counters, never reset, written into the WAITER's memory (fa_flag_write) and waited for by a one-thread kernel on the
    waiter's copy stream (fa_flag_wait_geq; traps after 20 s instead of hanging):
        KV_READY      written by the predecessor: number of K/V blocks it has made available (cumulative over all calls)
        KV_CONSUMED   written by the successor: number of pulls it has completed out of this rank's buffers
        DKV_READY / DKV_CONSUMED   the same for the fp32 dK/dV accumulator that follows its block.
    Calls are collective and identical on all ranks, so every rank derives the same counter values on the host."""

    def __init__(self, rt, world, kv_bytes, dkv_bytes):
        assert world > 1, "a ring of one has no peers (a process cannot map its own IPC handle)"
        self.rt, self.world = rt, world
        self.kv_bytes, self.dkv_bytes = kv_bytes, dkv_bytes
        self.kv_off = [_FLAG_BYTES, _FLAG_BYTES + kv_bytes]
        self.dkv_off = [_FLAG_BYTES + 2 * kv_bytes, _FLAG_BYTES + 2 * kv_bytes + dkv_bytes]
        self.mem = rt.alloc(_FLAG_BYTES + 2 * kv_bytes + 2 * dkv_bytes, _FLAG_BYTES)
        self.prev_ptr, self.next_ptr = rt.neighbours(self.mem)        # addresses of the neighbours' `mem`
        self.s_kv, self.s_dkv = rt.new_stream(), rt.new_stream()
        self.kv_ready = self.kv_consumed = self.dkv_ready = self.dkv_consumed = 0     # cumulative, same on all ranks

    def my_flag(self, word):
        return self.rt.addr(self.mem) + 64 * word

    def view(self, off, shape, dtype):
        return self.rt.view(self.mem, off, shape, dtype)


_links = {}


class _PeerKV:
    def __init__(self, link, k_rows, v_rows):
        L, rt = link, link.rt
        self.link = L
        shape = (2,) + tuple(k_rows.shape)
        self.buf = [L.view(L.kv_off[i], shape, k_rows.dtype) for i in (0, 1)]
        self.idx, self.arrived = 0, None
        self.R, self.C = L.kv_ready, L.kv_consumed
        L.kv_ready += L.world - 1                # a call announces world-1 blocks and pulls world-1 blocks
        L.kv_consumed += L.world - 1
        # buffer 0 may still be the source of the successor's last pull of the previous call
        rt.flag_wait_geq(L.my_flag(_KV_CONSUMED), self.C, L.s_kv)
        rt.wait_stream(rt.current_stream(), L.s_kv)
        rt.copy_(self.buf[0][0], k_rows)
        rt.copy_(self.buf[0][1], v_rows)

    k = property(lambda self: self.buf[self.idx][0])
    v = property(lambda self: self.buf[self.idx][1])

    def prefetch(self, t):
        """pull the block the predecessor holds at step t into the other buffer (it is this rank's block of step t+1)"""
        L, rt, s = self.link, self.link.rt, self.link.s_kv
        # the target buffer was read by the kernels of step t-1; at t = 0 the fill of buffer 0 must have finished
        rt.wait_stream(s, rt.current_stream())
        if t == 0:
            rt.flag_write(L.next_ptr + 64 * _KV_READY, self.R + 1, s)              # my own K/V is block #0
        rt.flag_wait_geq(L.my_flag(_KV_READY), self.R + t + 1, s)                  # the predecessor's block #t is there
        rt.flag_wait_geq(L.my_flag(_KV_CONSUMED), self.C + t, s)                   # the successor is done with my target buffer
        rt.peer_copy(rt.addr(self.buf[(t + 1) % 2]), L.prev_ptr + L.kv_off[t % 2], L.kv_bytes, s)
        if t + 1 <= L.world - 2:
            rt.flag_write(L.next_ptr + 64 * _KV_READY, self.R + t + 2, s)          # block #(t+1) can be pulled from me
        rt.flag_write(L.prev_ptr + 64 * _KV_CONSUMED, self.C + t + 1, s)
        self.arrived = rt.record(s)

    def advance(self):
        rt = self.link.rt
        rt.wait_event(rt.current_stream(), self.arrived)
        self.idx ^= 1


class _PeerDKV:
    def __init__(self, link, rows, h, d):
        L, rt = link, link.rt
        self.link = L
        self.buf = [L.view(L.dkv_off[i], (2, rows, h, d), torch.float32) for i in (0, 1)]
        self.idx, self.arrived = 0, None
        self.R, self.C = L.dkv_ready, L.dkv_consumed
        L.dkv_ready += L.world                   # world announcements and world pulls per call (the last one goes home)
        L.dkv_consumed += L.world
        rt.flag_wait_geq(L.my_flag(_DKV_CONSUMED), self.C, L.s_dkv)
        rt.wait_stream(rt.current_stream(), L.s_dkv)
        rt.zero_(self.buf[0])

    cur = property(lambda self: self.buf[self.idx])

    def depart(self, t):
        """step t's sums are in `cur`: announce them to the successor, and pull the predecessor's (the accumulator of the
        block this rank holds at step t+1) under the next step's kernel"""
        L, rt, s = self.link, self.link.rt, self.link.s_dkv
        rt.wait_stream(s, rt.current_stream())                                       # the accumulate kernels of step t
        rt.flag_write(L.next_ptr + 64 * _DKV_READY, self.R + t + 1, s)
        rt.flag_wait_geq(L.my_flag(_DKV_READY), self.R + t + 1, s)
        rt.flag_wait_geq(L.my_flag(_DKV_CONSUMED), self.C + t, s)                  # the successor has pulled what my target buffer held
        rt.peer_copy(rt.addr(self.buf[(t + 1) % 2]), L.prev_ptr + L.dkv_off[t % 2], L.dkv_bytes, s)
        rt.flag_write(L.prev_ptr + 64 * _DKV_CONSUMED, self.C + t + 1, s)
        self.arrived = rt.record(s)

    def arrive(self):
        rt = self.link.rt
        rt.wait_event(rt.current_stream(), self.arrived)
        self.idx ^= 1

    def home(self):
        self.arrive()
        return self.cur


class PeerTransport:
    name = "peer"

    def __init__(self, ring, device, rows, h, d, dtype, rt=None, links=None):
        es = torch.empty((), dtype=dtype).element_size()
        links = _links if links is None else links
        key = (id(ring.group), ring.world, ring.rank, getattr(device, "index", None), rows, h, d, dtype)
        if key not in links:
            links[key] = _PeerLink(rt or _CudaRt(ring, device), ring.world, 2 * rows * h * d * es, 2 * rows * h * d * 4)
        self.link = links[key]

    def kv_stream(self, k_rows, v_rows):
        return _PeerKV(self.link, k_rows, v_rows)

    def dkv_stream(self, rows, h, d, device):
        return _PeerDKV(self.link, rows, h, d)


default_transport = None      # None = "peer" for CUDA tensors on an NCCL group of more than one rank, else "dist"


def _make_transport(ring, q, rows, transport):
    if transport is not None and not isinstance(transport, str):
        return transport                    # a ready-made transport object (tests/test_ring_protocol_cpu.py)
    transport = transport or default_transport
    if transport is None:
        peer_ok = q.is_cuda and ring.world > 1 and dist.is_initialized() and dist.get_backend(ring.group) == "nccl"
        transport = "peer" if peer_ok else "dist"
    if transport == "peer" and ring.world > 1:
        return PeerTransport(ring, q.device, rows, q.shape[-2], q.shape[-1], q.dtype)
    assert transport in ("dist", "peer"), f"unknown ring transport {transport!r}"
    return DistTransport(ring)


# ---------------------------------------------------------------------------------------------------------
# forward / backward
# ---------------------------------------------------------------------------------------------------------
def ring_forward(q, k, v, softmax_scale, group=None, ops=None, transport=None):
    """q, k, v: local (b, 2, c, h, d) zig-zag shards.  -> out (b, 2, c, h, d) (dtype of q), lse (b, h, 2c) fp32."""
    ops = _ops_or_default(ops)
    ring = getattr(transport, "ring", None) or _Ring(group)
    b, two, c, h, d = q.shape
    assert two == 2 and k.shape == q.shape and v.shape == q.shape
    assert c % 16 == 0, "chunk length must be a multiple of 16"
    rows = b * 2 * c
    qf = q.reshape(rows, h, d)
    o_acc = torch.empty(rows, h, d, dtype=torch.float32, device=q.device)
    lse_acc = torch.empty(b, h, 2 * c, dtype=torch.float32, device=q.device)
    kvs = _make_transport(ring, q, rows, transport).kv_stream(k.reshape(rows, h, d), v.reshape(rows, h, d))
    for step in range(ring.world):
        more = step + 1 < ring.world
        if more:
            kvs.prefetch(step)          # the next block moves under this step's kernel
        plan = _step_plan(step, ring.rank, ring.world)
        with _span("fwd_kernels", step):
            o_blk, lse_blk = ops.fwd(qf, kvs.k, kvs.v, b, c, plan, softmax_scale)
            ops.merge(o_acc, lse_acc, o_blk, lse_blk, b, c, plan, step == 0)
        if more:
            with _span("fwd_wait_kv", step):
                kvs.advance()
    return o_acc.to(q.dtype).reshape(b, 2, c, h, d), lse_acc


def ring_backward(dout, q, k, v, out, lse, softmax_scale, group=None, ops=None, transport=None):
    """-> dq, dk, dv, each (b, 2, c, h, d) in the dtype of q.  `lse` is ring_forward's (b, h, 2c) final log-sum-exp."""
    ops = _ops_or_default(ops)
    ring = getattr(transport, "ring", None) or _Ring(group)
    b, two, c, h, d = q.shape
    rows = b * 2 * c
    qf, of, gf = (t.reshape(rows, h, d) for t in (q, out, dout.contiguous()))
    tr = _make_transport(ring, q, rows, transport)
    kvs = tr.kv_stream(k.reshape(rows, h, d), v.reshape(rows, h, d))
    acc = tr.dkv_stream(rows, h, d, q.device)       # follows its K/V block around the ring, one step behind
    dq_state = ops.dq_begin(qf, b, c)               # dQ stays local and accumulates in fp32 over all steps
    for step in range(ring.world):
        more = step + 1 < ring.world
        if more:
            kvs.prefetch(step)
        plan = _step_plan(step, ring.rank, ring.world)
        with _span("bwd_kernels", step):
            dkv_blk = ops.bwd(gf, qf, kvs.k, kvs.v, of, lse, b, c, plan, softmax_scale, dq_state)
        if step > 0:                    # the accumulator of the block held now left the previous rank a step ago
            with _span("bwd_wait_dkv", step):
                acc.arrive()
        with _span("bwd_accumulate", step):
            ops.accumulate_dkv(acc.cur, dkv_blk, b, c, plan[2])
        acc.depart(step)                # travels under the next step's kernel; after the last step it goes home
        if more:
            with _span("bwd_wait_kv", step):
                kvs.advance()
    dq = ops.dq_end(dq_state, qf, b, c, softmax_scale)
    with _span("bwd_wait_dkv", ring.world):
        dkv = acc.home()
    dk, dv = ops.dkv_end(dkv, q.dtype)
    shape = (b, 2, c, h, d)
    return dq.reshape(shape), dk.reshape(shape), dv.reshape(shape)


class RingFlashAttnFunc(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, softmax_scale, group, ops, transport):
        softmax_scale = q.shape[-1] ** (-0.5) if softmax_scale is None else softmax_scale
        out, lse = ring_forward(q, k, v, softmax_scale, group, ops, transport)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.softmax_scale, ctx.group, ctx.ops, ctx.transport = softmax_scale, group, ops, transport
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out, lse = ctx.saved_tensors
        dq, dk, dv = ring_backward(dout, q, k, v, out, lse, ctx.softmax_scale, ctx.group, ctx.ops, ctx.transport)
        return dq, dk, dv, None, None, None, None


def ring_flash_attn_func(q, k, v, softmax_scale=None, group=None, ops=None, transport=None):
    """Causal attention over a sequence sharded zig-zag across the ranks of `group`.
    q, k, v: local (b, 2, c, h, d) shards (see zigzag_shard).  Returns the local (b, 2, c, h, d) output shard.
    transport: "peer" (CUDA-IPC peer memory over NVLink; the default on GPUs), "dist" (torch.distributed P2P) or None."""
    return RingFlashAttnFunc.apply(q, k, v, softmax_scale, group, ops, transport)
