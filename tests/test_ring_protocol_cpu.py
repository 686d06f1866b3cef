"""Model check of the peer-memory ring transport (flash_attn/ring_attention.py: _PeerLink / _PeerKV / _PeerDKV) on CPU.

The gloo test (tests/test_ring_cpu.py) runs the torch.distributed transport; the transport the GPUs use by default is a
hand-written cross-process protocol -- double buffers in CUDA-IPC memory, copy-engine pulls on side streams, monotone flag
counters (READY / CONSUMED) -- whose failure modes are stream-ordering and buffer-reuse races that no data-substituting test
can see.  Here the REAL protocol code runs against a simulated device runtime (`SimRt`, same interface as `_CudaRt`):

  * N simulated ranks each enqueue two full forward + backward ring passes (so the counters continue across calls and the
    buffers are recycled); nothing executes while they enqueue, exactly like asynchronous CUDA launches;
  * a scheduler then executes the queued stream operations, at every step picking a random runnable one among all streams of
    all ranks (in-order per stream; events, flag waits and flag writes with their CUDA meaning; a peer copy moves the
    symbolic content of a buffer);
  * the compute operations check what they see: the attention step t of rank r must find K/V block (r - t) mod N in its
    buffer, a dK/dV accumulator must belong to that same block and must not already hold this rank's contribution, and the
    accumulator that comes home must hold every rank's contribution exactly once;
  * a deadlock (queues not empty, nothing runnable) fails the test;
  * MUTATION checks: with the waits on one flag word disabled the same harness must FAIL -- it does detect the races it
    is there to detect.
"""
import random

import pytest
import torch

from flash_attn import ring_attention as ring


class SimAddr:
    def __init__(self, rank, off):
        self.rank, self.off = rank, off

    def __add__(self, n):
        return SimAddr(self.rank, self.off + n)

    key = property(lambda self: (self.rank, self.off))


class SimMem:
    def __init__(self, rank):
        self.rank = rank


class SimView:
    """a buffer inside a rank's shared allocation (sub-views [0] / [1] address its K / V or dK / dV halves)"""

    def __init__(self, rank, off, part=None):
        self.rank, self.off, self.part = rank, off, part

    def __getitem__(self, i):
        return SimView(self.rank, self.off, i)

    key = property(lambda self: (self.rank, self.off))


class SimEvent:
    done = False


class SimStream:
    def __init__(self, name):
        self.name, self.q = name, []


class SimWorld:
    def __init__(self, n, ignore_waits_on=None):
        self.n, self.flags, self.content, self.streams = n, {}, {}, []
        self.ignore_waits_on = ignore_waits_on          # mutation: waits on this flag word always pass
        self.homecomings = 0

    def stream(self, name):
        s = SimStream(name)
        self.streams.append(s)
        return s

    def runnable(self, op):
        kind = op[0]
        if kind == "wait_event":
            return op[1].done
        if kind == "flag_wait":
            _, addr, value = op
            if self.ignore_waits_on is not None and addr.off == 64 * self.ignore_waits_on:
                return True
            return self.flags.get(addr.key, 0) >= value
        return True

    def execute(self, op):
        kind = op[0]
        if kind == "record":
            op[1].done = True
        elif kind == "flag_write":
            _, addr, value = op
            assert value >= self.flags.get(addr.key, 0), "a flag counter went backwards"
            self.flags[addr.key] = value
        elif kind == "peer_copy":
            _, dst, src = op
            c = self.content.get(src.key)
            self.content[dst.key] = None if c is None else {k: (list(v) if isinstance(v, list) else v) for k, v in c.items()}
        elif kind == "fn":
            op[1]()

    def run(self, rng):
        while True:
            ready = [s for s in self.streams if s.q and self.runnable(s.q[0])]
            if not ready:
                stuck = [(s.name, s.q[0][0]) for s in self.streams if s.q]
                assert not stuck, f"deadlock: {stuck[:6]}"
                return
            s = rng.choice(ready)
            self.execute(s.q.pop(0))


class SimRt:
    """the interface of ring_attention._CudaRt on the simulated world"""

    def __init__(self, world, rank):
        self.w, self.rank = world, rank
        self.compute = world.stream(f"r{rank}.compute")
        self._n = 0

    def alloc(self, nbytes, zero_head):
        return SimMem(self.rank)

    def view(self, mem, off, shape, dtype):
        return SimView(mem.rank, off)

    def addr(self, x):
        return SimAddr(x.rank, getattr(x, "off", 0))

    def neighbours(self, mem):
        return SimAddr((self.rank - 1) % self.w.n, 0), SimAddr((self.rank + 1) % self.w.n, 0)

    def copy_(self, dst, src):          # the fill of buffer 0 with this rank's own K (part 0) / V (part 1)
        def fn(key=dst.key, part="kv"[dst.part], me=self.rank):
            c = self.w.content.get(key)
            c = dict(c) if c is not None and "k" in c else {"k": None, "v": None}
            c[part] = me
            self.w.content[key] = c
        self.compute.q.append(("fn", fn))

    def zero_(self, view):              # a fresh accumulator: it belongs to this rank's own block and holds nothing yet
        self.compute.q.append(("fn", lambda key=view.key, me=self.rank: self.w.content.__setitem__(key, {"block": me, "contrib": []})))

    def new_stream(self):
        self._n += 1
        return self.w.stream(f"r{self.rank}.side{self._n}")

    def current_stream(self):
        return self.compute

    def wait_stream(self, waiter, awaited):
        ev = SimEvent()
        awaited.q.append(("record", ev))
        waiter.q.append(("wait_event", ev))

    def record(self, stream):
        ev = SimEvent()
        stream.q.append(("record", ev))
        return ev

    def wait_event(self, stream, ev):
        stream.q.append(("wait_event", ev))

    def flag_write(self, addr, value, stream):
        stream.q.append(("flag_write", addr, value))

    def flag_wait_geq(self, addr, value, stream):
        stream.q.append(("flag_wait", addr, value))

    def peer_copy(self, dst_addr, src_addr, nbytes, stream):
        stream.q.append(("peer_copy", dst_addr, src_addr))


class SimOps:
    """the ring's block operators: no arithmetic, they CHECK what the transport hands them (at execution time)"""

    def __init__(self, world, rt, rows, h, d):
        self.w, self.rt, self.rank, self.shape = world, rt, rt.rank, (rows, h, d)
        self.n_fwd = self.n_bwd = self.n_acc = 0

    def _check_kv(self, k_view, v_view, step, what):
        def fn(kk=k_view.key, vk=v_view.key, want=(self.rank - step) % self.w.n, me=self.rank):
            c = self.w.content.get(kk)
            assert kk == vk and c is not None and c.get("k") == want and c.get("v") == want, \
                f"rank {me} {what} step {step}: K/V buffer holds {c}, expected block {want}"
        self.rt.compute.q.append(("fn", fn))

    def fwd(self, q, k, v, b, c, plan, scale):
        self._check_kv(k, v, self.n_fwd % self.w.n, "forward")
        self.n_fwd += 1
        return None, None

    def merge(self, *a):
        pass

    def dq_begin(self, q, b, c):
        return None

    def bwd(self, dout, q, k, v, out, lse, b, c, plan, scale, dq_state):
        self._check_kv(k, v, self.n_bwd % self.w.n, "backward")
        self.n_bwd += 1
        return None

    def dq_end(self, dq_state, q, b, c, scale):
        return torch.zeros_like(q)

    def accumulate_dkv(self, acc, dkv_blk, b, c, k_len):
        step = self.n_acc % self.w.n
        self.n_acc += 1

        def fn(key=acc.key, want=(self.rank - step) % self.w.n, me=self.rank):
            cont = self.w.content.get(key)
            assert cont is not None and cont.get("block") == want, \
                f"rank {me} step {step}: accumulator holds {cont}, expected the one of block {want}"
            assert me not in cont["contrib"], f"rank {me} step {step}: already contributed to {cont}"
            cont["contrib"].append(me)
        self.rt.compute.q.append(("fn", fn))

    def dkv_end(self, dkv, dtype):
        def fn(key=dkv.key, me=self.rank):
            cont = self.w.content.get(key)
            assert cont is not None and cont.get("block") == me and sorted(cont["contrib"]) == list(range(self.w.n)), \
                f"rank {me}: the accumulator that came home is {cont}"
            self.w.homecomings += 1
        self.rt.compute.q.append(("fn", fn))
        z = torch.zeros(self.shape)
        return z, z


class _SimRing:
    group = None

    def __init__(self, world, rank):
        self.world, self.rank, self.next, self.prev = world, rank, (rank + 1) % world, (rank - 1) % world


def simulate(n, passes, seed, ignore_waits_on=None):
    world = SimWorld(n, ignore_waits_on)
    b, c, h, d = 1, 16, 1, 8
    rows = b * 2 * c
    ranks = []
    for r in range(n):
        rt = SimRt(world, r)
        tr = ring.PeerTransport(_SimRing(n, r), None, rows, h, d, torch.float32, rt=rt, links={})
        tr.ring = _SimRing(n, r)
        ranks.append((tr, SimOps(world, rt, rows, h, d)))
    x = torch.zeros(b, 2, c, h, d)
    for _ in range(passes):                     # host code of every rank runs to completion: everything is only ENQUEUED
        for tr, ops in ranks:
            out, lse = ring.ring_forward(x, x, x, 1.0, None, ops, tr)
            ring.ring_backward(x, x, x, x, out, lse, 1.0, None, ops, tr)
    world.run(random.Random(seed))
    assert world.homecomings == n * passes
    return world


@pytest.mark.parametrize("n", [2, 3, 4, 8])
def test_peer_protocol_is_correct_under_random_interleavings(n):
    """two passes per rank, 40 random schedules: right block at every step, every accumulator home complete, no deadlock"""
    for seed in range(40):
        simulate(n, passes=2, seed=seed)


@pytest.mark.parametrize("word,name", [(ring._KV_READY, "KV_READY"), (ring._KV_CONSUMED, "KV_CONSUMED"),
                                       (ring._DKV_READY, "DKV_READY"), (ring._DKV_CONSUMED, "DKV_CONSUMED")])
def test_the_harness_detects_a_missing_wait(word, name):
    """mutation check: with the waits on one flag word disabled, some schedule must expose the race (wrong block in a buffer,
    an accumulator overwritten or pulled too early)"""
    failures = 0
    for n in (2, 3, 4):
        for seed in range(60):
            try:
                simulate(n, passes=2, seed=seed, ignore_waits_on=word)
            except AssertionError:
                failures += 1
    assert failures > 0, f"dropping the {name} waits went unnoticed: the simulator is too weak to trust"
