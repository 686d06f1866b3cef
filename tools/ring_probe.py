"""Where the 64K ring spends its time at N GPUs (run under torchrun):  python -m torch.distributed.run ... tools/ring_probe.py
 1. per-phase timeline of one fwd+bwd ring pass (flash_attn.ring_attention._trace): kernels vs stalls on transfers
 2. what NCCL isend/irecv delivers for the ring's message sizes (rank -> rank+1, all ranks at once)
 3. the same transfer as a direct peer copy over NVLink from a CUDA-IPC view of the previous rank's buffer
Rank 0 prints; nothing here is imported by the product."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "hazyresearch_flash-attention_b200"))
import torch
import torch.distributed as dist

world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
from flash_attn import ring_attention as ring


def say(*a):
    if rank == 0:
        print(*a, flush=True)


def timed(fn, reps=10):
    for _ in range(2):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


# ---- 1. ring timeline ---------------------------------------------------------------------------------------
S, b, h, d = 65536, 1, 12, 128
c = S // (2 * world)
g = torch.Generator(device=dev).manual_seed(5 + rank)
q, k, v = [torch.randn(b, 2, c, h, d, device=dev, dtype=torch.bfloat16, generator=g).requires_grad_() for _ in range(3)]
dout = torch.randn(b, 2, c, h, d, device=dev, dtype=torch.bfloat16, generator=g)


def step():
    out = ring.ring_flash_attn_func(q, k, v)
    torch.autograd.grad(out, (q, k, v), dout)


ms = timed(step, 3)
say(f"== ring 64K at N={world}: {ms:.2f} ms per fwd+bwd (max over ranks)")
ring._trace = []  # noqa
dist.barrier()
torch.cuda.synchronize()
step()
torch.cuda.synchronize()
tot = {}
for tag, st, e0, e1 in ring._trace:
    tot[tag] = tot.get(tag, 0.0) + e0.elapsed_time(e1)
ring._trace = None
tags = sorted(tot)
mine = torch.tensor([tot[t] for t in tags], device=dev)
allr = [torch.empty_like(mine) for _ in range(world)]
dist.all_gather(allr, mine)
if rank == 0:
    print("   phase (ms, summed over the ring steps) " + " ".join(f"r{r:<6d}" for r in range(world)))
    for i, t in enumerate(tags):
        print(f"   {t:<38s} " + " ".join(f"{a[i].item():7.2f}" for a in allr), flush=True)

# ---- 2. NCCL ring transfer ----------------------------------------------------------------------------------
nxt, prv = (rank + 1) % world, (rank - 1) % world
for name, nbytes in (("K/V block (bf16)", 2 * 2 * c * h * d * 2), ("dK/dV accumulator (fp32)", 2 * 2 * c * h * d * 4)):
    src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dst = torch.empty_like(src)

    def xfer():
        reqs = dist.batch_isend_irecv([dist.P2POp(dist.isend, src, nxt), dist.P2POp(dist.irecv, dst, prv)])
        for r in reqs:
            r.wait()

    ms = timed(xfer)
    say(f"== NCCL isend/irecv ring, {name}: {nbytes / 2**20:.0f} MiB in {ms:.3f} ms = {nbytes / ms / 1e6:.0f} GB/s per GPU and direction")

# ---- 3. direct peer copy through CUDA IPC -------------------------------------------------------------------
try:
    from torch.multiprocessing.reductions import reduce_tensor
    nbytes = 2 * 2 * c * h * d * 4
    mine_buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    mine_buf.fill_(rank + 1)
    fn, args = reduce_tensor(mine_buf)
    handles = [None] * world
    dist.all_gather_object(handles, args)
    peer = fn(*handles[prv])          # view of the previous rank's buffer, living on ITS device
    local_dst = torch.empty_like(mine_buf)
    torch.cuda.synchronize()
    dist.barrier()
    local_dst.copy_(peer)
    torch.cuda.synchronize()
    ok = bool((local_dst[:: 1 << 20] == prv + 1).all())
    ms = timed(lambda: local_dst.copy_(peer, non_blocking=True))
    say(f"== CUDA-IPC peer copy (pull from rank-1 over NVLink), {nbytes / 2**20:.0f} MiB: {ms:.3f} ms = {nbytes / ms / 1e6:.0f} GB/s per GPU, "
        f"data correct: {ok}, peer tensor device {peer.device}")
    side = torch.cuda.Stream()

    def both():        # pull while a kernel-sized local job runs: do they overlap?
        with torch.cuda.stream(side):
            local_dst.copy_(peer, non_blocking=True)
        torch.cuda.current_stream().wait_stream(side)
    ms2 = timed(both)
    say(f"   same on a side stream: {ms2:.3f} ms")
    dist.barrier()
    del peer
except Exception as e:      # noqa: BLE001  (a probe: report and go on)
    say(f"== CUDA-IPC peer copy: NOT available here: {type(e).__name__}: {e}")
dist.barrier()
dist.destroy_process_group()
