"""Host logic of the causal K/V ring (flash_attn/ring_attention.py) at world_size 2 and 4 with the gloo backend on CPU.

The per-block attention / merge operators are injected: here they are an fp32 PyTorch restatement with exactly the
kernels' contract (block forward returns (O, lse); block backward takes the FINAL lse and D = rowsum(dO o O_final)),
so what is tested is the ring schedule, the zig-zag sharding, the P2P exchange and the gradient routing.  The result
is compared with the oracle's causal attention over the whole sequence.  (On a GPU box the same functions run with
CudaBlockOps -- tests/test_ring_gpu.py.)"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import attention_ref as orc


class TorchBlockOps:
    """fp32 block operators with the CUDA kernels' semantics (test infrastructure)."""

    @staticmethod
    def _scores(q, k, n, c, scale, causal):
        h, d = q.shape[1:]
        qq, kk = q.reshape(n, c, h, d).float(), k.reshape(n, c, h, d).float()
        s = torch.einsum('bthd,bshd->bhts', qq, kk) * scale
        if causal:
            s = s.masked_fill(orc.causal_mask_topleft(c, c, q.device), float('-inf'))
        return s

    def fwd(self, q, k, v, n, c, scale, causal):
        h, d = q.shape[1:]
        s = self._scores(q, k, n, c, scale, causal)
        lse = torch.logsumexp(s, dim=-1)
        p = torch.exp(s - lse[..., None])
        o = torch.einsum('bhts,bshd->bthd', p, v.reshape(n, c, h, d).float())
        return o.reshape(n * c, h, d).to(q.dtype), lse

    def bwd(self, dout, q, k, v, out, lse, n, c, scale, causal):
        h, d = q.shape[1:]
        s = self._scores(q, k, n, c, scale, causal)
        p = torch.exp(s - lse[..., None])                       # FINAL lse: probabilities of the whole row
        g = dout.reshape(n, c, h, d).float()
        D = (g * out.reshape(n, c, h, d).float()).sum(-1).permute(0, 2, 1)      # (n, h, c)
        dp = torch.einsum('bthd,bshd->bhts', g, v.reshape(n, c, h, d).float())
        ds = p * (dp - D[..., None])
        dq = torch.einsum('bhts,bshd->bthd', ds, k.reshape(n, c, h, d).float()) * scale
        dk = torch.einsum('bhts,bthd->bshd', ds, q.reshape(n, c, h, d).float()) * scale
        dv = torch.einsum('bhts,bthd->bshd', p, g)
        return tuple(t.reshape(n * c, h, d).to(q.dtype) for t in (dq, dk, dv))

    def merge(self, o_acc, lse_acc, lse_out, o_blk, lse_blk, n, c, first):
        h, d = o_acc.shape[1:]
        if first:
            o_acc.copy_(o_blk.float())
            lse_out.copy_(lse_blk)
            return
        new = torch.logaddexp(lse_acc, lse_blk)
        wa = torch.exp(lse_acc - new).permute(0, 2, 1).reshape(n * c, h, 1)
        wb = torch.exp(lse_blk - new).permute(0, 2, 1).reshape(n * c, h, 1)
        o_acc.mul_(wa).add_(o_blk.float() * wb)
        lse_out.copy_(new)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, b, S, h, d, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from flash_attn import ring_attention as ring
        g = torch.Generator().manual_seed(1234)
        q, k, v, dout = [torch.randn(b, S, h, d, generator=g) for _ in range(4)]
        ql, kl, vl = [ring.zigzag_shard(t, world, rank).requires_grad_() for t in (q, k, v)]
        gl = ring.zigzag_shard(dout, world, rank)
        out = ring.ring_flash_attn_func(ql, kl, vl, None, None, TorchBlockOps())
        dql, dkl, dvl = torch.autograd.grad(out, (ql, kl, vl), gl)
        gathered = []
        for t in (out.detach(), dql, dkl, dvl):
            parts = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(parts, t.contiguous())
            gathered.append(ring.zigzag_unshard(parts))
        if rank == 0:
            qf, kf, vf = [t.clone().requires_grad_() for t in (q, k, v)]
            ref, _ = orc.attention_ref(qf, kf, vf, None, None, 0.0, None, causal=True)
            rq, rk, rv = torch.autograd.grad(ref, (qf, kf, vf), dout)
            errs = [(a - r).abs().max().item() for a, r in zip(gathered, (ref, rq, rk, rv))]
            ret.put(errs)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_ring_matches_full_causal_attention(world):
    b, S, h, d = 2, 16 * 2 * world * 2, 2, 16
    ctx = mp.get_context("spawn")
    ret = ctx.SimpleQueue()
    mp.spawn(_worker, args=(world, _free_port(), b, S, h, d, ret), nprocs=world, join=True)
    errs = ret.get()
    assert max(errs) < 2e-5, errs


def test_block_plan_covers_the_causal_triangle_exactly_once():
    from flash_attn.ring_attention import _block_plan
    for world in (1, 2, 3, 4, 8):
        seen = {}
        for rank in range(world):
            chunk = {0: rank, 1: 2 * world - 1 - rank}
            for step in range(world):
                src = (rank - step) % world
                kchunk = {0: src, 1: 2 * world - 1 - src}
                for qh, kh, causal in _block_plan(step, rank, world):
                    pairs = [(0, 0), (1, 1)] if qh is None else [(qh, kh)]
                    for a, bb in pairs:
                        key = (chunk[a], kchunk[bb])
                        assert key not in seen
                        seen[key] = causal
                        assert causal == (key[0] == key[1])
        want = {(i, j) for i in range(2 * world) for j in range(2 * world) if j <= i}
        assert set(seen) == want
        # balance: every rank computes the same number of c x c blocks
