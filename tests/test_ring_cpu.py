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
    """fp32 block operators with the CUDA kernels' semantics (test infrastructure).  All tensors are (b*2c, h, d)
    local rows; a plan (q_off, q_len, k_len, causal) addresses chunk ranges of every 2c-row batch element."""

    @staticmethod
    def _parts(b, c, plan, *ts):
        q_off, q_len, k_len, _ = plan
        h, d = ts[0].shape[1:]
        out = []
        for t, is_q in ts[1:]:
            t4 = t.reshape(-1, h, d)
            pad = b * 2 * c - t4.shape[0]               # tensors may start at row q_off*c
            t4 = torch.cat([t4.new_zeros(pad, h, d), t4]).reshape(b, 2 * c, h, d).float()
            out.append(t4[:, q_off * c:(q_off + q_len) * c] if is_q else t4[:, :k_len * c])
        return out

    @staticmethod
    def _scores(qq, kk, scale, causal):
        s = torch.einsum('bthd,bshd->bhts', qq, kk) * scale
        if causal:
            s = s.masked_fill(orc.causal_mask_topleft(qq.shape[1], kk.shape[1], qq.device), float('-inf'))
        return s

    @staticmethod
    def _place(x, b, c, off, dtype):
        """(b, n, h, d) -> rows from off*c on of a (b*2c, h, d) tensor"""
        h, d = x.shape[2:]
        full = x.new_zeros(b, 2 * c, h, d)
        full[:, off * c:off * c + x.shape[1]] = x
        return full.reshape(b * 2 * c, h, d)[off * c:].to(dtype)

    def fwd(self, q, k, v, b, c, plan, scale):
        qq, kk, vv = self._parts(b, c, plan, q, (q, True), (k, False), (v, False))
        s = self._scores(qq, kk, scale, plan[3])
        lse = torch.logsumexp(s, dim=-1)
        o = torch.einsum('bhts,bshd->bthd', torch.exp(s - lse[..., None]), vv)
        return self._place(o, b, c, plan[0], q.dtype), lse

    def dq_begin(self, q, b, c):
        return torch.zeros(q.shape, dtype=torch.float32)

    def bwd(self, dout, q, k, v, out, lse, b, c, plan, scale, dq_state):
        q_off, q_len, k_len, causal = plan
        qq, gg, oo, kk, vv = self._parts(b, c, plan, q, (q, True), (dout, True), (out, True), (k, False), (v, False))
        lse_v = lse[:, :, q_off * c:(q_off + q_len) * c]
        p = torch.exp(self._scores(qq, kk, scale, causal) - lse_v[..., None])   # FINAL lse: probabilities of the whole row
        D = (gg * oo).sum(-1).permute(0, 2, 1)                                   # (b, h, t)
        dp = torch.einsum('bthd,bshd->bhts', gg, vv)
        ds = p * (dp - D[..., None])
        dq = torch.einsum('bhts,bshd->bthd', ds, kk) * scale
        dk = torch.einsum('bhts,bthd->bshd', ds, qq) * scale
        dv = torch.einsum('bhts,bthd->bshd', p, gg)
        h, d = q.shape[1:]
        dq_state.reshape(b, 2 * c, h, d)[:, q_off * c:(q_off + q_len) * c] += dq
        junk = float('nan')                                                      # rows the kernels leave unwritten
        dkv = torch.full((2, b, 2 * c, h, d), junk)
        dkv[0, :, :k_len * c], dkv[1, :, :k_len * c] = dk, dv
        return dkv.reshape(2, b * 2 * c, h, d).to(q.dtype)

    def dq_end(self, dq_state, q, b, c, scale):
        return dq_state.to(q.dtype)

    def merge(self, o_acc, lse_acc, o_blk, lse_blk, b, c, plan, first):
        q_off, q_len = plan[0], plan[1]
        h, d = o_acc.shape[1:]
        oa = o_acc.reshape(b, 2 * c, h, d)[:, q_off * c:(q_off + q_len) * c]
        ob = self._parts(b, c, plan, o_acc, (o_blk, True))[0]
        la = lse_acc[:, :, q_off * c:(q_off + q_len) * c]
        if first:
            oa.copy_(ob)
            la.copy_(lse_blk)
            return
        new = torch.logaddexp(la, lse_blk)
        wa = torch.exp(la - new).permute(0, 2, 1)[..., None]
        wb = torch.exp(lse_blk - new).permute(0, 2, 1)[..., None]
        oa.copy_(oa * wa + ob * wb)
        la.copy_(new)

    def dkv_end(self, dkv, dtype):
        return dkv[0].to(dtype), dkv[1].to(dtype)

    def accumulate_dkv(self, acc, dkv_blk, b, c, k_len):
        h, d = acc.shape[2:]
        a5, s5 = acc.reshape(2, b, 2 * c, h, d), dkv_blk.reshape(2, b, 2 * c, h, d).float()
        a5[:, :, :k_len * c] += s5[:, :, :k_len * c]


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


def test_step_plan_covers_the_causal_triangle_exactly_once():
    from flash_attn.ring_attention import _step_plan
    for world in (1, 2, 3, 4, 8):
        seen = {}
        for rank in range(world):
            chunk = {0: rank, 1: 2 * world - 1 - rank}
            for step in range(world):
                src = (rank - step) % world
                kchunk = {0: src, 1: 2 * world - 1 - src}
                q_off, q_len, k_len, causal = _step_plan(step, rank, world)
                assert q_len * k_len == 2 or (step == 0 and causal)      # two c x c blocks of work per step and rank
                for a in range(q_off, q_off + q_len):
                    for bb in range(k_len):
                        key = (chunk[a], kchunk[bb])
                        if causal and key[1] > key[0]:
                            continue                                      # above the diagonal: masked inside the launch
                        assert key not in seen
                        seen[key] = causal
                        assert causal or key[1] < key[0]                 # full products lie strictly below the diagonal
        want = {(i, j) for i in range(2 * world) for j in range(2 * world) if j <= i}
        assert set(seen) == want
