"""Model check of the ordered-dQ schedule (FA_FLAG_DETERMINISTIC; csrc/fa_bwd_sm100.cu, drain warps and CTA decode).

In that mode the key-block CTAs of a sequence add their part of dQ to a query tile one after the other: a counter per tile
holds how many are done, a CTA spins until the counter equals its RANK on that tile, adds, and publishes rank + 1.  Two things
have to hold for every shape, and neither shows up in a data test until it deadlocks a GPU:

  1. on every tile the ranks of the contributing CTAs are 0 .. n-1, each once (else a turn never comes, or two CTAs share one);
  2. a CTA only ever waits for CTAs with a LOWER blockIdx: the hardware hands blocks to SMs in blockIdx order, so those are
     running or finished however few SMs there are -- no deadlock (a spinning CTA keeps its SM).

This file restates the kernel's arithmetic (rank, tile walk, launch order -- the lines are cited) and checks both properties by
brute force over ragged shapes, then runs a scheduler with 1 .. 148 SMs.  Mutation checks: the orders that were NOT shipped
(ranks counting down from the diagonal with the longest block launched first; a rotated non-causal walk when not all blocks
are resident) must deadlock here -- the harness sees what it is there to see."""
import random

import pytest

BLK = 128


def decode_cta(lin, nx, n_bh, bh_group):
    """fa_common.cuh decode_cta: x-major inside groups of (batch, head) pairs"""
    per_group = bh_group * nx
    group, r = divmod(lin, per_group)
    gsize = min(bh_group, n_bh - group * bh_group)
    x = r // gsize
    return x, group * bh_group + (r - x * gsize)


def ctas(causal, n_kblocks, n_bh, bh_group, shortest_first=True):
    """launch order -> (bh, key block).  causal, ordered mode: n_blk = n_kblocks - 1 - x (shortest block first)."""
    out = []
    for lin in range(n_kblocks * n_bh):
        if causal:
            x, bh = decode_cta(lin, n_kblocks, n_bh, bh_group)
            out.append((bh, n_kblocks - 1 - x if shortest_first else x))
        else:
            out.append(divmod(lin, n_kblocks))       # block-major inside each pair
    return out


def visits(causal, j, seqlen_q, seqlen_k, countdown=True):
    """(tile, rank) in visiting order for key block j of one sequence; [] when the CTA exits at once"""
    nq, nk = -(-seqlen_q // BLK), -(-seqlen_k // BLK)
    if j * BLK >= seqlen_k:
        return []                                     # k0 >= seqlen_k: return
    if causal:                                        # m_start = n_blk; rank = min(m, last block) - n_blk
        return [(m, (min(m, nk - 1) - j) if countdown else j) for m in range(j, nq)]
    return [(m, j) for m in range(nq)]                # tiles in order, rank = n_blk


def run(causal, seqs, n_sm, bh_group=2, seed=0, shortest_first=True, countdown=True, plans=None):
    """blocks are handed out in launch order to n_sm SMs; a resident CTA performs its next visit when the tile's counter equals
    its rank.  Raises on deadlock; returns the per-tile rank lists otherwise."""
    rng = random.Random(seed)
    n_kblocks = max(-(-sk // BLK) for _, sk in seqs)
    order = ctas(causal, n_kblocks, len(seqs), bh_group, shortest_first)
    plan = plans or {(bh, j): visits(causal, j, *seqs[bh], countdown=countdown) for bh, j in order}
    counter, seen = {}, {}
    pending, resident = list(order), {}
    while pending or resident:
        while pending and len(resident) < n_sm:
            c = pending.pop(0)
            if plan[c]:
                resident[c] = 0                       # CTAs without work exit immediately
        if not resident:
            continue
        ready = [c for c, i in resident.items() if counter.get((c[0], plan[c][i][0]), 0) == plan[c][i][1]]
        if not ready:
            raise RuntimeError(f"deadlock with {n_sm} SMs: resident {sorted(resident)[:6]} ...")
        c = rng.choice(ready)
        m, r = plan[c][resident[c]]
        counter[(c[0], m)] = r + 1
        seen.setdefault((c[0], m), []).append(r)
        resident[c] += 1
        if resident[c] == len(plan[c]):
            del resident[c]
    return seen


SHAPES = [[(4096, 4096)] * 3, [(1, 1)], [(129, 128), (128, 129)], [(700, 2048), (130, 1500), (2048, 96)],
          [(2048, 700), (1500, 130), (96, 2048), (257, 257)], [(64, 5000), (5000, 64)]]


@pytest.mark.parametrize("causal", [False, True])
@pytest.mark.parametrize("seqs", SHAPES)
def test_every_tile_gets_each_rank_once_and_nobody_waits_forever(causal, seqs):
    for n_sm in (1, 2, 7, 148):
        for seed in range(3):
            seen = run(causal, seqs, n_sm, seed=seed)
            for (bh, m), ranks in seen.items():
                sq, sk = seqs[bh]
                n = min(m, -(-sk // BLK) - 1) + 1 if causal else -(-sk // BLK)
                assert ranks == list(range(n)), (bh, m, ranks)            # in turn order, each once
            assert {m for (bh, m) in seen if bh == 0} == (set(range(-(-seqs[0][0] // BLK))) if seqs[0][1] > 0 else set())


def test_countdown_ranks_need_the_shortest_block_first():
    """mutation: the same ranks with the default (longest first) launch order wait for CTAs that are not resident yet"""
    seqs = [(4096, 4096)] * 4
    run(True, seqs, 3, shortest_first=True)
    with pytest.raises(RuntimeError, match="deadlock"):
        run(True, seqs, 3, shortest_first=False)
    run(True, seqs, 3, shortest_first=False, countdown=False)            # block j after j - 1 fits the longest-first order


def test_rotated_non_causal_walk_needs_every_block_resident():
    """the schedule that was built, measured (no gain) and removed: block j starts at tile j mod n and walks round; its ranks
    are a permutation too, but block j waits for block j + 1 -- fine with all blocks of a sequence resident, a deadlock otherwise"""
    nq = nk = 8
    plans = {}
    for bh in range(3):
        for j in range(nk):
            plans[(bh, j)] = [((j + i) % nq, i) for i in range(nq)]      # lockstep: the i-th visit is the i-th turn
    seqs = [(nq * BLK, nk * BLK)] * 3
    run(False, seqs, 8, plans=plans)
    run(False, seqs, 20, plans=plans)
    with pytest.raises(RuntimeError, match="deadlock"):
        run(False, seqs, 7, plans=plans)
