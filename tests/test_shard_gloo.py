"""N>1 host logic on CPU: contiguous shards (specs and history segments), and the rank protocol bench.py uses
(barrier + max-over-ranks) under torch.distributed with the gloo backend, world_size 2."""
import os
import socket

import numpy as np
import pytest

from cron_operator_b200 import shard, synth


def test_shard_ranges_cover_and_align():
    for n in (0, 1, 2, 7, 1000, 1_000_001):
        for world in (1, 2, 3, 4, 8):
            spans = [shard.shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c and a <= b
            for r, (a, b) in enumerate(spans):
                assert a % 2 == 0 and (b % 2 == 0 or r == world - 1)


def test_history_cuts_balance_on_segment_boundaries():
    rng = np.random.default_rng(0)
    lens = rng.integers(0, 17, 10_000)
    seg = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    for world in (1, 2, 4, 8):
        cuts = shard.history_cuts(seg, world)
        assert cuts[0] == 0 and cuts[-1] == len(lens) and np.all(np.diff(cuts) >= 0)
        per = [int(seg[cuts[g + 1]] - seg[cuts[g]]) for g in range(world)]
        assert max(per) - min(per) <= 32 and sum(per) == int(seg[-1])
    assert shard.history_cuts(np.array([0, 0, 0], np.uint64), 2).tolist() == [0, 0, 2]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, n, q):
    import torch.distributed as dist
    from oracle import pyoracle as po
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = shard.shard_range(n, rank, world)
        # a shard generated on its own equals the same slice of the whole synthetic spec list (counter-based RNG)
        masks = synth.gen_specs(hi - lo, seed=5, start=lo)
        frm = synth.gen_times(hi - lo, synth.ts(2024), synth.ts(2034), seed=5, start=lo)
        out = po.batch_next_n(masks, frm, 2, threads=1)  # CPU stand-in for the per-rank kernel call
        dist.barrier()
        t = shard.max_over_ranks([1.0 + rank, 5.0 - rank])
        q.put((rank, lo, hi, masks, frm, out, t))
    finally:
        dist.destroy_process_group()


def test_two_ranks_gloo_shards_equal_whole():
    import torch.multiprocessing as mp
    from oracle import pyoracle as po
    n, world = 20_001, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted([q.get(timeout=240) for _ in range(world)], key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    whole_m = synth.gen_specs(n, seed=5)
    whole_f = synth.gen_times(n, synth.ts(2024), synth.ts(2034), seed=5)
    whole = po.batch_next_n(whole_m, whole_f, 2, threads=2)
    assert np.array_equal(np.concatenate([g[3] for g in got], axis=1), whole_m)
    assert np.array_equal(np.concatenate([g[4] for g in got]), whole_f)
    assert np.array_equal(np.concatenate([g[5] for g in got]), whole)
    assert got[0][1] == 0 and got[0][2] == got[1][1] and got[1][2] == n
    for g in got:
        assert g[6] == [2.0, 5.0]  # max over ranks, identical on every rank
