"""The library's own multi-device context (b200cron_ctx_create with several devices) on a box with >= 2 GPUs: specs
shard contiguously over the devices, history is cut on segment boundaries so that no Cron's records straddle two
devices, and nothing is exchanged between devices.  Every result must equal the single-device
answer and the oracle's.  Skipped on a one-GPU box; run under `gpurun --gpus 2` (log kept in profiles/)."""
import numpy as np
import pytest

from cron_operator_b200 import api, synth
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
ts = synth.ts
KEYS = ("last", "next", "missed", "status", "decision")


def _gpus():
    import torch
    return torch.cuda.device_count()


@pytest.fixture(scope="module")
def ctxs():
    g = _gpus()
    if g < 2:
        pytest.skip("one GPU visible")
    one, many = api.Context(devices=[0]), api.Context(devices=list(range(g)))
    assert many.device_count == g
    yield one, many
    one.close()
    many.close()


def _dst(ctx, names):
    return np.array([z for z in range(1, len(names)) if (ctx.zone_periods(z)[0] > ts(2030)).any()])


def test_next_n_shards_match_single_device_and_oracle(ctxs):
    one, many = ctxs
    names = one.zone_names
    for n in (1, 3, 300_001, 1_000_000):
        masks = synth.gen_specs(n, seed=8, with_seconds=True)
        zid = synth.gen_zone_ids(n, len(names), _dst(one, names), seed=8)
        frm = synth.gen_times(n, ts(2024), ts(2036), seed=8)
        for k in (1, 4):
            got = many.next_n(masks, frm, k, zid)
            assert np.array_equal(got, one.next_n(masks, frm, k, zid)), (n, k)
        idx = np.arange(0, n, 97)
        want = po.batch_next_n(np.ascontiguousarray(masks[:, idx]), frm[idx], 4, zid[idx], names)
        assert np.array_equal(got[idx], want), n
    # pinned buffers: every device works on its shard of the host arrays in place
    n = 400_002
    masks = synth.gen_specs(n, seed=9)
    frm = synth.gen_times(n, ts(2024), ts(2034), seed=9)
    P = api.PinnedArray
    hm, hf, ho = P((6, n), np.uint64), P(n, np.int64), P((n, 1), np.int64)
    hm.array[:] = masks; hf.array[:] = frm
    many.next_n(hm.array, hf.array, 1, out=ho.array, five_field=True)
    assert np.array_equal(ho.array, po.batch_next_n(masks, frm, 1))


def test_unmet_since_and_spec_set_over_devices(ctxs):
    one, many = ctxs
    names = one.zone_names
    n = 500_001
    masks = synth.gen_specs(n, seed=41, dense_minutes=0.5)
    zid = synth.gen_zone_ids(n, len(names), _dst(one, names), seed=41)
    now = ts(2026, 3, 29, 12)
    earliest = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=41)
    rng = np.random.default_rng(41)
    policy, active = rng.integers(0, 3, n).astype(np.uint8), rng.integers(0, 3, n).astype(np.uint32)
    flags = (rng.random(n) < 0.02).astype(np.uint8)
    got = many.unmet_since(masks, earliest, now, 0, policy, active, flags, None, zid)
    ref = one.unmet_since(masks, earliest, now, 0, policy, active, flags, None, zid)
    for k in KEYS:
        assert np.array_equal(got[k], ref[k]), k
    idx = np.arange(0, n, 499)
    want = po.batch_unmet(np.ascontiguousarray(masks[:, idx]), earliest[idx], now, 0, policy[idx], active[idx], flags[idx],
                          None, zid[idx], names)
    for k in KEYS:
        assert np.array_equal(got[k][idx], want[k]), k
    st = many.specset(masks, zid)
    try:
        got2 = st.unmet_since(earliest, now, 0, policy, active, flags)
        for k in KEYS:
            assert np.array_equal(got2[k], ref[k]), k
        frm = synth.gen_times(n, ts(2024), ts(2034), seed=5)
        assert np.array_equal(st.next_n(frm, 2), one.next_n(masks, frm, 2, zid))
        new_m, new_z = synth.gen_specs(4001, seed=77), np.zeros(4001, np.int32)
        lo = n // 2 - 2000                       # straddles the boundary between two devices' shards
        st.update(lo, new_m, new_z)
        m2, z2 = masks.copy(), zid.copy()
        m2[:, lo:lo + 4001] = new_m
        z2[lo:lo + 4001] = 0
        assert np.array_equal(st.next_n(frm, 1), one.next_n(m2, frm, 1, z2))
    finally:
        st.close()


def test_history_cut_on_segment_boundaries(ctxs):
    one, many = ctxs
    rng = np.random.default_rng(3)
    # skewed lengths: the record-balanced cut lands far from the middle segment, and long segments sit at the cuts
    lens = np.concatenate([rng.integers(0, 4, 200_000), rng.integers(0, 60, 30_000), [5000, 0, 0, 3000], rng.integers(0, 17, 100_000)])
    seg_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    key = (1_700_000_000 - rng.integers(0, 400, int(seg_off[-1]))).astype(np.int64)
    limit = rng.integers(-1, 20, len(lens)).astype(np.int64)
    perm, keep = many.history_prune(seg_off, key, limit)
    p1, k1 = one.history_prune(seg_off, key, limit)
    assert np.array_equal(perm, p1) and np.array_equal(keep, k1)
    wperm, wkeep = po.batch_history(seg_off, key, limit)
    assert np.array_equal(perm, wperm) and np.array_equal(keep, wkeep)
    # degenerate shapes: fewer segments than devices, all records in one segment
    for so in ([0, 7], [0, 0, 9], [0, 3, 3, 3]):
        so = np.array(so, np.uint64)
        kk = rng.integers(0, 5, int(so[-1])).astype(np.int64)
        p, k = many.history_prune(so, kk, None)
        wp, wk = po.batch_history(so, kk, None)
        assert np.array_equal(p, wp) and np.array_equal(k, wk)


def test_C5_shape_through_one_multi_device_context(ctxs):
    """BASELINE config 5's shape (UTC specs, one fire time, plus per-Cron history with historyLimit 10) at a size whose
    host arrays fit comfortably: 16M specs and 2M history segments through ONE context that shards them itself"""
    one, many = ctxs
    n = 16_000_000
    P = api.PinnedArray
    hm, hf, ho = P((6, n), np.uint64), P(n, np.int64), P((n, 1), np.int64)
    step = 2_000_000
    for a in range(0, n, step):
        hm.array[:, a:a + step] = synth.gen_specs(step, start=a)
        hf.array[a:a + step] = synth.gen_times(step, ts(2024), ts(2034), start=a)
    many.next_n(hm.array, hf.array, 1, out=ho.array, five_field=True)
    idx = np.arange(0, n, 9_973)
    want = po.batch_next_n(np.ascontiguousarray(hm.array[:, idx]), hf.array[idx], 1)
    assert np.array_equal(ho.array[idx], want)
    g = many.device_count
    for d in range(1, g):                       # the specs either side of every shard boundary
        b = (n * d // g) & ~1
        sl = slice(b - 64, b + 64)
        assert np.array_equal(ho.array[sl], po.batch_next_n(np.ascontiguousarray(hm.array[:, sl]), hf.array[sl], 1))
    rng = np.random.default_rng(5)
    ns = 2_000_000
    lens = rng.integers(0, 17, ns)
    seg_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    key = (1_760_000_000 - rng.integers(0, 86_400, int(seg_off[-1]))).astype(np.int64)
    limit = np.full(ns, 10, np.int64)
    perm, keep = many.history_prune(seg_off, key, limit)
    wperm, wkeep = po.batch_history(seg_off, key, limit)
    assert np.array_equal(perm, wperm) and np.array_equal(keep, wkeep)
