"""GPU parity tests proper: every call goes through the C ABI of libb200cron.so (cron_operator_b200.api) and is
compared bit-exactly with the CPU oracle on the same seeded inputs, with the committed golden vectors, and -- at
full BASELINE sizes -- through size-independent properties."""
import json
import os

import numpy as np
import pytest

from cron_operator_b200 import api, synth
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
ts = synth.ts
V = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "vectors.json")))
KEYS = ("last", "next", "missed", "status", "decision")


@pytest.fixture(scope="module")
def ctx():
    c = api.Context()
    yield c
    c.close()


@pytest.fixture(scope="module")
def names(ctx):
    return ctx.zone_names


def _masks_of(texts, with_seconds=False):
    specs = [api._parse(t, with_seconds) for t in texts]
    return api.pack_specs(specs)


def _dst_zones(ctx, names):
    return np.array([z for z in range(1, len(names)) if (ctx.zone_periods(z)[0] > ts(2030)).any()])


# ---------------------------------------------------------------------------------------------------------------
def test_golden_vectors(ctx, names):
    for c in V["next"]:
        m = np.array([[int(x)] for x in c["masks"]], dtype=np.uint64)
        z = np.array([names.index(c["zone"])], np.int32)
        for f, want in zip(c["from"], c["want"]):
            assert ctx.next_n(m, [f], 4, z)[0].tolist() == want, (c["spec"], c["zone"], f)
    for c in V["unmet"]:
        m = np.array([[int(x)] for x in c["masks"]], dtype=np.uint64)
        z = np.array([names.index(c["zone"])], np.int32)
        dl = [po.ZERO if c["deadline"] is None else c["deadline"]]
        got = ctx.unmet_since(m, [c["earliest"]], c["now"], 0, [c["policy"]], [c["active"]], [c["flags"]], dl, z)
        for k, v in c["want"].items():
            assert int(got[k][0]) == v, (c["spec"], k)
    for c in V["history"]:
        perm, keep = ctx.history_prune([0, len(c["keys"])], np.array(c["keys"], dtype=np.int64),
                                       [-1 if c["limit"] is None else c["limit"]])
        assert perm.tolist() == c["perm"] and keep.tolist() == c["keep"]


def test_upstream_robfig_vectors(ctx, names):
    """robfig/cron v3.0.1's own TestNext / TestActivation rows ([UPSTREAM-RECALLED], tests/golden/upstream_robfig.json)
    through b200cron_parse_* and b200cron_next_n on the GPU: expected values from outside this repository"""
    U = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "upstream_robfig.json")))
    specs = [api._parse(c["spec"], True) for c in U["next"]]
    zid = np.array([names.index(s.Location or c["zone"]) for s, c in zip(specs, U["next"])], np.int32)
    frm = np.array([c["from"] for c in U["next"]], np.int64)
    want = np.array([po.ZERO if c["want"] is None else c["want"] for c in U["next"]], np.int64)
    got = ctx.next_n(api.pack_specs(specs), frm, 1, zid)[:, 0]
    bad = np.nonzero(got != want)[0]
    assert len(bad) == 0, [(U["next"][b]["time"], U["next"][b]["spec"], int(got[b]), int(want[b])) for b in bad[:5]]
    specs = [api._parse(c["spec"], False) for c in U["activation"]]
    at = np.array([c["at"] for c in U["activation"]], np.int64)
    got = ctx.next_n(api.pack_specs(specs), at - 1, 1)[:, 0]
    assert ((got == at) == np.array([c["expected"] for c in U["activation"]])).all()


def test_odd_transition_known_answers_and_ordinary_dst_scan(ctx, names):
    """the hand-traced upstream behaviours at odd transitions (tests/test_oracle_independent_dst.py QUIRKS) and a sample
    of the algorithm-free ordinary-DST scan, through b200cron_next_n on the GPU"""
    import datetime as dt
    import random
    import zoneinfo
    from tests import test_oracle_independent_dst as ind
    rows = [q for q in ind.QUIRKS if q[0] in names]
    assert len(rows) >= 4
    specs = [api._parse(q[1], False) for q in rows]
    frm = np.array([ind._local(q[0], *q[2]) for q in rows], np.int64)
    want = np.array([int(dt.datetime(*q[3], tzinfo=zoneinfo.ZoneInfo(q[0]), fold=1).timestamp()) for q in rows], np.int64)
    zid = np.array([names.index(q[0]) for q in rows], np.int32)
    got = ctx.next_n(api.pack_specs(specs), frm, 1, zid)[:, 0]
    assert np.array_equal(got, want), (got.tolist(), want.tolist())
    # ordinary whole-hour transitions: Next() == earliest matching minute found by scanning with python's zoneinfo
    rng = random.Random(77)
    cands = [(names.index(ind.NAMES[zi]), tr) for zi, tr in ind._ordinary_transitions() if ind.NAMES[zi] in names]
    cases = [(*rng.choice(cands), rng.choice(ind.SPECS)) for _ in range(600)]
    specs = [api._parse(c[2], False) for c in cases]
    frm = np.array([c[1] + rng.randrange(-2 * 86400, 86400) for c in cases], np.int64)
    zid = np.array([c[0] for c in cases], np.int32)
    got = ctx.next_n(api.pack_specs(specs), frm, 1, zid)[:, 0]
    checked = 0
    for i, c in enumerate(cases):
        try:
            z = zoneinfo.ZoneInfo(names[c[0]])
        except Exception:
            continue
        m, r, t = specs[i].masks(), int(got[i]), int(frm[i])
        assert r != po.ZERO and r > t
        assert ind._first_match_by_scan(m, z, t, min(r, t + 10 * 86400)) is None, (names[c[0]], c[2], t, r)
        if r <= t + 10 * 86400:
            assert ind._matches(m, dt.datetime.fromtimestamp(r, z)), (names[c[0]], c[2], t, r)
        checked += 1
    assert checked > 500


def test_reference_shapes(ctx):
    """the reference's own known-answer tests, written against the mirrored interface"""
    now = ts(2026, 1, 1, 12)
    with pytest.raises(ValueError, match="unparsable cron"):      # K1 cron_controller_test.go:170-186
        ctx.getNextSchedule("60 31 30 2 *", now - 300, now)
    with pytest.raises(ValueError, match="unschedulable cron"):   # K2 :188-204
        ctx.getNextSchedule("0 0 30 2 *", now - 300, now)
    last, nxt = ctx.getNextSchedule("*/1 * * * *", now - 300, now)  # K3 :206-222
    assert last == now and nxt == now + 60
    assert ctx.sortByCreationTimestamp([now - 600, now - 1200, now - 300]).tolist() == [1, 0, 2]  # K4
    s = api.ParseStandard("CRON_TZ=Asia/Tokyo 0 9 * * *")
    assert ctx.Next(s, ts(2025, 6, 1)) == ts(2025, 6, 2, 0)
    assert ctx.Next(api.ParseStandard("@every 1h30m"), 1000) == 1000 + 5400
    deleted, kept = ctx.syncCronHistory([5, 1, 4, 2, 3], 3)
    assert deleted.tolist() == [1, 3] and kept.tolist() == [4, 2, 0]


def test_C2_full_parity(ctx):
    """BASELINE config 2: 1M random 5-field UTC specs, next single fire time -- 100% of indices vs the oracle"""
    n = 1_000_000
    masks = synth.gen_specs(n)
    frm = synth.gen_times(n, ts(2024), ts(2034))
    got = ctx.next_n(masks, frm, 1)
    want = po.batch_next_n(masks, frm, 1)
    bad = np.nonzero(got != want)[0]
    assert len(bad) == 0, (bad[:5], got[bad[:5]], want[bad[:5]])
    assert 500 < (want == po.ZERO).sum() < 2000


def test_C3_shape_parity(ctx, names):
    """config 3 shape: seconds field, mixed IANA zones incl. DST, next 16 fire times"""
    n = 400_000
    masks = synth.gen_specs(n, seed=77, with_seconds=True)
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=77)
    frm = synth.gen_times(n, ts(2024), ts(2036), seed=77)
    # 1%: re-draw to within +-2h of a transition of that spec's zone
    rng = np.random.default_rng(77)
    for i in rng.choice(n, n // 100, replace=False):
        when = ctx.zone_periods(int(zid[i]))[0]
        w = when[(when > ts(2024)) & (when < ts(2036))]
        if len(w):
            frm[i] = int(rng.choice(w)) + int(rng.integers(-7200, 7200))
    got = ctx.next_n(masks, frm, 16, zid)
    want = po.batch_next_n(masks, frm, 16, zid, names)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert len(bad) == 0, [(names[zid[b]], int(frm[b]), got[b][:3], want[b][:3]) for b in bad[:3]]


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 8, 16, 32, 33, 64])
def test_every_k_and_ragged_n(ctx, names, k):
    for n in (0, 1, 2, 63, 64, 65, 255, 257, 1001):
        masks = synth.gen_specs(n, seed=k, with_seconds=True)
        zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=k)
        frm = synth.gen_times(n, ts(2024), ts(2036), seed=k)
        got = ctx.next_n(masks, frm, k, zid)
        want = po.batch_next_n(masks, frm, k, zid, names)
        assert np.array_equal(got, want), (k, n)


def test_bad_arguments(ctx):
    m = synth.gen_specs(4)
    with pytest.raises(api.B200CronError):
        ctx.next_n(m, np.zeros(4, np.int64), 0)
    with pytest.raises(api.B200CronError):
        ctx.next_n(m, np.zeros(4, np.int64), 65)
    with pytest.raises(api.B200CronError) as e:
        ctx.next_n(m, np.array([0, 0, 0, 2**62], np.int64), 1)
    assert e.value.code == api.ERANGE
    with pytest.raises(api.B200CronError) as e:
        ctx.next_n(m, np.zeros(4, np.int64), 1, np.array([0, 1, 2, 100000], np.int32))
    assert e.value.code == api.ERANGE
    z = ctx.zone_id("Europe/Berlin")
    with pytest.raises(api.B200CronError) as e:  # beyond the expanded-footer horizon for a DST zone
        ctx.next_n(m, np.full(4, ts(2197), np.int64), 1, np.full(4, z, np.int32))
    assert e.value.code == api.ERANGE
    assert ctx.next_n(m, np.full(4, ts(2197), np.int64), 1).shape == (4, 1)  # fine in UTC
    with pytest.raises(api.ParseError):
        ctx.zone_id("Nope/Nowhere")


def test_transitions_through_abi(ctx, names):
    texts = ["* * * * *", "*/5 * * * *", "30 2 * * *", "0 0 * * *", "0 1 * * *", "59 23 * * *", "0 0 1 * *",
             "15 3 * * 0", "0 */2 * * *", "30 0 * * *", "45 1-3 25-31 * 0", "0 3 * * *", "*/30 * * * *"]
    base = _masks_of(texts)
    for name in ["America/New_York", "Europe/London", "Australia/Lord_Howe", "Africa/Cairo", "America/Havana",
                 "America/Santiago", "Atlantic/Azores", "Asia/Tehran", "Pacific/Apia", "Asia/Gaza"]:
        z = names.index(name)
        when = ctx.zone_periods(z)[0]
        trans = [int(w) for w in when if ts(2015) < w < ts(2046)]
        offs = [-1, -61, -3601, -7201, -86401, -40 * 86400, 0, 1, 3600]
        starts = np.array([t + o for t in trans for o in offs], dtype=np.int64)
        S, T = len(texts), len(starts)
        masks = np.ascontiguousarray(np.repeat(base, T, axis=1))
        frm = np.tile(starts, S)
        zid = np.full(S * T, z, dtype=np.int32)
        got = ctx.next_n(masks, frm, 4, zid)
        want = po.batch_next_n(masks, frm, 4, zid, names)
        assert np.array_equal(got, want), name


def test_device_pointer_path_and_unaligned(ctx, names):
    import torch
    n = 100_003
    masks = synth.gen_specs(n + 1, seed=5, with_seconds=True)
    zid = synth.gen_zone_ids(n + 1, len(names), _dst_zones(ctx, names), seed=5)
    frm = synth.gen_times(n + 1, ts(2024), ts(2036), seed=5)
    dm = torch.from_numpy(masks.view(np.int64)).cuda()
    dz, df = torch.from_numpy(zid).cuda(), torch.from_numpy(frm).cuda()
    for shift in (0, 1):  # shift 1: 8-byte-aligned-only pointers -> scalar-load kernel
        for k in (1, 16):
            out = torch.empty((n, k), dtype=torch.int64, device="cuda")
            planes = [dm[p, shift:shift + n] for p in range(6)]
            ctx.next_n_dev(planes, df[shift:shift + n], out, k, dz[shift:shift + n],
                           stream=torch.cuda.current_stream().cuda_stream)
            ctx.sync_check(stream=torch.cuda.current_stream().cuda_stream)
            want = po.batch_next_n(np.ascontiguousarray(masks[:, shift:shift + n]), frm[shift:shift + n], k,
                                   zid[shift:shift + n], names)
            assert np.array_equal(out.cpu().numpy(), want), (shift, k)


def _cmp(got, want, ctx_=""):
    for k in KEYS:
        bad = np.nonzero(got[k] != want[k])[0]
        assert len(bad) == 0, (ctx_, k, bad[:5], got[k][bad[:5]], want[k][bad[:5]])


def test_unmet_C4_shape(ctx, names):
    n = 60_000
    masks = synth.gen_specs(n, seed=41, dense_minutes=0.5)
    dense = np.array([bin(int(x) & ((1 << 60) - 1)).count("1") for x in masks[1]]) > 6
    masks[2] = np.where(dense, masks[2] & np.uint64(0x800000000000000E), masks[2])  # keeps the literal oracle affordable
    masks[2] = np.where(masks[2] & np.uint64(0xFFFFFF), masks[2], np.uint64(1 << 2))
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=41)
    now = ts(2026, 3, 29, 12)
    earliest = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=41)
    rng = np.random.default_rng(41)
    policy = rng.integers(0, 3, n).astype(np.uint8)
    active = rng.choice([0, 1, 2, 3, 4], size=n, p=[0.6, 0.3, 0.04, 0.03, 0.03]).astype(np.uint32)
    flags = (rng.random(n) < 0.02).astype(np.uint8)
    deadline = np.where(rng.random(n) < 0.01, now - 100, po.ZERO).astype(np.int64)
    got = ctx.unmet_since(masks, earliest, now, 0, policy, active, flags, deadline, zid)
    want = po.batch_unmet(masks, earliest, now, 0, policy, active, flags, deadline, zid, names)
    _cmp(got, want)
    assert want["missed"].max() > 100
    # optional inputs NULL
    got = ctx.unmet_since(masks[:, :1000].copy(), earliest[:1000], now)
    want = po.batch_unmet(masks[:, :1000].copy(), earliest[:1000], now)
    _cmp(got, want)


def test_unmet_across_transitions(ctx, names):
    texts = ["*/20 * * * *", "0 * * * *", "30 2 * * *", "0 0 * * *", "0 1 * * *", "59 23 * * *", "0 0 1 * *",
             "15 3 * * 0", "0 */2 * * *", "30 0 * * *", "45 1-3 25-31 * 0", "0 3 * * *", "0 0 2 * *", "*/30 0-4 * * *"]
    base = _masks_of(texts)
    wins = [(-3 * 86400, 3 * 86400), (-40 * 86400, 1), (-5, 86400 * 2), (-86400, -1), (-7200, 7200), (0, 86400),
            (-20 * 86400, 20 * 86400), (-3600, 0), (1, 5 * 86400)]
    for name in ["America/New_York", "Europe/London", "Australia/Lord_Howe", "Africa/Cairo", "America/Havana",
                 "Atlantic/Azores", "Asia/Tehran", "America/Santiago"]:
        z = names.index(name)
        when = ctx.zone_periods(z)[0]
        for t in [int(w) for w in when if ts(2021) < w < ts(2031)][:10]:
            for a, b in wins:
                S = len(texts)
                zz = np.full(S, z, np.int32)
                got = ctx.unmet_since(base, np.full(S, t + a), t + b, zone_id=zz)
                want = po.batch_unmet(base, np.full(S, t + a), t + b, zone_id=zz, zone_names=names)
                _cmp(got, want, (name, t, a, b))


def test_history_parity(ctx):
    rng = np.random.default_rng(9)
    for nseg, maxlen in [(1, 3), (1000, 16), (100_000, 16), (257, 40), (3, 5000), (600, 30)]:
        lens = rng.integers(0, maxlen + 1, nseg)
        seg_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        key = (1_700_000_000 - rng.integers(0, 400, int(seg_off[-1]))).astype(np.int64)  # many ties
        for limit in (None, np.full(nseg, 10, np.int64), rng.integers(-1, 20, nseg).astype(np.int64)):
            perm, keep = ctx.history_prune(seg_off, key, limit)
            wperm, wkeep = po.batch_history(seg_off, key, limit)
            assert np.array_equal(perm, wperm) and np.array_equal(keep, wkeep), (nseg, maxlen)
    perm, keep = ctx.history_prune(np.array([0, 0, 0], np.uint64), np.zeros(0, np.int64), np.array([1, 1], np.int64))
    assert len(perm) == 0 and len(keep) == 0


def test_properties_at_full_size(ctx, names):
    """size-independent properties on a C3-sized slice (device-resident, no oracle): rows strictly increasing,
    first > from, chaining (NextN(k) == Next applied k times), determinism across launches"""
    import torch
    n, k = 4_000_000, 16
    masks = synth.gen_specs(n, seed=3, with_seconds=True)
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=3)
    frm = synth.gen_times(n, ts(2024), ts(2036), seed=3)
    dm = torch.from_numpy(masks.view(np.int64)).cuda()
    dz, df = torch.from_numpy(zid).cuda(), torch.from_numpy(frm).cuda()
    st = torch.cuda.current_stream().cuda_stream
    out = torch.empty((n, k), dtype=torch.int64, device="cuda")
    ctx.next_n_dev(dm, df, out, k, dz, stream=st)
    ctx.sync_check(stream=st)
    zero = out == po.ZERO
    assert bool(((out[:, 1:] > out[:, :-1]) | zero[:, 1:]).all())       # strictly increasing until the zero time
    assert bool((zero[:, 1:] | ~zero[:, :-1]).all())                    # once zero, always zero
    assert bool(((out[:, 0] > df) | zero[:, 0]).all())
    one = torch.empty((n, 1), dtype=torch.int64, device="cuda")
    cur = df.clone()
    for j in range(3):                                                   # chaining: column j == Next(column j-1)
        ok = cur != po.ZERO
        ctx.next_n_dev(dm, torch.where(ok, cur, torch.zeros_like(cur)), one, 1, dz, stream=st)
        ctx.sync_check(stream=st)
        cur = torch.where(ok, one[:, 0], cur)
        assert bool((cur == out[:, j]).all()), j
    out2 = torch.empty_like(out)
    ctx.next_n_dev(dm, df, out2, k, dz, stream=st)
    ctx.sync_check(stream=st)
    assert bool((out2 == out).all())
    # sampled oracle parity on every 997th index
    idx = np.arange(0, n, 997)
    want = po.batch_next_n(np.ascontiguousarray(masks[:, idx]), frm[idx], k, zid[idx], names)
    assert np.array_equal(out[torch.from_numpy(idx).cuda()].cpu().numpy(), want)


def test_multi_device_context_matches_single(ctx, names):
    import torch
    g = torch.cuda.device_count()
    if g < 2:
        pytest.skip("one GPU visible")
    n = 300_001
    masks = synth.gen_specs(n, seed=8, with_seconds=True)
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=8)
    frm = synth.gen_times(n, ts(2024), ts(2036), seed=8)
    c2 = api.Context(devices=list(range(g)))
    try:
        assert np.array_equal(c2.next_n(masks, frm, 4, zid), ctx.next_n(masks, frm, 4, zid))
    finally:
        c2.close()


def test_unmet_list_parity(ctx, names):
    """capped list of the missed fire times: prefix-sum offsets + emit kernel vs the oracle's literal chain"""
    import torch
    for n in (1, 5, 1023, 1024, 1025, 50_000):
        masks = synth.gen_specs(n, seed=52, dense_minutes=0.4)
        zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=52)
        now = ts(2026, 3, 29, 12)
        earliest = now - synth.gen_times(n, 0, 2 * 86400, seed=52)
        rng = np.random.default_rng(n)
        flags = (rng.random(n) < 0.05).astype(np.uint8)
        u = ctx.unmet_since(masks, earliest, now, flags=flags, zone_id=zid)
        for cap in (1, 100):
            off, lst = ctx.unmet_list(masks, earliest, u["missed"], u["decision"], cap, zid)
            woff, wlst = po.batch_unmet_list(masks, earliest, u["missed"], u["decision"], cap, zid, names)
            assert np.array_equal(off, woff), (n, cap)
            assert np.array_equal(lst, wlst), (n, cap)
            # every listed run is <= now, ascending, and the last listed run of an uncapped Cron is lastMissed
            cnt = np.diff(off.astype(np.int64))
            full = (cnt > 0) & (u["missed"] <= cap)
            ends = off[1:][full].astype(np.int64) - 1
            assert np.array_equal(lst[ends], u["last"][full])
    # device-pointer variant, two steps
    n = 20_000
    masks = synth.gen_specs(n, seed=53, dense_minutes=0.4)
    now = ts(2026, 3, 29, 12)
    earliest = now - synth.gen_times(n, 0, 86400, seed=53)
    u = ctx.unmet_since(masks, earliest, now)
    st = torch.cuda.current_stream().cuda_stream
    dm, de = torch.from_numpy(masks.view(np.int64)).cuda(), torch.from_numpy(earliest).cuda()
    dmi, dd = torch.from_numpy(u["missed"].view(np.int32)).cuda(), torch.from_numpy(u["decision"]).cuda()
    doff = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    ctx.unmet_list_offsets_dev(dmi, dd, doff, 100, stream=st)
    ctx.sync_check(stream=st)
    total = int(doff[-1].item())
    dl = torch.empty(total, dtype=torch.int64, device="cuda")
    ctx.unmet_list_emit_dev(dm, de, doff, dl, stream=st)
    ctx.sync_check(stream=st)
    woff, wlst = po.batch_unmet_list(masks, earliest, u["missed"], u["decision"], 100)
    assert np.array_equal(doff.cpu().numpy().view(np.uint64), woff) and np.array_equal(dl.cpu().numpy(), wlst)
    # capacity too small -> ENOMEM with the needed size in list_off[n]
    with pytest.raises(api.B200CronError) as e:
        off = np.empty(n + 1, np.uint64)
        tiny = np.empty(1, np.int64)
        ctx._check(api.lib().b200cron_unmet_list(ctx._h, api._planes(masks)[0], None, earliest.ctypes.data,
                                                 u["missed"].ctypes.data, u["decision"].ctypes.data, n, 100,
                                                 off.ctypes.data, tiny.ctypes.data, 1))
    assert e.value.code == api.ENOMEM


def test_device_generator_matches_numpy(ctx, names):
    """the device-side synthetic workload generator is bit-identical to synth.py for any index window"""
    import torch
    st = torch.cuda.current_stream().cuda_stream
    dst = _dst_zones(ctx, names)
    pool = torch.from_numpy(synth.zone_pool(len(names), dst)).cuda()
    for (n, start, ws, dense, seed) in [(100_003, 0, False, 0.0, 0xC0FFEE), (65_537, 987_654_321, True, 0.0, 3),
                                         (50_000, 123, False, 0.5, 41)]:
        dm = torch.empty((6, n), dtype=torch.int64, device="cuda")
        df = torch.empty(n, dtype=torch.int64, device="cuda")
        dz = torch.empty(n, dtype=torch.int32, device="cuda")
        ctx.synth_specs_dev(dm, n, seed, start, ws, dense, 0.001, stream=st)
        ctx.synth_times_dev(df, n, ts(2024), ts(2036), seed, start, stream=st)
        ctx.synth_zones_dev(dz, n, pool, seed, start, stream=st)
        ctx.sync_check(stream=st)
        assert np.array_equal(dm.cpu().numpy().view(np.uint64), synth.gen_specs(n, seed, start, ws, dense)), (n, start)
        assert np.array_equal(df.cpu().numpy(), synth.gen_times(n, ts(2024), ts(2036), seed, start))
        assert np.array_equal(dz.cpu().numpy(), synth.gen_zone_ids(n, len(names), dst, seed, start))
    # index-array regeneration used for sampled parity at full size
    idx = np.array([5, 77, 100_002], dtype=np.uint64)
    assert np.array_equal(synth.gen_specs(0, idx=idx), synth.gen_specs(100_003)[:, idx.astype(np.int64)])


def test_C3_full_size_sampled_parity(ctx, names):
    """BASELINE config 3 at its full size: 100M specs with a seconds field and mixed IANA zones generated on the
    device, next 16 fire times each; oracle parity on every 9973rd index (regenerated on the host by index) plus the
    size-independent row properties over all 1.6e9 results."""
    import torch
    free, _ = torch.cuda.mem_get_info()
    n, k = 100_000_000, 16
    if free < 30 * (1 << 30):
        pytest.skip("needs ~20 GB of free device memory")
    st = torch.cuda.current_stream().cuda_stream
    dst = _dst_zones(ctx, names)
    pool = torch.from_numpy(synth.zone_pool(len(names), dst)).cuda()
    dm = torch.empty((6, n), dtype=torch.int64, device="cuda")
    df = torch.empty(n, dtype=torch.int64, device="cuda")
    dz = torch.empty(n, dtype=torch.int32, device="cuda")
    ctx.synth_specs_dev(dm, n, 3, 0, True, stream=st)
    ctx.synth_times_dev(df, n, ts(2024), ts(2036), 3, 0, stream=st)
    ctx.synth_zones_dev(dz, n, pool, 3, 0, stream=st)
    out = torch.empty((n, k), dtype=torch.int64, device="cuda")
    ctx.next_n_dev(dm, df, out, k, dz, stream=st)
    ctx.sync_check(stream=st)
    idx = np.arange(0, n, 9973, dtype=np.uint64)
    want = po.batch_next_n(synth.gen_specs(0, 3, with_seconds=True, idx=idx), synth.gen_times(0, ts(2024), ts(2036), 3, idx=idx),
                           k, synth.gen_zone_ids(0, len(names), dst, 3, idx=idx), names)
    got = out[torch.from_numpy(idx.astype(np.int64)).cuda()].cpu().numpy()
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert len(bad) == 0, (idx[bad][:5], got[bad[:2]], want[bad[:2]])
    zero = out == po.ZERO
    assert bool((zero[:, 1:] | ~zero[:, :-1]).all())        # once the zero time, always the zero time
    assert bool(((out[:, 0] > df) | zero[:, 0]).all())      # first fire is after `from`
    # rows increase except where upstream itself steps back across a DST overlap: rare
    back = ((out[:, 1:] <= out[:, :-1]) & ~zero[:, 1:]).sum().item()
    assert back < n * k * 1e-6, back


def test_device_parse_matches_host_parser(ctx):
    """batched device ParseStandard vs the host parser (itself checked against the oracle) on the corpus + fuzz"""
    from tests.test_abi_cpu import CORPUS, _fuzz
    for with_seconds in (False, True):
        texts = CORPUS + _fuzz(6000, 11 + with_seconds)
        texts = [t.encode("utf-8", errors="surrogateescape") if isinstance(t, str) else t for t in texts]
        masks, status, tz = ctx.parse_batch(texts, with_seconds)
        n_ok = 0
        for i, t in enumerate(texts):
            try:
                want = api._parse(t, with_seconds)
            except api.ParseError:
                want = None
            if status[i] == 2:  # an "@every ..." the device hands to the host parser, which validates the duration
                assert b"@every " in t and (want is None or want.Every), t
                continue
            if want is None:
                assert status[i] == 1, (t, status[i])
                assert not masks[:, i].any()
            elif want.Every:  # plain h/m/s durations are parsed on the device, into the EVERY_TAG encoding
                assert status[i] == 0 and tuple(int(x) for x in masks[:, i]) == want.masks(), t
            else:
                assert status[i] == 0, (t, status[i])
                assert tuple(int(x) for x in masks[:, i]) == want.masks(), t
                assert (tz[i] or "") == want.Location, (t, tz[i], want.Location)
                n_ok += 1
        assert n_ok > 200
    # the mirrored interface, incl. error propagation and zone validation
    got = ctx.ParseStandardBatch(["*/5 * * * *", "CRON_TZ=Asia/Tokyo 0 9 * * *", "@every 90s", "@daily"])
    assert got[0] == api.ParseStandard("*/5 * * * *") and got[1].Location == "Asia/Tokyo" and got[2].Every == 90
    with pytest.raises(api.ParseError):
        ctx.ParseStandardBatch(["* * * * *", "60 31 30 2 *"])  # K1
    with pytest.raises(api.ParseError):
        ctx.ParseStandardBatch(["TZ=Nope/Nowhere * * * * *"])
    assert ctx.parse_batch([])[0].shape == (6, 0)


def test_zero_copy_and_null_seconds_plane(ctx, names):
    """pinned host buffers are used in place (zero-copy); a NULL seconds plane means second 0 (ParseStandard's output)"""
    n = 300_001
    masks = synth.gen_specs(n, seed=9)
    assert (masks[0] == 1).all()
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=9)
    frm = synth.gen_times(n, ts(2024), ts(2034), seed=9)
    want = po.batch_next_n(masks, frm, 3, zid, names)
    hm, hf = api.PinnedArray((6, n), np.uint64), api.PinnedArray(n, np.int64)
    hz, ho = api.PinnedArray(n, np.int32), api.PinnedArray((n, 3), np.int64)
    hm.array[:] = masks; hf.array[:] = frm; hz.array[:] = zid
    for five in (False, True):
        ho.array[:] = 0
        ctx.next_n(hm.array, hf.array, 3, hz.array, out=ho.array, five_field=five)        # pinned: in place
        assert np.array_equal(ho.array, want), five
        assert np.array_equal(ctx.next_n(masks, frm, 3, zid, five_field=five), want), five   # pageable: staged copies
    w1 = po.batch_next_n(masks, frm, 1)
    ho1 = api.PinnedArray((n, 1), np.int64)
    ctx.next_n(hm.array, hf.array, 1, out=ho1.array, five_field=True)
    assert np.array_equal(ho1.array, w1)
    with pytest.raises(api.B200CronError):  # error words still come back on the zero-copy path
        hf.array[0] = 2**62
        ctx.next_n(hm.array, hf.array, 1, out=ho1.array)


def test_unmet_zero_copy_on_pinned_buffers(ctx, names):
    n = 50_001
    masks = synth.gen_specs(n, seed=61, dense_minutes=0.3)
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=61)
    now = ts(2026, 3, 29, 12)
    earliest = now - synth.gen_times(n, 0, 2 * 86400, seed=61)
    want = ctx.unmet_since(masks, earliest, now, zone_id=zid)  # pageable: staged
    P = api.PinnedArray
    hm, he, hz = P((6, n), np.uint64), P(n, np.int64), P(n, np.int32)
    hm.array[:] = masks; he.array[:] = earliest; hz.array[:] = zid
    owners = dict(last=P(n, np.int64), next=P(n, np.int64), missed=P(n, np.uint32), status=P(n, np.uint8),
                  decision=P(n, np.uint8))  # keep the pinned allocations alive while their views are in use
    out = {k: v.array for k, v in owners.items()}
    got = ctx.unmet_since(hm.array, he.array, now, zone_id=hz.array, out=out)  # pinned: in place
    for k in KEYS:
        assert np.array_equal(got[k], want[k]), k


def test_overlapping_independent_launches(ctx, names):
    """b200cron_ctx_set_overlap: back-to-back k == 1 launches over independent batches may overlap on the GPU
    (programmatic dependent launch); every batch must still get exactly the oracle's answers"""
    import torch
    nb, n = 6, 150_001
    stream = torch.cuda.Stream()
    st = stream.cuda_stream
    masks = [synth.gen_specs(n, seed=70 + b) for b in range(nb)]
    frm = [synth.gen_times(n, ts(2024), ts(2034), seed=70 + b) for b in range(nb)]
    zid = [synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=70 + b) if b % 2 else None for b in range(nb)]
    with torch.cuda.stream(stream):
        dm = [torch.from_numpy(m.view(np.int64)).cuda() for m in masks]
        df = [torch.from_numpy(f).cuda() for f in frm]
        dz = [None if z is None else torch.from_numpy(z).cuda() for z in zid]
        out = [torch.full((n, 1), 7, dtype=torch.int64, device="cuda") for _ in range(nb)]
    stream.synchronize()
    ctx.set_overlap(True)
    try:
        for rep in range(3):
            for b in range(nb):
                ctx.next_n_dev(dm[b], df[b], out[b], 1, dz[b], stream=st)
        ctx.sync_check(stream=st)
    finally:
        ctx.set_overlap(False)
    for b in range(nb):
        want = po.batch_next_n(masks[b], frm[b], 1, zid[b], names)
        assert np.array_equal(out[b].cpu().numpy(), want), b


def test_unmet_kernel_pair_on_concurrent_streams(ctx, names):
    """the easy/hard unmet kernel pair hands Crons over through a small pool of scratch lists: calls on different
    streams (more calls than lists, growing batch sizes) must not disturb each other"""
    import torch
    sizes = [3_001, 20_000, 777, 45_003, 12_345, 60_000, 5]
    now = ts(2026, 3, 29, 12)
    calls = []
    for c, n in enumerate(sizes):
        masks = synth.gen_specs(n, seed=90 + c, dense_minutes=0.5)
        dense = np.array([bin(int(x) & ((1 << 60) - 1)).count("1") for x in masks[1]]) > 6
        masks[2] = np.where(dense, masks[2] & np.uint64(0x800000000000000E), masks[2])  # affordable for the oracle
        masks[2] = np.where(masks[2] & np.uint64(0xFFFFFF), masks[2], np.uint64(1 << 2))
        zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=90 + c)
        earliest = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=90 + c)
        stream = torch.cuda.Stream()
        with torch.cuda.stream(stream):
            dm, dz, de = torch.from_numpy(masks.view(np.int64)).cuda(), torch.from_numpy(zid).cuda(), torch.from_numpy(earliest).cuda()
            o = dict(last=torch.empty(n, dtype=torch.int64, device="cuda"), next=torch.empty(n, dtype=torch.int64, device="cuda"),
                     missed=torch.empty(n, dtype=torch.int32, device="cuda"), status=torch.empty(n, dtype=torch.uint8, device="cuda"),
                     decision=torch.empty(n, dtype=torch.uint8, device="cuda"))
        calls.append((masks, zid, earliest, stream, dm, dz, de, o))
    torch.cuda.synchronize()
    for rep in range(2):
        for masks, zid, earliest, stream, dm, dz, de, o in calls:
            ctx.unmet_since_dev(dm, de, now, o, 0, None, None, None, None, dz, stream=stream.cuda_stream)
    for _, _, _, stream, *_ in calls:
        ctx.sync_check(stream=stream.cuda_stream)
    for masks, zid, earliest, stream, dm, dz, de, o in calls:
        want = po.batch_unmet(masks, earliest, now, zone_id=zid, zone_names=names)
        got = {k: v.cpu().numpy().view(want[k].dtype) for k, v in o.items()}
        _cmp(got, want, len(earliest))


def test_every_schedules_on_the_device(ctx, names):
    """ConstantDelaySchedule ("@every <dur>", robfig constantdelay.go; reached by the reference at
    cron_controller.go:392,405,409,436) as a schedule kind inside ordinary batches: next_n (k = 1 and chained),
    unmet_since (pageable and pinned buffers), the unmet list and the device parser"""
    from tests.test_unmet_host import every_batch
    masks, is_every, delay, earliest, now = every_batch(50_001, seed=7)
    n = masks.shape[1]
    zid = np.where(np.arange(n) % 3 == 0, names.index("Europe/Berlin"), 0).astype(np.int32)
    for k in (1, 4, 5):
        got = ctx.next_n(masks, earliest, k, zid)
        assert np.array_equal(got, po.batch_next_n(masks, earliest, k, zid, names)), k
        assert np.array_equal(got[is_every], earliest[is_every, None] + delay[is_every, None] * np.arange(1, k + 1))
    assert np.array_equal(ctx.next_n(masks, earliest, 1), po.batch_next_n(masks, earliest, 1))  # all-UTC kernel
    rng = np.random.default_rng(1)
    policy, active = rng.integers(0, 3, n).astype(np.uint8), rng.integers(0, 3, n).astype(np.uint32)
    flags = (rng.random(n) < 0.05).astype(np.uint8)
    for z in (zid, None):
        got = ctx.unmet_since(masks, earliest, now, 0, policy, active, flags, None, z)
        _cmp(got, po.batch_unmet(masks, earliest, now, 0, policy, active, flags, None, z, names if z is not None else None))
    off, lst = ctx.unmet_list(masks, earliest, got["missed"], got["decision"], 100)
    woff, wlst = po.batch_unmet_list(masks, earliest, got["missed"], got["decision"], 100)
    assert np.array_equal(off, woff) and np.array_equal(lst, wlst)
    # the reference's scalar shapes through the same path
    s = api.ParseStandard("@every 1h30m")
    assert ctx.Next(s, 1000) == 1000 + 5400
    assert ctx.getNextSchedule("@every 90s", now - 200, now) == (now - 20, now + 90)
    # device parser: plain h/m/s durations come back encoded, the rest goes to the host parser
    texts = ["@every 90s", "@every 1h30m", "@every 0s", "@every 2h0m1s", "@every 1.5h", "@every 500ms", "@every -5s",
             "@every 1h 30m", "@every 12345678901s", "@every 999999999h", "@every", "@every 5", "@every 5x", "@every 0"]
    m, status, _ = ctx.parse_batch(texts)
    for i, t in enumerate(texts):
        try:
            want = api.ParseStandard(t)
        except api.ParseError:
            want = None
        if status[i] == 0:
            assert want is not None and tuple(int(x) for x in m[:, i]) == want.masks(), t
        else:
            assert status[i] == 2 if t.startswith("@every ") else status[i] == 1, (t, status[i])
    assert status[:4].tolist() == [0, 0, 0, 0] and int(m[3, 2]) == 1 and int(m[3, 3]) == 7201


def test_unmet_null_seconds_plane_and_every_second_specs(ctx, names):
    """masks[0] == NULL means second 0 for unmet_since too; and the chains of specs that fire every second are
    fast-forwarded hour by hour next to a transition (they used to be walked one fire at a time)"""
    n = 30_001
    masks = synth.gen_specs(n, seed=17, dense_minutes=0.3)
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=17)
    now = ts(2026, 3, 29, 12)
    earliest = now - synth.gen_times(n, 0, 3 * 86400, seed=17)
    want = po.batch_unmet(masks, earliest, now, zone_id=zid, zone_names=names)
    out = dict(last=np.empty(n, np.int64), next=np.empty(n, np.int64), missed=np.empty(n, np.uint32),
               status=np.empty(n, np.uint8), decision=np.empty(n, np.uint8))
    planes = api._planes(masks, five_field=True)[0]
    ctx._check(api.lib().b200cron_unmet_since(ctx._h, planes, zid.ctypes.data, earliest.ctypes.data, now, 0, None, None,
                                              None, None, n, out["last"].ctypes.data, out["next"].ctypes.data,
                                              out["missed"].ctypes.data, out["status"].ctypes.data,
                                              out["decision"].ctypes.data))
    _cmp(out, want)
    texts = ["* * 0-5,9,22-23 * * *", "*/7 * 1-3 * * *", "59 59 * * * *", "* 59 0,2,4,23 * * *", "* * * * * *",
             "30-59 * 3-4 * * *", "* * 12 * * *"]
    base = _masks_of(texts, with_seconds=True)
    zs = [names.index(z) for z in ("Europe/Oslo", "Europe/Chisinau", "America/New_York", "Australia/Lord_Howe",
                                   "Asia/Tehran", "America/Havana", "Pacific/Chatham", "UTC")]
    S, Z = len(texts), len(zs)
    m = np.ascontiguousarray(np.tile(base, (1, Z)))
    z = np.repeat(np.array(zs, np.int32), S)
    for back in (6 * 3600, 86400 + 1800, 3 * 86400):
        e = np.full(S * Z, now - back)
        got = ctx.unmet_since(m, e, now, zone_id=z)
        _cmp(got, po.batch_unmet(m, e, now, zone_id=z, zone_names=names), back)


def test_resident_spec_set(ctx, names):
    """b200cron_specset_*: masks and zone ids stay in HBM, a tick moves only times in and results out; answers must be
    those of the host-pointer calls (and so the oracle's), through pageable and pinned per-tick buffers, after an
    update of a range, and for 5-field sets without a seconds plane"""
    n = 300_001
    masks = synth.gen_specs(n, seed=21, dense_minutes=0.2)
    zid = synth.gen_zone_ids(n, len(names), _dst_zones(ctx, names), seed=21)
    frm = synth.gen_times(n, ts(2024), ts(2034), seed=21)
    P = api.PinnedArray
    hf, ho1, ho4 = P(n, np.int64), P((n, 1), np.int64), P((n, 4), np.int64)
    hf.array[:] = frm
    for five, z in ((False, zid), (True, None)):
        st = ctx.specset(masks, z, five_field=five)
        try:
            assert len(st) == n
            want1 = po.batch_next_n(masks, frm, 1, z, names if z is not None else None)
            want4 = po.batch_next_n(masks, frm, 4, z, names if z is not None else None)
            assert np.array_equal(st.next_n(frm, 1), want1)                      # pageable: staged
            assert np.array_equal(st.next_n(hf.array, 1, out=ho1.array), want1)  # pinned: in place
            assert np.array_equal(st.next_n(hf.array, 4, out=ho4.array), want4)
            assert np.array_equal(st.next_n(frm, 4), want4)
            now = ts(2026, 3, 29, 12)
            m = 60_000
        finally:
            st.close()
    # unmet over a set, then an update of a range in the middle
    m = 80_001
    mm = np.ascontiguousarray(masks[:, :m])
    now = ts(2026, 3, 29, 12)
    earliest = now - synth.gen_times(m, 0, 2 * 86400, seed=22)
    rng = np.random.default_rng(3)
    policy, active = rng.integers(0, 3, m).astype(np.uint8), rng.integers(0, 3, m).astype(np.uint32)
    st = ctx.specset(mm, zid[:m])
    try:
        want = ctx.unmet_since(mm, earliest, now, 0, policy, active, None, None, zid[:m])
        _cmp(st.unmet_since(earliest, now, 0, policy, active), want, "pageable")
        he, hp, ha = P(m, np.int64), P(m, np.uint8), P(m, np.uint32)
        he.array[:] = earliest; hp.array[:] = policy; ha.array[:] = active
        owners = dict(last=P(m, np.int64), next=P(m, np.int64), missed=P(m, np.uint32), status=P(m, np.uint8), decision=P(m, np.uint8))
        out = {k: v.array for k, v in owners.items()}
        _cmp(st.unmet_since(he.array, now, 0, hp.array, ha.array, out=out), want, "pinned")
        lo, cnt = 12_345, 3_001
        new_m = synth.gen_specs(cnt, seed=99)
        new_z = np.zeros(cnt, np.int32)
        st.update(lo, new_m, new_z)
        mm2, z2 = mm.copy(), zid[:m].copy()
        mm2[:, lo:lo + cnt] = new_m
        z2[lo:lo + cnt] = 0
        _cmp(st.unmet_since(earliest, now, 0, policy, active), po.batch_unmet(mm2, earliest, now, 0, policy, active, None, None, z2, names), "updated")
        with pytest.raises(api.B200CronError):
            st.update(m - 5, new_m, new_z)  # range outside the set
    finally:
        st.close()


def _c4_inputs(n4, now):
    """the C4 workload exactly as bench.py generates it (seed 41, dense minutes 0.5, 30-day backlog)"""
    rng = np.random.default_rng(41)
    pol = rng.integers(0, 3, n4).astype(np.uint8)
    act = rng.choice([0, 1, 2], size=n4, p=[.6, .3, .1]).astype(np.uint32)
    fl = (rng.random(n4) < 0.02).astype(np.uint8)
    dl = np.where(rng.random(n4) < 0.01, now - 100, -2**63).astype(np.int64)
    return pol, act, fl, dl


def test_C4_full_size_sampled_parity(ctx, names):
    """BASELINE config 4 at its full size and on the benched distribution (no thinning of the dense specs): 10M Crons
    with a 30-day backlog, now inside a European DST day (getNextSchedule, cron_controller.go:389-437).  Oracle parity
    on every 997th Cron plus a sample of the Crons the closed form gave up on (the general-walk path)."""
    import torch
    n4 = 10_000_000
    now = ts(2026, 3, 29, 12)
    st = torch.cuda.current_stream().cuda_stream
    dst = _dst_zones(ctx, names)
    pool = torch.from_numpy(synth.zone_pool(len(names), dst)).cuda()
    dm = torch.empty((6, n4), dtype=torch.int64, device="cuda")
    dz = torch.empty(n4, dtype=torch.int32, device="cuda")
    de = torch.empty(n4, dtype=torch.int64, device="cuda")
    ctx.synth_specs_dev(dm, n4, 41, 0, False, 0.5, stream=st)
    ctx.synth_zones_dev(dz, n4, pool, 41, 0, stream=st)
    ctx.synth_times_dev(de, n4, -86400, 86400, 41, 0, stream=st)
    de += now - 30 * 86400
    pol, act, fl, dl = _c4_inputs(n4, now)
    t = lambda a: torch.from_numpy(a).cuda()  # noqa: E731
    o = dict(last=torch.empty(n4, dtype=torch.int64, device="cuda"), next=torch.empty(n4, dtype=torch.int64, device="cuda"),
             missed=torch.empty(n4, dtype=torch.int32, device="cuda"), status=torch.empty(n4, dtype=torch.uint8, device="cuda"),
             decision=torch.empty(n4, dtype=torch.uint8, device="cuda"))
    ctx.unmet_since_dev(dm, de, now, o, 0, t(pol), t(act), t(fl), t(dl), dz, stream=st)
    ctx.sync_check(stream=st)
    deferred, ndef = ctx.last_deferred(stream=st)
    assert 0 < ndef < n4 // 10 and len(deferred) == ndef
    idx = np.unique(np.concatenate([np.arange(0, n4, 997), np.sort(deferred)[::97].astype(np.int64)])).astype(np.uint64)
    ms = synth.gen_specs(0, 41, dense_minutes=0.5, idx=idx)
    zs = synth.gen_zone_ids(0, len(names), dst, 41, idx=idx)
    es = now - 30 * 86400 + synth.gen_times(0, -86400, 86400, 41, idx=idx)
    ii = idx.astype(np.int64)
    want = po.batch_unmet(ms, es, now, 0, pol[ii], act[ii], fl[ii], dl[ii], zs, names)
    gi = torch.from_numpy(ii).cuda()
    got = {k: v[gi].cpu().numpy().view(want[k].dtype) for k, v in o.items()}
    _cmp(got, want, "C4 sample")
    assert want["missed"].max() >= 43_000      # the `* * * * *` chains were really walked by the oracle
    # size-independent properties over all 10M: next > now unless failed/gated; last <= now; decision consistent
    nxt, last, dec, miss = o["next"], o["last"], o["decision"], o["missed"].to(torch.int64)
    live = dec >= api.DEC_REQUEUE_ONLY
    assert bool(((nxt > now) | ~live).all()) and bool(((last <= now) | (last == po.ZERO)).all())
    assert bool(((miss > 0) == ((last != po.ZERO) & live))[live].all())
    assert bool(((dec == api.DEC_REQUEUE_ONLY) == (live & (last == po.ZERO)))[live].all())


def test_history_chunking_wide_keys_and_length_limit(ctx):
    """history_prune works through a tile in chunks of whole segments (sized by records), keeps 64-bit keys when a
    chunk's keys do not fit 32-bit differences, ranks a long segment from global memory and rejects an absurd one"""
    import torch
    rng = np.random.default_rng(19)
    for nseg, maxlen, wide in [(700, 60, False), (5, 2500, False), (300, 16, True), (40, 300, True), (2, 20_000, False)]:
        lens = rng.integers(0, maxlen + 1, nseg)
        seg_off = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        m = int(seg_off[-1])
        if wide:  # nanosecond-like and full-range keys, with ties
            key = rng.integers(-2**62, 2**62, m, dtype=np.int64)
            key[rng.random(m) < 0.3] = key[0] if m else 0
        else:
            key = (1_700_000_000 - rng.integers(0, 500, m)).astype(np.int64)
        for limit in (None, rng.integers(-1, 30, nseg).astype(np.int64), np.full(nseg, 2**40, np.int64)):
            perm, keep = ctx.history_prune(seg_off, key, limit)
            wperm, wkeep = po.batch_history(seg_off, key, limit)
            assert np.array_equal(perm, wperm) and np.array_equal(keep, wkeep), (nseg, maxlen, wide)
    seg_off = np.array([0, 3, 3 + 65_537, 3 + 65_537 + 2], np.uint64)
    key = rng.integers(0, 1000, int(seg_off[-1])).astype(np.int64)
    with pytest.raises(api.B200CronError) as e:
        ctx.history_prune(seg_off, key, None)
    assert e.value.code == api.EINVAL and "65536" in str(e.value)
    st = torch.cuda.current_stream().cuda_stream
    ds, dk = torch.from_numpy(seg_off.view(np.int64)).cuda(), torch.from_numpy(key).cuda()
    perm = torch.empty(len(key), dtype=torch.int32, device="cuda")
    keep = torch.empty(len(key), dtype=torch.uint8, device="cuda")
    ctx.history_prune_dev(ds, dk, None, perm, keep, stream=st)
    with pytest.raises(api.B200CronError) as e:
        ctx.sync_check(stream=st)
    assert e.value.code == api.EINVAL
    p = perm.cpu().numpy()
    assert np.array_equal(p[3:3 + 65_537], np.arange(65_537)) and bool((keep[3:3 + 65_537] == 1).all())  # left as is
    w0, _ = po.history_prune(key[:3], None)
    assert np.array_equal(p[:3].view(np.uint32), w0)                                                     # neighbours fine


def test_unmet_period_clock_class_on_the_device(ctx, names):
    """the unmet kernels answer a zoned Cron whose whole window lies in one zone period with the UTC machine on that
    period's clock, and hand it to the general path when a Next() leaves the period (tests/test_unmet_host.py has the
    host-side twin): mixed zones, windows ending around transitions, sparse specs with a far next fire"""
    rng = np.random.default_rng(77)
    zones = ["Asia/Tokyo", "Asia/Kolkata", "Asia/Kathmandu", "Europe/Berlin", "America/New_York", "Australia/Sydney",
             "America/Santiago", "Pacific/Chatham", "Africa/Cairo", "Atlantic/Azores", "Asia/Tehran", "America/Sao_Paulo",
             "Australia/Lord_Howe", "UTC", "Etc/GMT+12", "Pacific/Apia"]
    n = 60_001
    masks = synth.gen_specs(n, seed=78, dense_minutes=0.3)
    zid = np.array([names.index(z) for z in zones], np.int32)[rng.integers(0, len(zones), n)]
    for now in (ts(2026, 3, 29, 12), ts(2026, 7, 15, 3, 7, 9), ts(2026, 10, 25, 0, 59, 59), ts(2027, 1, 1)):
        earliest = now - synth.gen_times(n, -7200, 20 * 86400, seed=int(now % 1000))
        got = ctx.unmet_since(masks, earliest, now, zone_id=zid)
        _cmp(got, po.batch_unmet(masks, earliest, now, zone_id=zid, zone_names=names), now)
    z = names.index("Europe/Berlin")
    when = ctx.zone_periods(z)[0]
    texts = ["* * * * *", "*/5 * * * *", "30 2 * * *", "0 0 * * *", "0 0 1 * *", "15 3 * * 0", "59 23 31 12 *", "0 12 1 1 *"]
    base = _masks_of(texts)
    S = len(texts)
    for t in [int(w) for w in when if ts(2025) < w < ts(2027)]:
        for back, ahead in [(40 * 86400, -33 * 86400), (40 * 86400, -1), (86400, -1), (3600, 3600), (10, 40 * 86400), (0, 86400)]:
            zz = np.full(S, z, np.int32)
            e, now = np.full(S, t - back), t + ahead
            if now < t - back:
                continue
            _cmp(ctx.unmet_since(base, e, now, zone_id=zz), po.batch_unmet(base, e, now, zone_id=zz, zone_names=names), (t, back, ahead))


def test_reserve_unmet_scratch_means_no_allocation_in_the_tick(names):
    """b200cron_ctx_reserve_unmet sizes the index lists up front: unmet_since_dev calls up to that size then leave the
    device's free memory untouched (and still answer like the oracle)"""
    import torch
    c = api.Context(devices=[0])
    try:
        n = 150_001
        masks = synth.gen_specs(n, seed=33, dense_minutes=0.3)
        zid = synth.gen_zone_ids(n, len(names), np.arange(1, len(names)), seed=33)
        now = ts(2026, 3, 29, 12)
        earliest = now - synth.gen_times(n, 0, 3 * 86400, seed=33)
        st = torch.cuda.current_stream().cuda_stream
        dm, dz, de = torch.from_numpy(masks.view(np.int64)).cuda(), torch.from_numpy(zid).cuda(), torch.from_numpy(earliest).cuda()
        o = dict(last=torch.empty(n, dtype=torch.int64, device="cuda"), next=torch.empty(n, dtype=torch.int64, device="cuda"),
                 missed=torch.empty(n, dtype=torch.int32, device="cuda"), status=torch.empty(n, dtype=torch.uint8, device="cuda"),
                 decision=torch.empty(n, dtype=torch.uint8, device="cuda"))
        c.reserve_unmet(200_000)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        for _ in range(4):  # more calls than scratch buffers
            c.unmet_since_dev(dm, de, now, o, zone_id=dz, stream=st)
        c.sync_check(stream=st)
        assert torch.cuda.mem_get_info()[0] == free0
        want = po.batch_unmet(masks, earliest, now, zone_id=zid, zone_names=names)
        _cmp({k: v.cpu().numpy().view(want[k].dtype) for k, v in o.items()}, want)
    finally:
        c.close()
