"""CPU checks of the kernels' walk source (host build, tests/shim.py) against the oracle: zone table vs Go-style
lookup, Next / NextN in UTC and across every IANA zone incl. DST edges."""
import numpy as np
import pytest

from cron_operator_b200 import synth
from oracle import pyoracle as po
from tests import shim

ts = synth.ts
NAMES = shim.zone_names()
EDGE_ZONES = ["America/New_York", "Europe/Berlin", "Europe/London", "Australia/Lord_Howe", "Asia/Kathmandu",
              "Pacific/Chatham", "Africa/Cairo", "Asia/Beirut", "America/Havana", "America/Santiago",
              "America/Sao_Paulo", "Atlantic/Azores", "Asia/Tehran", "America/Nuuk", "Pacific/Apia",
              "Africa/Casablanca", "America/Asuncion", "Asia/Gaza", "Antarctica/Troll", "Australia/Sydney",
              "Pacific/Auckland", "America/St_Johns", "Asia/Amman", "Europe/Chisinau", "America/Scoresbysund"]


def test_zone_table_matches_go_lookup():
    rng = np.random.default_rng(1)
    assert NAMES[0] == "UTC" and len(NAMES) > 400
    for z, name in enumerate(NAMES):
        if z == 0:
            continue
        loc = po.load_location(name)
        when, off = shim.zone_entries(z)
        assert when[0] == po.ZERO and np.all(np.diff(when.astype(np.float64)) > 0)
        probes = list(rng.integers(ts(1850), ts(2194), size=40))
        for w in when[1:]:
            if ts(1850) < w < ts(2194):
                probes += [int(w) - 1, int(w)]
        probes = np.array(probes[:400], dtype=np.int64)
        got = off[np.searchsorted(when, probes, side="right") - 1]
        want = np.array([po.lookup(loc, int(p))[0] for p in probes])
        bad = np.nonzero(got != want)[0]
        assert len(bad) == 0, (name, probes[bad][:5], got[bad][:5], want[bad][:5])


def _masks_of(texts, with_seconds=False):
    specs = [po.parse_standard(t, with_seconds)[0] for t in texts]
    return np.ascontiguousarray(
        np.array([[getattr(s, f) for s in specs] for f in ("second", "minute", "hour", "dom", "month", "dow")],
                 dtype=np.uint64))


def test_next_utc_random_specs():
    n = 20000
    masks = synth.gen_specs(n, seed=11)
    frm = synth.gen_times(n, ts(2024), ts(2034), seed=11)
    got = shim.next_n(masks, frm, k=1)
    want = po.batch_next_n(masks, frm, k=1)
    bad = np.nonzero(got != want)[0]
    assert len(bad) == 0, (bad[:5], masks[:, bad[:1]], frm[bad[:5]], got[bad[:5]], want[bad[:5]])
    assert (want == po.ZERO).sum() > 0  # planted unschedulable specs are present


def test_next_n_utc_with_seconds():
    n = 4000
    masks = synth.gen_specs(n, seed=12, with_seconds=True)
    frm = synth.gen_times(n, ts(2024), ts(2036), seed=12)
    got = shim.next_n(masks, frm, k=16)
    want = po.batch_next_n(masks, frm, k=16)
    assert np.array_equal(got, want)


def test_utc_specialisation_equals_general_walk():
    """no zone ids -> the all-UTC instantiation (next1_kernel<true>); zone id 0 everywhere -> the general walk bound to
    UTC.  Same answers, and the oracle's, for single and chained calls."""
    n = 30000
    for with_seconds, k, seed in [(False, 1, 21), (True, 5, 22)]:
        masks = synth.gen_specs(n, seed=seed, with_seconds=with_seconds)
        frm = synth.gen_times(n, ts(2024), ts(2036), seed=seed)
        a = shim.next_n(masks, frm, k=k)
        b = shim.next_n(masks, frm, k=k, zone_id=np.zeros(n, np.int32))
        assert np.array_equal(a, b)
        assert np.array_equal(a[:3000], po.batch_next_n(masks[:, :3000].copy(), frm[:3000], k=k))


def test_next_extreme_times_utc():
    texts = ["* * * * *", "0 0 29 2 *", "59 23 31 12 *", "0 0 1 1 *", "*/7 */5 * * 1"]
    masks = _masks_of(texts)
    for t in [ts(1970) - 1, -1, 0, ts(1900), ts(2099, 12, 31, 23, 59, 59), ts(2100, 2, 28, 23, 59, 59), ts(1, 1, 1) + 86400,
              ts(9990), ts(2400, 2, 29), -62135596800 + 400 * 86400]:
        frm = np.full(len(texts), t, dtype=np.int64)
        assert np.array_equal(shim.next_n(masks, frm, k=3), po.batch_next_n(masks, frm, k=3)), t


def test_degenerate_masks():
    # empty fields (the "," quirk), garbage high bits, bit 0 of dom/month set
    m = _masks_of(["* * * * *"] * 6)
    m[1, 0] = 0
    m[2, 1] = 0
    m[3, 2] = 0          # dom empty with star on dow -> AND rule -> never
    m[4, 3] = 0
    m[5, 4] = np.uint64(0xFF80)   # only invalid dow bits, no star; dom has star -> AND -> never
    m[3, 5] = np.uint64((1 << 63) | 1 | (1 << 40))  # only bits upstream never tests
    frm = np.full(6, ts(2025, 6, 1), dtype=np.int64)
    got = shim.next_n(m, frm, k=2)
    want = po.batch_next_n(m, frm, k=2)
    assert np.array_equal(got, want) and np.all(got == po.ZERO)


@pytest.mark.parametrize("name", EDGE_ZONES)
def test_next_n_around_every_transition(name):
    """every transition of the zone in 2015-2040 (+ the footer era 2038-2045): start a few seconds/hours/days before it"""
    z = NAMES.index(name)
    when, _ = shim.zone_entries(z)
    trans = [int(w) for w in when if ts(2015) < w < ts(2046)]
    texts = ["* * * * *", "*/5 * * * *", "30 2 * * *", "0 0 * * *", "0 1 * * *", "59 23 * * *", "0 0 1 * *",
             "15 3 * * 0", "0 0 31 3,10 *", "0 */2 * * *", "30 0 * * *", "0 0 * 4,11 *", "45 1-3 25-31 * 0",
             "0 12 1 1 *", "0 0 30 2 *", "0 3 * * *", "0 0 2 * *", "*/30 * * * *"]
    base = _masks_of(texts)
    offs = [-1, -2, -61, -3599, -3601, -7201, -86399, -86401, -2 * 86400 - 5, -40 * 86400, 0, 1, 1799, 3600, 86400]
    starts = np.array([t + o for t in trans for o in offs], dtype=np.int64)
    S, T = len(texts), len(starts)
    masks = np.ascontiguousarray(np.repeat(base, T, axis=1))
    frm = np.tile(starts, S)
    zid = np.full(S * T, z, dtype=np.int32)
    got = shim.next_n(masks, frm, k=4, zone_id=zid)
    want = po.batch_next_n(masks, frm, k=4, zone_id=zid, zone_names=NAMES)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert len(bad) == 0, (name, [(texts[b // T], int(frm[b]), got[b].tolist(), want[b].tolist()) for b in bad[:3]])


def test_next_n_seconds_field_around_transitions():
    texts = ["*/10 * * * * *", "59 59 * * * *", "59 59 23 * * *", "0 30 0 * * *", "30 */15 1-3 * * *", "59 59 1 * * 0"]
    base = _masks_of(texts, with_seconds=True)
    for name in ["America/New_York", "Australia/Lord_Howe", "Africa/Cairo", "America/Havana", "Atlantic/Azores",
                 "Europe/London", "America/Santiago"]:
        z = NAMES.index(name)
        when, _ = shim.zone_entries(z)
        trans = [int(w) for w in when if ts(2022) < w < ts(2040)]
        starts = np.array([t + o for t in trans for o in (-7205, -3601, -61, -2, -1, 0, 1, 3599, 86399)], dtype=np.int64)
        S, T = len(texts), len(starts)
        masks = np.ascontiguousarray(np.repeat(base, T, axis=1))
        frm = np.tile(starts, S)
        zid = np.full(S * T, z, dtype=np.int32)
        got = shim.next_n(masks, frm, k=5, zone_id=zid)
        want = po.batch_next_n(masks, frm, k=5, zone_id=zid, zone_names=NAMES)
        assert np.array_equal(got, want), name


def test_next_n_all_zones_random():
    """C3-shaped: random specs with a seconds field, mixed zones, K=16"""
    n = 30000
    masks = synth.gen_specs(n, seed=21, with_seconds=True)
    dst = [z for z in range(1, len(NAMES)) if (shim.zone_entries(z)[0] > ts(2030)).any()]
    zid = synth.gen_zone_ids(n, len(NAMES), np.array(dst), seed=21)
    frm = synth.gen_times(n, ts(2024), ts(2036), seed=21)
    got = shim.next_n(masks, frm, k=16, zone_id=zid)
    want = po.batch_next_n(masks, frm, k=16, zone_id=zid, zone_names=NAMES)
    bad = np.nonzero((got != want).any(axis=1))[0]
    assert len(bad) == 0, [(NAMES[zid[b]], int(frm[b]), [hex(int(x)) for x in masks[:, b]]) for b in bad[:3]]


def test_next_historic_and_far_future_zones():
    """pre-1972 second-granular LMT offsets, the 1800s, and the expanded-footer era"""
    texts = ["* * * * *", "0 0 * * *", "30 2 * * 0", "0 0 1 * *"]
    base = _masks_of(texts)
    rng = np.random.default_rng(4)
    for name in ["Africa/Monrovia", "Europe/Amsterdam", "Asia/Manila", "America/Juneau", "Europe/Dublin",
                 "America/New_York", "Australia/Lord_Howe", "Pacific/Kiritimati", "Asia/Kolkata"]:
        z = NAMES.index(name)
        starts = np.concatenate([rng.integers(ts(1880), ts(1975), size=60), rng.integers(ts(2037), ts(2190), size=60)])
        S, T = len(texts), len(starts)
        masks = np.ascontiguousarray(np.repeat(base, T, axis=1))
        frm = np.tile(starts, S).astype(np.int64)
        zid = np.full(S * T, z, dtype=np.int32)
        got = shim.next_n(masks, frm, k=3, zone_id=zid)
        want = po.batch_next_n(masks, frm, k=3, zone_id=zid, zone_names=NAMES)
        assert np.array_equal(got, want), name


def test_calendar_primitives_exhaustive():
    """the kernels' civil-from-days (Neri-Schneider) and days-before-month against python's calendar: every day of
    years 1..9999 through the host build (next fire of '0 0 * * *' from 23:59:59 UTC is the next midnight, and
    '0 0 1 * *' exercises month lengths / leap years)"""
    import datetime as dt
    first = dt.date(1, 1, 2).toordinal() - dt.date(1970, 1, 1).toordinal()
    last = dt.date(9993, 12, 31).toordinal() - dt.date(1970, 1, 1).toordinal()
    days = np.arange(first, last, dtype=np.int64)
    frm = days * 86400 - 1                                   # 23:59:59 of the previous day
    daily = _masks_of(["0 0 * * *"])
    got = shim.next_n(np.ascontiguousarray(np.repeat(daily, len(days), axis=1)), frm, 1)[:, 0]
    assert np.array_equal(got, days * 86400)
    # monthly: the next fire must be the first of the next month per python's calendar
    monthly = _masks_of(["0 0 1 * *"])
    got = shim.next_n(np.ascontiguousarray(np.repeat(monthly, len(days), axis=1)), days * 86400, 1)[:, 0]
    ords = days + dt.date(1970, 1, 1).toordinal()
    want = np.empty(len(days), dtype=np.int64)
    cache = {}
    for i, o in enumerate(ords[::1]):
        d = dt.date.fromordinal(int(o))
        key = (d.year, d.month)
        if key not in cache:
            ny, nm = (d.year + 1, 1) if d.month == 12 else (d.year, d.month + 1)
            cache[key] = (dt.date(ny, nm, 1).toordinal() - dt.date(1970, 1, 1).toordinal()) * 86400
        want[i] = cache[key]
    assert np.array_equal(got, want)
