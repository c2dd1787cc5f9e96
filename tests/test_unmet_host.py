"""CPU checks of the kernels' unmet-enumeration source (host build) against the oracle's literal getNextSchedule
chain (cron_controller.go:389-437) + gates/decision (:163-220)."""
import numpy as np
import pytest

from cron_operator_b200 import synth
from oracle import pyoracle as po
from tests import shim
from tests.test_walk_host import EDGE_ZONES, NAMES, _masks_of

ts = synth.ts
KEYS = ("last", "next", "missed", "status", "decision")


def _cmp(got, want, ctx=""):
    for k in KEYS:
        bad = np.nonzero(got[k] != want[k])[0]
        assert len(bad) == 0, (ctx, k, bad[:5], got[k][bad[:5]], want[k][bad[:5]])


def test_K3_through_kernel_source():
    now = ts(2026, 1, 1, 12)
    m = _masks_of(["*/1 * * * *", "0 0 30 2 *", "*/5 * * * *"])
    got = shim.unmet(m, np.full(3, now - 300), now)
    assert got["last"].tolist() == [now, po.ZERO, now]
    assert got["next"].tolist() == [now + 60, po.ZERO, now + 300]
    assert got["missed"].tolist() == [5, 0, 1]
    assert got["status"].tolist() == [0, 1, 0]
    assert got["decision"].tolist() == [po.DEC_CREATE, po.DEC_ERR_UNSCHEDULABLE, po.DEC_CREATE]


def test_gates_and_policies():
    now = ts(2026, 1, 1, 12)
    m = _masks_of(["*/5 * * * *"] * 10)
    earliest = np.array([now - 300] * 8 + [now + 10, now - 1], dtype=np.int64)
    policy = np.array([0, 1, 1, 2, 2, 0, 0, 0, 0, 0], dtype=np.uint8)
    active = np.array([3, 0, 2, 0, 4, 0, 0, 0, 0, 0], dtype=np.uint32)
    flags = np.array([0, 0, 0, 0, 0, 1, 2, 0, 0, 0], dtype=np.uint8)
    deadline = np.array([po.ZERO] * 7 + [now - 1, po.ZERO, now], dtype=np.int64)
    for nsec in (0, 5):
        got = shim.unmet(m, earliest, now, nsec, policy, active, flags, deadline)
        want = po.batch_unmet(m, earliest, now, nsec, policy, active, flags, deadline)
        _cmp(got, want)
    got = shim.unmet(m, earliest, now, 0, policy, active, flags, deadline)
    assert got["decision"].tolist() == [po.DEC_CREATE, po.DEC_CREATE, po.DEC_SKIP_FORBID, po.DEC_REPLACE_CREATE,
                                        po.DEC_REPLACE_CREATE, po.DEC_GATE_SUSPENDED, po.DEC_GATE_DELETED,
                                        po.DEC_GATE_DEADLINE, po.DEC_REQUEUE_ONLY, po.DEC_CREATE]
    # deadline == now: only a sub-second "now" is after it
    got = shim.unmet(m, earliest, now, 1, policy, active, flags, deadline)
    assert got["decision"][9] == po.DEC_GATE_DEADLINE


def test_unmet_utc_random_short_backlog():
    n = 6000
    masks = synth.gen_specs(n, seed=31, dense_minutes=0.3)
    now = ts(2026, 3, 29, 12)
    earliest = now - synth.gen_times(n, 0, 3 * 86400, seed=31)
    got = shim.unmet(masks, earliest, now)
    want = po.batch_unmet(masks, earliest, now)
    _cmp(got, want)
    assert want["missed"].max() > 100 and (want["status"] == 1).sum() > 0


def test_unmet_utc_long_backlog_sparse():
    """multi-year backlogs on specs that fire at most a few times a day (keeps the literal oracle affordable)"""
    n = 1500
    masks = synth.gen_specs(n, seed=32)
    masks[1] = np.uint64(1) << (synth.gen_times(n, 0, 60, seed=32, stream=400).astype(np.uint64))  # one minute value
    masks[2] &= np.uint64(0x800000000000000F) | (np.uint64(1) << np.uint64(13))                      # few hours
    masks[2] = np.where(masks[2] & np.uint64(0xFFFFFF), masks[2], np.uint64(1 << 7))
    now = ts(2031, 7, 4, 9, 30)
    earliest = now - synth.gen_times(n, 0, 9 * 366 * 86400, seed=33)
    got = shim.unmet(masks, earliest, now)
    want = po.batch_unmet(masks, earliest, now)
    _cmp(got, want)


def test_unmet_year_limit_inside_backlog():
    # "0 0 29 2 *" from 2095: fires 2096-02-29, then the next one (2104) is beyond the 5-year limit -> the chain
    # hits the zero time and the whole call is "unschedulable" even though one run was missed
    m = _masks_of(["0 0 29 2 *", "0 0 29 2 *", "0 0 29 2 *"])
    earliest = np.array([ts(2095), ts(2095), ts(2019)], dtype=np.int64)
    now_cases = [ts(2097), ts(2104, 3, 1)]
    for now in now_cases:
        got = shim.unmet(m, earliest, now)
        want = po.batch_unmet(m, earliest, now)
        _cmp(got, want, now)
    assert shim.unmet(m, earliest, ts(2104, 3, 1))["status"].tolist() == [1, 1, 1]


@pytest.mark.parametrize("name", EDGE_ZONES)
def test_unmet_windows_across_transitions(name):
    z = NAMES.index(name)
    when, _ = shim.zone_entries(z)
    trans = [int(w) for w in when if ts(2021) < w < ts(2041)][:24]
    texts = ["*/20 * * * *", "0 * * * *", "30 2 * * *", "0 0 * * *", "0 1 * * *", "59 23 * * *", "0 0 1 * *",
             "15 3 * * 0", "0 */2 * * *", "30 0 * * *", "45 1-3 25-31 * 0", "0 3 * * *", "0 0 2 * *", "*/30 0-4 * * *"]
    base = _masks_of(texts)
    # (earliest offset, now offset) relative to the transition
    wins = [(-3 * 86400, 3 * 86400), (-40 * 86400, 1), (-5, 86400 * 2), (-86400, -1), (-7200, 7200), (0, 86400),
            (-20 * 86400, 20 * 86400), (-3600, 0), (1, 5 * 86400)]
    cases = [(t + a, t + b) for t in trans for a, b in wins]
    for now_off in range(len(wins)):
        sel = [c for i, c in enumerate(cases) if i % len(wins) == now_off]
        for (e, now) in sel[:6]:
            S = len(texts)
            got = shim.unmet(base, np.full(S, e), now, zone_id=np.full(S, z, np.int32))
            want = po.batch_unmet(base, np.full(S, e), now, zone_id=np.full(S, z, np.int32), zone_names=NAMES)
            _cmp(got, want, (name, e, now))


def test_unmet_seconds_field_across_transitions():
    texts = ["*/20 * * * * *", "59 59 * * * *", "59 59 23 * * *", "0 30 0 * * *", "59 59 1 * * 0", "30 */15 1-3 * * *"]
    base = _masks_of(texts, with_seconds=True)
    for name in ["America/New_York", "Australia/Lord_Howe", "Africa/Cairo", "America/Havana", "Atlantic/Azores"]:
        z = NAMES.index(name)
        when, _ = shim.zone_entries(z)
        for t in [int(w) for w in when if ts(2024) < w < ts(2028)]:
            for a, b in [(-7200, 7200), (-86400, 3600), (-3, 3), (-90000, 90000)]:
                S = len(texts)
                got = shim.unmet(base, np.full(S, t + a), t + b, zone_id=np.full(S, z, np.int32))
                want = po.batch_unmet(base, np.full(S, t + a), t + b, zone_id=np.full(S, z, np.int32), zone_names=NAMES)
                _cmp(got, want, (name, t, a, b))


def test_unmet_C4_shape_mixed_zones():
    """C4-shaped sample: 30-day backlog, now inside a European DST day, mixed zones, policies, gates"""
    n = 3000
    masks = synth.gen_specs(n, seed=41, dense_minutes=0.5)
    # keep the literal oracle affordable: dense-minute specs fire in a few hours only
    dense = (masks[1] & np.uint64((1 << 60) - 1)).astype(np.uint64)
    many = np.array([bin(int(x)).count("1") for x in dense]) > 6
    masks[2] = np.where(many, masks[2] & np.uint64(0x800000000000000E), masks[2])
    masks[2] = np.where(masks[2] & np.uint64(0xFFFFFF), masks[2], np.uint64(1 << 2))
    dst = [z for z in range(1, len(NAMES)) if (shim.zone_entries(z)[0] > ts(2030)).any()]
    zid = synth.gen_zone_ids(n, len(NAMES), np.array(dst), seed=41)
    now = ts(2026, 3, 29, 12)
    earliest = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=41)
    rng = np.random.default_rng(41)
    policy = rng.integers(0, 3, n).astype(np.uint8)
    active = rng.choice([0, 1, 2, 3, 4], size=n, p=[0.6, 0.3, 0.04, 0.03, 0.03]).astype(np.uint32)
    flags = (rng.random(n) < 0.02).astype(np.uint8)
    deadline = np.where(rng.random(n) < 0.01, now - 100, po.ZERO).astype(np.int64)
    got = shim.unmet(masks, earliest, now, 0, policy, active, flags, deadline, zone_id=zid)
    want = po.batch_unmet(masks, earliest, now, 0, policy, active, flags, deadline, zone_id=zid, zone_names=NAMES)
    _cmp(got, want)


def test_upstream_hangs_are_reported_not_reproduced():
    """Two inputs on which upstream never returns (both found by tests/stress_walk.py); oracle and kernels agree on a
    terminating answer (DESIGN.md 'Deliberate divergences')."""
    # (1) robfig day loop: Kiritimati skipped 1994-12-31, AddDate(0,0,1) from Dec 30 resolves back to Dec 30
    z = NAMES.index("Pacific/Kiritimati")
    m = np.array([[0x1], [0x8FFFFFFFFFFFFFFF], [0x8000000000FFFFFF], [0x2AAAAAA], [0x8000000000001FFE], [0x49]],
                 dtype=np.uint64)
    got = shim.next_n(m, np.array([788858203]), 3, np.array([z], np.int32))
    want = po.batch_next_n(m, np.array([788858203]), 3, np.array([z], np.int32), NAMES)
    assert np.array_equal(got, want) and got[0, 0] == po.ZERO
    # (2) getNextSchedule loop: Chatham's repeated 02:45-03:45 hour sends a chained Next() back to the first 03:00
    z = NAMES.index("Pacific/Chatham")
    m = np.array([[0x1], [0x800010000100001], [0x8], [0x80000000FFFFFFFE], [0x8000000000001FFE], [0x11]], dtype=np.uint64)
    zz = np.array([z], np.int32)
    got = shim.unmet(m, np.array([1838162695]), 1839851648, zone_id=zz)
    want = po.batch_unmet(m, np.array([1838162695]), 1839851648, zone_id=zz, zone_names=NAMES)
    _cmp(got, want)
    assert got["status"][0] == po.ST_NONTERMINATING and got["decision"][0] == po.DEC_ERR_UNSCHEDULABLE


@pytest.mark.parametrize("easy_rounds", [1, 3, 5])
def test_easy_hard_split_equals_general_machine(easy_rounds):
    """unmet_kernel<easy> + unmet_hard_kernel: closed-form rounds first, the general machine (from scratch) for the
    Crons left over -- same answers as the general machine alone, which the tests above pin to the oracle"""
    n = 60000
    masks = synth.gen_specs(n, seed=41, dense_minutes=0.5)
    dst = [z for z in range(1, len(NAMES)) if (shim.zone_entries(z)[0] > ts(2030)).any()]
    zid = synth.gen_zone_ids(n, len(NAMES), np.array(dst), seed=41)
    now = ts(2026, 3, 29, 12)
    earliest = now - 30 * 86400 + synth.gen_times(n, -86400, 86400, seed=41)
    earliest[::97] = now + 5  # earliest after now: straight to Next(now)
    rng = np.random.default_rng(5)
    policy = rng.integers(0, 3, n).astype(np.uint8)
    active = rng.integers(0, 3, n).astype(np.uint32)
    flags = (rng.random(n) < 0.03).astype(np.uint8) * rng.integers(1, 3, n).astype(np.uint8)
    deadline = np.where(rng.random(n) < 0.01, now - 100, po.ZERO).astype(np.int64)
    want = shim.unmet(masks, earliest, now, 0, policy, active, flags, deadline, zone_id=zid)
    got = shim.unmet_two_phase(masks, earliest, now, easy_rounds, 0, policy, active, flags, deadline, zone_id=zid)
    _cmp(got, want)
    frac = got["deferred"].mean()
    assert 0 < frac < (1.0 if easy_rounds == 1 else 0.2), frac
    # transition windows of the edge zones, and the cycle upstream hangs on
    texts = ["*/20 * * * *", "30 2 * * *", "0 0 * * *", "59 23 * * *", "0 0 1 * *", "*/30 0-4 * * *"]
    base = _masks_of(texts)
    for name in EDGE_ZONES:
        z = NAMES.index(name)
        when, _ = shim.zone_entries(z)
        for t in [int(w) for w in when if ts(2021) < w < ts(2041)][:8]:
            for a, b in [(-3 * 86400, 3 * 86400), (-40 * 86400, 1), (-7200, 7200), (-20 * 86400, 20 * 86400)]:
                zz = np.full(len(texts), z, np.int32)
                _cmp(shim.unmet_two_phase(base, np.full(len(texts), t + a), t + b, easy_rounds, zone_id=zz),
                     shim.unmet(base, np.full(len(texts), t + a), t + b, zone_id=zz), (name, t, a, b))
    z = NAMES.index("Pacific/Chatham")
    m = np.array([[0x1], [0x800010000100001], [0x8], [0x80000000FFFFFFFE], [0x8000000000001FFE], [0x11]], dtype=np.uint64)
    zz = np.array([z], np.int32)
    got = shim.unmet_two_phase(m, np.array([1838162695]), 1839851648, easy_rounds, zone_id=zz)
    assert got["status"][0] == po.ST_NONTERMINATING and got["deferred"][0] == 1


def test_every_second_specs_around_transitions_fast_forward_by_the_hour():
    """specs that can fire at :59:59 and exclude some hour (the only ones whose chained calls can hit a Date() hour
    reset): near a transition the fast-forward works hour by hour instead of giving up; same answers as the literal
    chain, and the general machine needs few steps (it used to take one step per fire)"""
    texts = ["* * 0-5,9,22-23 * * *", "*/7 * 1-3 * * *", "59 59 * * * *", "* 59 0,2,4,23 * * *", "* * * * * *",
             "30-59 * 3-4 * * *", "* * 12 * * *"]
    base = _masks_of(texts, with_seconds=True)
    for name in ["Europe/Oslo", "Europe/Chisinau", "America/New_York", "Australia/Lord_Howe", "Asia/Tehran",
                 "America/Havana", "Pacific/Chatham"]:
        z = NAMES.index(name)
        when = shim.zone_entries(z)[0]
        for t in [int(w) for w in when if ts(2024) < w < ts(2028)][:4]:
            for a, b in [(-6 * 3600, 5 * 3600), (-86400, 11 * 3600), (-2 * 86400, -1), (-100, 7200), (0, 3 * 3600)]:
                S = len(texts)
                zz = np.full(S, z, np.int32)
                got = shim.unmet_two_phase(base, np.full(S, t + a), t + b, 0, zone_id=zz)
                want = po.batch_unmet(base, np.full(S, t + a), t + b, zone_id=zz, zone_names=NAMES)
                _cmp(got, want, (name, t, a, b))
                ok = got["status"] == 0
                assert got["rounds"][ok].max() < 120, (name, t, a, b, got["rounds"])


def test_utc_machine_skips_the_final_call_exactly():
    """unmet_step_utc answers Next(now) with the Next() that ended the chain; pin that against the literal restatement
    where the two calls have different five-year windows (Feb 29 across 2100, year ends) and on random specs"""
    texts = ["0 0 29 2 *", "0 0 31 12 *", "59 23 31 12 *", "0 0 1 1 *", "*/5 * * * *", "0 0 29 2 1", "0 12 30 4,6 *"]
    base = _masks_of(texts)
    S = len(texts)
    for now in [ts(2099, 6, 1), ts(2100, 3, 1), ts(2101, 3, 1), ts(2104, 2, 29), ts(2096, 2, 29) - 1, ts(2096, 2, 29),
                ts(2031, 1, 1), ts(2030, 12, 31, 23, 59, 59), ts(2026, 1, 1), ts(2025, 12, 31, 23, 59)]:
        for back in [1, 59, 86400, 400 * 86400, 5 * 366 * 86400, 9 * 366 * 86400]:
            got = shim.unmet(base, np.full(S, now - back), now)
            want = po.batch_unmet(base, np.full(S, now - back), now)
            _cmp(got, want, (now, back))
    n = 20000
    masks = synth.gen_specs(n, seed=5)
    now = ts(2027, 12, 31, 23, 59, 30)
    earliest = now - synth.gen_times(n, 0, 40 * 86400, seed=6)
    _cmp(shim.unmet(masks, earliest, now), po.batch_unmet(masks, earliest, now))


def every_batch(n=4000, seed=123):
    """a mixed batch: SpecSchedules and ConstantDelaySchedules ("@every", EVERY_TAG encoding) side by side, with the
    expected answers for the latter computed here in numpy (constantdelay.go: Next(t) = t + delay)"""
    rng = np.random.default_rng(seed)
    masks = synth.gen_specs(n, seed=seed, dense_minutes=0.3)
    is_every = rng.random(n) < 0.4
    delay = rng.choice([1, 7, 60, 90, 3600, 5400, 86400, 40 * 86400, 9_000_000_000], size=n).astype(np.int64)
    masks[:, is_every] = 0
    masks[3, is_every] = delay[is_every].astype(np.uint64)
    masks[4, is_every] = np.uint64(1 << 62)
    now = ts(2026, 3, 29, 12)
    earliest = now - synth.gen_times(n, -3600, 3 * 86400, seed=seed)  # some after now
    return masks, is_every, delay, earliest, now


def test_every_schedules_in_mixed_batches():
    masks, is_every, delay, earliest, now = every_batch()
    n = masks.shape[1]
    zid = np.where(np.arange(n) % 3 == 0, NAMES.index("Europe/Berlin"), 0).astype(np.int32)  # ignored by @every
    for k in (1, 5):
        got = shim.next_n(masks, earliest, k, zid)
        want = po.batch_next_n(masks, earliest, k, zid, NAMES)
        assert np.array_equal(got, want)
        assert np.array_equal(got[is_every], earliest[is_every, None] + delay[is_every, None] * np.arange(1, k + 1))
    rng = np.random.default_rng(1)
    policy, active = rng.integers(0, 3, n).astype(np.uint8), rng.integers(0, 3, n).astype(np.uint32)
    flags = (rng.random(n) < 0.05).astype(np.uint8)
    got = shim.unmet(masks, earliest, now, 0, policy, active, flags, None, zid)
    want = po.batch_unmet(masks, earliest, now, 0, policy, active, flags, None, zid, NAMES)
    _cmp(got, want)
    e = is_every & (flags == 0)
    after = earliest > now
    k = np.where(after, 0, (now - earliest) // delay)
    assert np.array_equal(got["missed"][e], k[e].astype(np.uint32))
    assert np.array_equal(got["next"][e], np.full(e.sum(), now) + delay[e])
    assert np.array_equal(got["last"][e], np.where(k[e] > 0, earliest[e] + k[e] * delay[e], po.ZERO))
    assert np.array_equal(got["status"][e], np.where(after[e], 2, 0))
    # the parsers hand out the same encoding
    for text, d in (("@every 90s", 90), ("@every 1h30m", 5400), ("@every 500ms", 1), ("TZ=Asia/Tokyo @every 2h", 7200)):
        s, kind, every, _ = po.parse_standard(text)
        assert (kind, every) == (1, d) and s.tuple() == (0, 0, 0, d, 1 << 62, 0)


def test_period_clock_machine_matches_literal_chain():
    """class A of the unmet kernels: a zoned Cron whose whole window lies in one zone period is answered by the UTC
    machine on that period's local clock, and rejected into the general machine when a Next() leaves the period; both
    must give the literal chain's answers.  Mixed zones (fixed-offset, 30/45-minute, DST on both hemispheres), windows
    that end days, hours or seconds before / after a transition, sparse specs whose next fire is months away."""
    rng = np.random.default_rng(77)
    zones = ["Asia/Tokyo", "Asia/Kolkata", "Asia/Kathmandu", "Europe/Berlin", "America/New_York", "Australia/Sydney",
             "America/Santiago", "Pacific/Chatham", "Africa/Cairo", "Atlantic/Azores", "Asia/Tehran", "America/Sao_Paulo",
             "Australia/Lord_Howe", "UTC", "Etc/GMT+12", "Pacific/Apia"]
    n = 40_000
    masks = synth.gen_specs(n, seed=78, dense_minutes=0.3)
    zid = np.array([NAMES.index(z) for z in zones], np.int32)[rng.integers(0, len(zones), n)]
    klass = np.zeros(n, np.uint8)
    seen = np.zeros(4, np.int64)
    for now in (ts(2026, 3, 29, 12), ts(2026, 7, 15, 3, 7, 9), ts(2026, 10, 25, 0, 59, 59), ts(2026, 11, 1, 6), ts(2027, 1, 1)):
        earliest = now - synth.gen_times(n, -7200, 20 * 86400, seed=int(now % 1000))
        got = shim.unmet(masks, earliest, now, zone_id=zid, klass_out=klass)
        want = po.batch_unmet(masks, earliest, now, zone_id=zid, zone_names=NAMES)
        _cmp(got, want, now)
        seen += np.bincount(klass, minlength=4)
    assert seen[1] > n and seen[2] > 1000 and seen[3] > 100, seen  # every path was exercised
    # windows placed right against transitions of a DST zone
    z = NAMES.index("Europe/Berlin")
    when = shim.zone_entries(z)[0]
    texts = ["* * * * *", "*/5 * * * *", "30 2 * * *", "0 0 * * *", "0 0 1 * *", "15 3 * * 0", "59 23 31 12 *", "0 12 1 1 *"]
    base = _masks_of(texts)
    S = len(texts)
    for t in [int(w) for w in when if ts(2024) < w < ts(2028)]:
        for back, ahead in [(40 * 86400, -33 * 86400), (40 * 86400, -1), (86400, -1), (3600, 3600), (10, 40 * 86400), (0, 86400)]:
            zz = np.full(S, z, np.int32)
            e, now = np.full(S, t - back), t + ahead
            if now < t - back:
                continue
            _cmp(shim.unmet(base, e, now, zone_id=zz), po.batch_unmet(base, e, now, zone_id=zz, zone_names=NAMES), (t, back, ahead))
