"""An independent statement of what Next() must return around ordinary DST transitions, against the oracle and the
kernels' walk source (host build).

robfig/cron's SpecSchedule.Next (called at internal/controller/cron_controller.go:405,409,436) walks local civil
fields.  Around a transition that moves the clock by a whole hour, at a whole local hour, away from midnight (01:00-04:00
on both sides: every daylight-saving rule in force today), its answer is simply the EARLIEST instant after t whose
local fields match the masks -- a statement that needs no knowledge of the algorithm.  The checker below finds that
instant by scanning minute by minute with python's zoneinfo (an independent TZif reader), over every such transition
of every zone between 1980 and 2037.

Transitions that touch local midnight, shift by 30/45 minutes or happen off the hour are different: there upstream's
loops (no re-check of the hour after the minute loop, a single Truncate) return instants a scan would not.  Six of
those were traced by hand through upstream's source and are kept as known answers (`QUIRKS`), so the literal behaviour
stays pinned as well."""
import datetime as dt
import random
import zoneinfo

import numpy as np
import pytest

from cron_operator_b200.synth import ts
from oracle import pyoracle as po
from tests import shim

NAMES = shim.zone_names()
SPECS = ["* * * * *", "*/5 * * * *", "59 23 * * *", "0 0 * * *", "30 2 * * *", "30 1 * * *", "0 3 * * *", "15 0 * * *",
         "0 */2 * * *", "*/20 1-3 * * *", "0 0 * * 0", "0 0 1 * *", "45 23 * * 6", "0 12 * * *", "30 0 * * *", "0 1 * * *",
         "59 0-3 * * *", "0 2-3 * * *", "1 0 * * *", "0 23 * * *", "10,50 */3 * * *", "0 2 * * 0", "30 1-2 1-7 * *",
         "*/15 2 * 3,10,11 0", "0 1,3 * * 1-5", "29,31 1-3 * * *"]


def _matches(m, d):
    if not (m[0] >> d.second & 1 and m[1] >> d.minute & 1 and m[2] >> d.hour & 1 and m[4] >> d.month & 1):
        return False
    dom_ok, dow_ok = bool(m[3] >> d.day & 1), bool(m[5] >> ((d.weekday() + 1) % 7) & 1)
    return (dom_ok and dow_ok) if ((m[3] | m[5]) >> 63) else (dom_ok or dow_ok)


def _ordinary_transitions():
    out = []
    for zi in range(1, len(NAMES)):
        when, off = shim.zone_entries(zi)
        for j in range(1, len(when)):
            if not (ts(1980) < when[j] < ts(2037)):
                continue
            before, after = int(off[j - 1]), int(off[j])
            if abs(after - before) != 3600 or before % 3600:
                continue
            wall = (int(when[j]) + before) % 86400  # the clock reading at which it jumps
            if wall % 3600 == 0 and 3600 <= wall <= 4 * 3600 and 3600 <= wall + after - before <= 4 * 3600:
                out.append((zi, int(when[j])))
    return out


def _first_match_by_scan(m, z, t, stop):
    cur = (t // 60 + 1) * 60  # 5-field schedules fire at second 0
    while cur < stop:
        if _matches(m, dt.datetime.fromtimestamp(cur, z)):
            return cur
        cur += 60
    return None


def test_next_is_the_earliest_matching_instant_around_ordinary_dst_transitions():
    cands = _ordinary_transitions()
    assert len(cands) > 10000
    rng = random.Random(2026)
    horizon = 10 * 86400
    cases = []
    for _ in range(2500):
        zi, tr = rng.choice(cands)
        cases.append((zi, tr + rng.randrange(-2 * 86400, 86400), rng.choice(SPECS)))
    # the kernels' walk source on the same cases, three chained calls each
    parsed = {text: po.parse_standard(text)[0] for text in SPECS}
    masks = np.ascontiguousarray(np.array([parsed[c[2]].tuple() for c in cases], dtype=np.uint64).T)
    frm = np.array([c[1] for c in cases], dtype=np.int64)
    zid = np.array([c[0] for c in cases], dtype=np.int32)
    walk = shim.next_n(masks, frm, 3, zid)
    checked = 0
    for i, (zi, t, text) in enumerate(cases):
        try:
            z = zoneinfo.ZoneInfo(NAMES[zi])
        except Exception:  # a zone python's tzdata does not carry
            continue
        spec, m, loc = parsed[text], parsed[text].tuple(), po.load_location(NAMES[zi])
        cur = t
        for j in range(3):
            r = po.spec_next(spec, loc, cur)
            assert r == walk[i, j], (NAMES[zi], text, t, j)
            assert r != po.ZERO and r > cur
            stop = min(r, cur + horizon)
            assert _first_match_by_scan(m, z, cur, stop) is None, (NAMES[zi], text, cur, r)  # nothing earlier matches
            if r <= cur + horizon:
                assert _matches(m, dt.datetime.fromtimestamp(r, z)), (NAMES[zi], text, cur, r)
            cur = r
        checked += 1
    assert checked > 2000


def _local(name, *ymdhms):
    """Unix time of an unambiguous local civil time"""
    return int(dt.datetime(*ymdhms, tzinfo=zoneinfo.ZoneInfo(name)).timestamp())


# (zone, spec, from (local), want (local), why) -- [UPSTREAM-TRACED]: hand-traced through robfig/cron v3.0.1 spec.go
QUIRKS = [
    # 1982-01-01 00:00 +07:30 became 00:30 +08:00.  The hour loop's Add(1h) from 23:00 lands on 00:30 (hour 0: WRAP),
    # the next Add(1h) gives 01:30; the minute loop then runs to 02:00 and wraps: Jan 1 01:00 is never visited.
    ("Asia/Singapore", "0 1 * * *", (1981, 12, 31, 1, 5, 18), (1982, 1, 2, 1, 0, 0)),
    # 2021-10-03 00:01 +08 became 03:01 +11.  Hour 0 matches at 00:00 +08; the minute loop crosses the jump and stops
    # at minute 30 without looking at the hour again: 03:30 +11.
    ("Antarctica/Casey", "30 0 * * *", (2021, 10, 2, 14, 13, 44), (2021, 10, 3, 3, 30, 0)),
    # 2000-10-29 02:30 +06 became 01:30 +05: hour 2 matches at 02:00 +06, the minute loop reaches minute 30 at the
    # transition instant, which reads 01:30 +05.
    ("Asia/Bishkek", "30 2 * * *", (2000, 10, 28, 19, 34, 45), (2000, 10, 29, 1, 30, 0)),
    # 1983-09-25 01:00 +01 became 00:00 +00: midnight happens twice.  The day loop's Date(y, m, d+1, 0, ...) looks the
    # zone up at "00:00 as if UTC", which is the transition instant itself: offset 0, so the SECOND midnight.
    ("Europe/Lisbon", "0 0 * * 0", (1983, 9, 24, 1, 53, 17), (1983, 9, 25, 0, 0, 0)),
    # 1998-04-27 00:00 -08:30 became 00:30 -08:00.  From Sunday Apr 26 the hour loop wraps at 00:30; AddDate then keeps
    # 00:30 through the week (the single Truncate is spent), Sunday May 3 is entered at 00:30, its minute loop wraps at
    # 01:00, the hour loop runs to Monday 00:00, and only May 10 is reached on the hour.
    ("Pacific/Pitcairn", "0 0 * * 0", (1998, 4, 26, 9, 37, 14), (1998, 5, 10, 0, 0, 0)),
    # 2015-08-15 00:00 +09 became 23:30 +08:30 (Aug 14).  23:00 + 1h reads 23:30, + 1h reads 00:30 (WRAP), + 1h reads
    # 01:30: the hour matches off the hour, the minute loop wraps at 02:00, and Aug 15 01:00 is never visited.
    ("Asia/Pyongyang", "0 1 * * *", (2015, 8, 14, 4, 24, 44), (2015, 8, 16, 1, 0, 0)),
]


@pytest.mark.parametrize("zone,text,frm,want", QUIRKS)
def test_upstream_literal_behaviour_at_odd_transitions(zone, text, frm, want):
    if zone not in NAMES:
        pytest.skip("zone not in this box's tzdata")
    spec, *_ = po.parse_standard(text)
    t = _local(zone, *frm)
    z = zoneinfo.ZoneInfo(zone)
    # the wanted instant is written in post-transition local time; fold=1 picks the later of two readings
    w = int(dt.datetime(*want, tzinfo=z, fold=1).timestamp())
    assert po.spec_next(spec, po.load_location(zone), t) == w
    m = np.ascontiguousarray(np.array([spec.tuple()], dtype=np.uint64).T)
    got = shim.next_n(m, np.array([t], np.int64), 1, np.array([NAMES.index(zone)], np.int32))
    assert int(got[0, 0]) == w


def test_get_next_schedule_matches_an_algorithm_free_scan():
    """getNextSchedule (internal/controller/cron_controller.go:389-437): lastMissed, Next(now) and the number of missed
    runs, against a scan that simply lists the matching minutes of (earliest, now] -- UTC and windows holding an
    ordinary DST transition; oracle and the kernels' classed unmet machine (host build)"""
    cands = _ordinary_transitions()
    rng = random.Random(405)
    parsed = {text: po.parse_standard(text)[0] for text in SPECS}
    n = 400
    zid, texts, earliest = [], [], []
    # one `now` per call (as one reconcile tick has): take a transition, then Crons in zones that switch near it
    zi0, tr0 = rng.choice([c for c in cands if ts(2015) < c[1] < ts(2030)])
    now = tr0 + rng.randrange(3600, 2 * 86400)
    near = [c for c in cands if abs(c[1] - tr0) < 3 * 86400] or [(zi0, tr0)]
    for i in range(n):
        zid.append(0 if i % 4 == 0 else rng.choice(near)[0])
        texts.append(rng.choice(SPECS))
        earliest.append(now - rng.randrange(0, 4 * 86400))
    masks = np.ascontiguousarray(np.array([parsed[t].tuple() for t in texts], dtype=np.uint64).T)
    zid = np.array(zid, dtype=np.int32)
    earliest = np.array(earliest, dtype=np.int64)
    want = po.batch_unmet(masks, earliest, now, zone_id=zid, zone_names=NAMES)
    got = shim.unmet(masks, earliest, now, zone_id=zid)
    for key in ("last", "next", "missed", "status", "decision"):
        assert np.array_equal(got[key], want[key]), key
    checked = 0
    for i in range(n):
        try:
            z = dt.timezone.utc if zid[i] == 0 else zoneinfo.ZoneInfo(NAMES[zid[i]])
        except Exception:
            continue
        m = parsed[texts[i]].tuple()
        fires = []
        cur = (int(earliest[i]) // 60 + 1) * 60
        while cur <= now:
            if _matches(m, dt.datetime.fromtimestamp(cur, z)):
                fires.append(cur)
            cur += 60
        nxt = _first_match_by_scan(m, z, now, now + 40 * 86400)
        assert int(want["missed"][i]) == len(fires), (NAMES[zid[i]], texts[i], int(earliest[i]), now)
        assert int(want["last"][i]) == (fires[-1] if fires else po.ZERO)
        assert nxt is not None and int(want["next"][i]) == nxt
        checked += 1
    assert checked > 350 and int(want["missed"].max()) > 100  # the TooManyMissedTimes range is in the sample


def test_seconds_field_next_around_ordinary_dst_transitions_by_second_scan():
    """the seconds extension (6-field specs, BASELINE config C3) on the same footing: second-by-second scan"""
    cands = _ordinary_transitions()
    rng = random.Random(99)
    texts6 = ["*/10 * * * * *", "5,35 */2 1-3 * * *", "59 59 * * * *", "0 30 1-2 * * *", "13 * 2 * * *", "*/7 */9 * * * *",
              "30 29,31 1-3 * * 0,6", "1 0 */2 * * *"]
    parsed = {t: po.parse_standard(t, True)[0] for t in texts6}
    cases = []
    for _ in range(160):
        zi, tr = rng.choice(cands)
        cases.append((zi, tr + rng.randrange(-5 * 3600, 3 * 3600), rng.choice(texts6)))
    masks = np.ascontiguousarray(np.array([parsed[c[2]].tuple() for c in cases], dtype=np.uint64).T)
    walk = shim.next_n(masks, np.array([c[1] for c in cases], np.int64), 2, np.array([c[0] for c in cases], np.int32))
    checked = 0
    for i, (zi, t, text) in enumerate(cases):
        try:
            z = zoneinfo.ZoneInfo(NAMES[zi])
        except Exception:
            continue
        spec, m, loc = parsed[text], parsed[text].tuple(), po.load_location(NAMES[zi])
        cur = t
        for j in range(2):
            r = po.spec_next(spec, loc, cur)
            assert r == walk[i, j] and r != po.ZERO and r > cur
            stop = min(r, cur + 8 * 3600)
            s = cur + 1
            while s < stop:
                assert not _matches(m, dt.datetime.fromtimestamp(s, z)), (NAMES[zi], text, cur, s, r)
                s += 1
            if r <= cur + 8 * 3600:
                assert _matches(m, dt.datetime.fromtimestamp(r, z)), (NAMES[zi], text, cur, r)
            cur = r
        checked += 1
    assert checked > 120
