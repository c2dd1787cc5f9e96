"""Generates tests/golden/*.json from the CPU oracle (oracle/cron_oracle.c).  The reference itself is Go and cannot
run on this box (no Go toolchain), so these vectors pin the ORACLE's answers -- including the reference's own
known-answer tests K1-K4 -- and let the GPU tests run without /root/reference or the zone files of this container
mattering: each zoned vector carries the flattened zone periods it was computed with.

    python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from cron_operator_b200 import synth  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

ts = synth.ts

NEXT_CASES = [
    # (spec text, with_seconds, zone, from times)
    ("*/1 * * * *", False, "UTC", [ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12)]),
    ("*/5 * * * *", False, "UTC", [ts(2026, 1, 1, 11, 55), ts(2025, 12, 31, 23, 59, 59)]),
    ("0 0 30 2 *", False, "UTC", [ts(2026, 1, 1, 11, 55)]),
    ("0 0 29 2 *", False, "UTC", [ts(2023, 3, 1), ts(2096, 3, 1), ts(2099, 1, 1)]),
    ("0 12 1,15 * 1", False, "UTC", [ts(2025, 6, 1), ts(2025, 6, 16, 12)]),
    ("0 0 13 * 5", False, "UTC", [ts(2025, 1, 1)]),
    ("59 59 23 31 12 *", True, "UTC", [ts(2025, 3, 1, 0, 0, 1)]),
    ("30 2 * * *", False, "America/New_York", [ts(2025, 3, 8, 12), ts(2025, 11, 1, 12)]),
    ("0 1 * * *", False, "America/New_York", [ts(2025, 11, 2, 4, 30)]),
    ("0 0 * * *", False, "Africa/Cairo", [ts(2025, 4, 23), ts(2025, 10, 29)]),
    ("0 0 * * *", False, "America/Santiago", [ts(2025, 9, 5), ts(2026, 4, 3)]),
    ("0 3 * * *", False, "Australia/Lord_Howe", [ts(2025, 4, 4), ts(2025, 10, 3)]),
    ("45 1-3 25-31 * 0", False, "Asia/Tehran", [ts(2021, 3, 18, 20, 30)]),
    ("0 0 1 * *", False, "Atlantic/Azores", [ts(2025, 3, 29)]),
    ("15 10 * * *", False, "Asia/Kathmandu", [ts(2025, 6, 1)]),
    ("*/10 * * * * *", True, "Europe/London", [ts(2025, 3, 30, 0, 59, 45), ts(2025, 10, 26, 0, 59, 45)]),
    ("0 0 * * *", False, "Europe/Berlin", [ts(2040, 3, 24), ts(2100, 10, 30)]),
]

UNMET_CASES = [
    # (spec, zone, earliest, now, policy, active, flags, deadline)
    ("*/1 * * * *", "UTC", ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12), 0, 0, 0, None),   # K3
    ("0 0 30 2 *", "UTC", ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12), 0, 0, 0, None),    # K2
    ("*/5 * * * *", "UTC", ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12), 1, 1, 0, None),   # C1 Forbid, one active
    ("*/5 * * * *", "UTC", ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12), 2, 3, 0, None),
    ("*/5 * * * *", "UTC", ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12), 0, 0, 1, None),
    ("*/5 * * * *", "UTC", ts(2026, 1, 1, 11, 55), ts(2026, 1, 1, 12), 0, 0, 0, ts(2026, 1, 1, 11)),
    ("*/5 * * * *", "UTC", ts(2026, 1, 1, 12, 1), ts(2026, 1, 1, 12), 0, 0, 0, None),
    ("* * * * *", "Europe/Berlin", ts(2026, 2, 27, 12), ts(2026, 3, 29, 12), 0, 0, 0, None),
    ("0 2 * * *", "America/New_York", ts(2025, 3, 1), ts(2025, 3, 20), 0, 0, 0, None),
    ("0 0 * * *", "Africa/Cairo", ts(2025, 4, 20), ts(2025, 4, 30), 0, 0, 0, None),
    ("0 0 29 2 *", "UTC", ts(2095, 1, 1), ts(2104, 3, 1), 0, 0, 0, None),
]


def zone_periods(name):
    from tests import shim
    names = shim.zone_names()
    when, off = shim.zone_entries(names.index(name))
    return [int(w) for w in when], [int(o) for o in off]


def main():
    out = {"next": [], "unmet": [], "history": []}
    for text, ws, zone, froms in NEXT_CASES:
        spec, *_ = po.parse_standard(text, ws)
        loc = None if zone == "UTC" else po.load_location(zone)
        rows = []
        for f in froms:
            t, row = f, []
            for _ in range(4):
                t = po.spec_next(spec, loc, t) if t != po.ZERO else t
                row.append(t)
            rows.append(row)
        out["next"].append(dict(spec=text, with_seconds=ws, zone=zone, masks=[str(x) for x in spec.tuple()],
                                **{"from": froms}, want=rows))
    for text, zone, e, now, pol, act, fl, dl in UNMET_CASES:
        spec, *_ = po.parse_standard(text)
        loc = None if zone == "UTC" else po.load_location(zone)
        d, st, last, nxt, missed = po.reconcile_decide(spec, loc, e, now, 0, pol, act, fl, po.ZERO if dl is None else dl)
        out["unmet"].append(dict(spec=text, zone=zone, masks=[str(x) for x in spec.tuple()], earliest=e, now=now,
                                 policy=pol, active=act, flags=fl, deadline=dl, want=dict(
                                     decision=d, status=st, last=last, next=nxt, missed=missed)))
    now = ts(2026, 1, 1, 12)
    for keys, limit in [([now - 600, now - 1200, now - 300], None),  # K4
                        ([5, 1, 4, 2, 3], 3), ([7, 7, 7, 1, 7], 2), ([], 10), ([9], 0), ([3, 2, 1], 10)]:
        perm, keep = po.history_prune(keys, limit)
        out["history"].append(dict(keys=keys, limit=limit, perm=perm.tolist(), keep=keep.tolist()))
    with open(os.path.join(HERE, "vectors.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote vectors.json:", {k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
