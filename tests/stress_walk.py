"""Heavy randomized stress of the kernels' walk source (host build) vs the oracle.  Not collected by pytest (no
test_ prefix); run by hand:  python tests/stress_walk.py [rounds [seed]]  (seed 0 = the sequence the round logs quote)"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from cron_operator_b200 import synth  # noqa: E402
from oracle import pyoracle as po  # noqa: E402
from tests import shim  # noqa: E402

ts = synth.ts
NAMES = shim.zone_names()


def near_transition_times(rng, zid, lo, hi, spread):
    out = np.empty(len(zid), dtype=np.int64)
    cache = {}
    for i, z in enumerate(zid):
        if z not in cache:
            w = shim.zone_entries(int(z))[0]
            cache[z] = w[(w > lo) & (w < hi)]
        w = cache[z]
        if len(w) == 0:
            out[i] = rng.integers(lo, hi)
        else:
            out[i] = int(rng.choice(w)) + int(rng.integers(-spread, spread))
    return out


def main(rounds, salt=0):
    rng = np.random.default_rng(12345 + salt)
    tot = 0
    t0 = time.time()
    for r in range(rounds):
        n = 40000
        seed = 1000 + r + 100003 * salt
        ws = bool(r % 2)
        masks = synth.gen_specs(n, seed=seed, with_seconds=ws, dense_minutes=0.3 if r % 3 == 0 else 0.0)
        zid = rng.integers(1, len(NAMES), n).astype(np.int32)
        spread = [3 * 3600, 86400 * 2, 86400 * 40, 400 * 86400][r % 4]
        frm = near_transition_times(rng, zid, ts(1968), ts(2046), spread)
        k = 6
        got = shim.next_n(masks, frm, k, zid)
        want = po.batch_next_n(masks, frm, k, zid, NAMES)
        bad = np.nonzero((got != want).any(axis=1))[0]
        tot += n * k
        if len(bad):
            b = bad[0]
            print("MISMATCH next", NAMES[zid[b]], int(frm[b]), [hex(int(x)) for x in masks[:, b]], got[b], want[b])
            return 1
        # unmet: windows around transitions (keep the literal oracle affordable: sparse hours for dense minutes)
        m2 = masks.copy()
        pc = lambda a: np.array([bin(int(x) & ((1 << 60) - 1)).count("1") for x in a])  # noqa: E731
        sub = slice(0, 4000)
        m2 = np.ascontiguousarray(m2[:, sub])
        # bound the literal oracle: <= 3 hours x <= 4 minutes x <= 2 seconds per day
        m2[2] = np.where(pc(m2[2]) > 3, m2[2] & np.uint64(0x800000000000000E), m2[2])
        m2[2] = np.where(m2[2] & np.uint64(0xFFFFFF), m2[2], np.uint64(1 << 2))
        m2[1] = np.where(pc(m2[1]) > 4, np.uint64((1 << 0) | (1 << 20) | (1 << 40) | (1 << 59)), m2[1])
        m2[0] = np.where(pc(m2[0]) > 2, np.uint64((1 << 0) | (1 << 59)), m2[0])
        now = int(rng.integers(ts(2020), ts(2030)))
        zz = zid[sub]
        e = near_transition_times(rng, zz, now - 20 * 86400, now, 86400 * 3)
        print(f"round {r}: next ok, unmet...", flush=True)
        g = shim.unmet(m2, e, now, zone_id=zz)
        w = po.batch_unmet(m2, e, now, zone_id=zz, zone_names=NAMES)
        for key in ("last", "next", "missed", "status", "decision"):
            bad = np.nonzero(g[key] != w[key])[0]
            if len(bad):
                b = bad[0]
                print("MISMATCH unmet", key, NAMES[zz[b]], int(e[b]), now, [hex(int(x)) for x in m2[:, b]], g[key][b], w[key][b])
                return 1
        # the kernels' easy/hard split (unmet_kernel<easy> + unmet_hard_kernel) over the full-density masks: the general
        # machine (pinned to the oracle just above on the thinned masks) is the reference here
        for er in (2, 5):
            full = np.ascontiguousarray(masks[:, :20000])
            e2 = near_transition_times(rng, zid[:20000], now - 20 * 86400, now, 86400 * 3)
            g1 = shim.unmet(full, e2, now, zone_id=zid[:20000])
            g2 = shim.unmet_two_phase(full, e2, now, er, zone_id=zid[:20000])
            for key in ("last", "next", "missed", "status", "decision"):
                bad = np.nonzero(g1[key] != g2[key])[0]
                if len(bad):
                    b = bad[0]
                    print("MISMATCH two-phase", er, key, NAMES[zid[b]], int(e2[b]), now, [hex(int(x)) for x in full[:, b]])
                    return 1
        print(f"round {r}: ok ({tot} evals, {time.time() - t0:.0f}s)", flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main(int(sys.argv[1]) if len(sys.argv) > 1 else 8, int(sys.argv[2]) if len(sys.argv) > 2 else 0))
