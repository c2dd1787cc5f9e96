"""Synthetic workloads of BASELINE.json configs C2-C5, generated with a counter-based RNG:
every draw is splitmix64(seed ^ f(index, field, draw)), so any index can be regenerated independently (CPU oracle
sampling, per-rank shards) and a shard [lo, hi) of a big workload equals the same slice of the whole.

Masks are produced directly in robfig/cron's SpecSchedule layout (bit i = value i allowed, bit 63 = starBit, set
exactly when ParseStandard would set it: `*`, `?` and `*/1`; cleared by a step > 1).
"""
from __future__ import annotations

import numpy as np

STAR = np.uint64(1 << 63)
U64 = np.uint64

# (lo, hi) per field in plane order: second, minute, hour, dom, month, dow
BOUNDS = ((0, 59), (0, 59), (0, 23), (1, 31), (1, 12), (0, 6))


def splitmix64(x: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        z = (x + U64(0x9E3779B97F4A7C15)).astype(np.uint64)
        z = (z ^ (z >> U64(30))) * U64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> U64(27))) * U64(0x94D049BB133111EB)
        return z ^ (z >> U64(31))


def _draw(seed: int, idx: np.ndarray, stream: int) -> np.ndarray:
    """one uint64 per index for the given stream id"""
    with np.errstate(over="ignore"):
        return splitmix64(splitmix64(idx ^ U64(seed)) + U64(stream) * U64(0xD1342543DE82EF95))


def _uniform_int(r: np.ndarray, lo, hi) -> np.ndarray:
    """integers in [lo, hi] (hi may be an array); slight modulo bias is irrelevant here"""
    span = (np.asarray(hi, dtype=np.int64) - np.asarray(lo, dtype=np.int64) + 1).astype(np.uint64)
    return (r % span).astype(np.int64) + lo


def form_thresholds(p_star: float) -> np.ndarray:
    """the 5 cumulative thresholds that split u in [0,1) into the 6 field forms (also handed to the device generator,
    so that both sides compare against the very same doubles)"""
    rest = 1.0 - p_star
    cuts = p_star + rest * np.array([0.25, 0.25 + 0.3333, 0.25 + 0.3333 + 0.1667, 0.25 + 0.3333 + 0.1667 + 0.0833])
    return np.concatenate(([p_star], cuts)).astype(np.float64)


def gen_field(seed: int, idx: np.ndarray, field: int, p_star: float = 0.40) -> np.ndarray:
    """One mask per index.  Forms: `*` p_star, `*/k` 15%, single 20%, `a-b` 10%, `a-b/k` 5%, list of 2-4 rest."""
    lo, hi = BOUNDS[field]
    n = idx.shape[0]
    base = field * 16
    u = (_draw(seed, idx, base + 0) >> U64(11)).astype(np.float64) / float(1 << 53)
    form = np.searchsorted(form_thresholds(p_star), u, side="right")  # 0 star,1 */k,2 single,3 a-b,4 a-b/k,5 list
    span = hi - lo + 1
    kmax = max(2, span // 2)
    k = _uniform_int(_draw(seed, idx, base + 1), 2, kmax)
    a = _uniform_int(_draw(seed, idx, base + 2), lo, hi)
    b = _uniform_int(_draw(seed, idx, base + 3), a, np.full(n, hi))
    start = np.where((form == 0) | (form == 1), lo, a)
    end = np.where((form == 0) | (form == 1), hi, np.where(form == 2, a, b))
    step = np.where((form == 1) | (form == 4), k, 1)
    bits = np.zeros(n, dtype=np.uint64)
    for v in range(lo, hi + 1):
        on = (v >= start) & (v <= end) & (((v - start) % step) == 0) & (form != 5)
        bits |= np.where(on, U64(1 << v), U64(0))
    # list of 2-4 singles
    cnt = _uniform_int(_draw(seed, idx, base + 4), 2, 4)
    lst = np.zeros(n, dtype=np.uint64)
    for j in range(4):
        v = _uniform_int(_draw(seed, idx, base + 5 + j), lo, hi).astype(np.uint64)
        lst |= np.where(j < cnt, U64(1) << v, U64(0))
    bits = np.where(form == 5, lst, bits)
    bits |= np.where(form == 0, STAR, U64(0))
    return bits


def gen_specs(n: int, seed: int = 0xC0FFEE, start: int = 0, with_seconds: bool = False, dense_minutes: float = 0.0,
              plant_unschedulable: float = 0.001, idx: np.ndarray | None = None) -> np.ndarray:
    """(6, n) uint64 SoA mask planes for spec indices [start, start + n), or for the explicit index array `idx`."""
    if idx is None:
        idx = np.arange(start, start + n, dtype=np.uint64)
    idx = np.asarray(idx, dtype=np.uint64)
    n = len(idx)
    m = np.empty((6, n), dtype=np.uint64)
    m[0] = gen_field(seed, idx, 0) if with_seconds else U64(1)
    m[1] = gen_field(seed, idx, 1)
    m[2] = gen_field(seed, idx, 2)
    m[3] = gen_field(seed, idx, 3, p_star=0.60)
    m[4] = gen_field(seed, idx, 4)
    m[5] = gen_field(seed, idx, 5, p_star=0.60)
    if dense_minutes > 0:  # C4: minute field forced to `*` or `*/k` in this fraction of specs
        u = (_draw(seed, idx, 100) >> U64(11)).astype(np.float64) / float(1 << 53)
        k = _uniform_int(_draw(seed, idx, 101), 1, 6)
        dense = np.zeros(n, dtype=np.uint64)
        for v in range(60):
            dense |= np.where((v % k) == 0, U64(1 << v), U64(0))
        dense |= np.where(k == 1, STAR, U64(0))
        m[1] = np.where(u < dense_minutes, dense, m[1])
    if plant_unschedulable > 0:
        u = (_draw(seed, idx, 102) >> U64(11)).astype(np.float64) / float(1 << 53)
        which = _draw(seed, idx, 103) & U64(1)
        planted = u < plant_unschedulable
        full = lambda f: U64(sum(1 << v for v in range(BOUNDS[f][0], BOUNDS[f][1] + 1))) | STAR  # noqa: E731
        # "0 0 30 2 *"  and  "0 0 31 4,6,9,11 *"
        m[0] = np.where(planted, U64(1), m[0])
        m[1] = np.where(planted, U64(1), m[1])
        m[2] = np.where(planted, U64(1), m[2])
        m[3] = np.where(planted, np.where(which == 0, U64(1 << 30), U64(1 << 31)), m[3])
        m[4] = np.where(planted, np.where(which == 0, U64(1 << 2), U64((1 << 4) | (1 << 6) | (1 << 9) | (1 << 11))), m[4])
        m[5] = np.where(planted, full(5), m[5])
    return np.ascontiguousarray(m)


def gen_times(n: int, lo: int, hi: int, seed: int = 0xC0FFEE, start: int = 0, stream: int = 200,
              idx: np.ndarray | None = None) -> np.ndarray:
    if idx is None:
        idx = np.arange(start, start + n, dtype=np.uint64)
    idx = np.asarray(idx, dtype=np.uint64)
    n = len(idx)
    return _uniform_int(_draw(seed, idx, stream), lo, np.full(n, hi - 1)).astype(np.int64)


def zone_pool(nzones: int, dst_zone_ids: np.ndarray) -> np.ndarray:
    return np.concatenate([np.arange(1, nzones, dtype=np.int32), np.asarray(dst_zone_ids, dtype=np.int32)])


def gen_zone_ids(n: int, nzones: int, dst_zone_ids: np.ndarray, seed: int = 0xC0FFEE, start: int = 0,
                 p_utc: float = 0.5, idx: np.ndarray | None = None) -> np.ndarray:
    """C3: 50% UTC (id 0), the rest uniform over zones 1..nzones-1 with DST-rule zones double-weighted."""
    if idx is None:
        idx = np.arange(start, start + n, dtype=np.uint64)
    idx = np.asarray(idx, dtype=np.uint64)
    u = (_draw(seed, idx, 300) >> U64(11)).astype(np.float64) / float(1 << 53)
    pool = zone_pool(nzones, dst_zone_ids)
    pick = pool[(_draw(seed, idx, 301) % U64(len(pool))).astype(np.int64)]
    return np.where(u < p_utc, 0, pick).astype(np.int32)


def ts(y, mo=1, d=1, h=0, mi=0, s=0) -> int:
    import datetime as dt
    return int(dt.datetime(y, mo, d, h, mi, s, tzinfo=dt.timezone.utc).timestamp())
