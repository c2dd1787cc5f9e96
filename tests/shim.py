"""ctypes access to tests/host_shim/libwalk_host.so: a TEST-ONLY g++ build of the kernels' per-spec walk
(cron_operator_b200/csrc/cron_walk.cuh, unmet_walk.cuh) and of the zone-table builder.  Lets the CPU suite check the
exact source the kernels compile against the oracle on a box without a GPU.  Not used by the product."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(ROOT, "cron_operator_b200", "csrc")
_LIB = None
_TABLES = {}


def lib():
    global _LIB
    if _LIB is None:
        # B200CRON_SHIM_FLAGS: extra -D flags for the walk, built into its own .so (test-only tuning hook)
        extra = os.environ.get("B200CRON_SHIM_FLAGS", "").split()
        tag = "".join(c if c.isalnum() else "_" for c in "".join(extra))
        so = os.path.join(HERE, "host_shim", f"libwalk_host{tag}.so")
        srcs = [os.path.join(HERE, "host_shim", "walk_host.cpp"), os.path.join(CSRC, "zone_table.cpp")]
        deps = srcs + [os.path.join(CSRC, f) for f in ("cron_walk.cuh", "unmet_walk.cuh", "parse_kernels.cuh", "zone_table.h")]
        if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
            subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I" + CSRC] + extra + ["-o", so] + srcs)
        L = C.CDLL(so)
        L.shim_table_build.restype = C.c_void_p
        L.shim_table_build.argtypes = [C.c_char_p, C.c_int]
        L.shim_zone_count.argtypes = [C.c_void_p]
        L.shim_zone_name.restype = C.c_char_p
        L.shim_zone_name.argtypes = [C.c_void_p, C.c_int]
        L.shim_zone_find.argtypes = [C.c_void_p, C.c_char_p]
        L.shim_zone_entries.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]
        L.shim_total_entries.restype = C.c_int64
        L.shim_total_entries.argtypes = [C.c_void_p]
        L.shim_next_n.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
        L.shim_unmet.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p, C.c_void_p]
        L.shim_set_klass_out.argtypes = [C.c_void_p]
        L.shim_unmet_two_phase.argtypes = ([C.c_void_p] * 4 + [C.c_int64, C.c_int32] + [C.c_void_p] * 4 +
                                           [C.c_size_t, C.c_int] + [C.c_void_p] * 7)
        L.shim_parse_batch.argtypes = [C.c_char_p, C.c_void_p, C.c_size_t, C.c_int] + [C.c_void_p] * 4
        _LIB = L
    return _LIB


def table(zoneinfo_dir="/usr/share/zoneinfo", horizon_year=2200):
    key = (zoneinfo_dir, horizon_year)
    if key not in _TABLES:
        h = lib().shim_table_build(zoneinfo_dir.encode(), horizon_year)
        assert h, "zone table build failed"
        _TABLES[key] = h
    return _TABLES[key]


def zone_names(t=None):
    t = t or table()
    return [lib().shim_zone_name(t, i).decode() for i in range(lib().shim_zone_count(t))]


def zone_entries(z, t=None):
    t = t or table()
    w, o = C.c_void_p(), C.c_void_p()
    n = lib().shim_zone_entries(t, z, C.byref(w), C.byref(o))
    when = np.ctypeslib.as_array(C.cast(w, C.POINTER(C.c_int64)), (n,)).copy()
    off = np.ctypeslib.as_array(C.cast(o, C.POINTER(C.c_int32)), (n,)).copy()
    return when, off


def _planes(masks):
    assert masks.dtype == np.uint64 and masks.shape[0] == 6 and masks.flags.c_contiguous
    return (C.c_void_p * 6)(*[masks[i].ctypes.data for i in range(6)])


def _p(a):
    return None if a is None else a.ctypes.data


def next_n(masks, from_, k=1, zone_id=None, t=None):
    t = t or table()
    n = masks.shape[1]
    from_ = np.ascontiguousarray(from_, dtype=np.int64)
    zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
    out = np.empty((n, k), dtype=np.int64)
    lib().shim_next_n(t, _planes(masks), _p(zid), from_.ctypes.data, n, k, out.ctypes.data)
    return out


def unmet(masks, earliest, now, now_nsec=0, policy=None, active=None, flags=None, deadline=None, zone_id=None, t=None,
          klass_out=None):
    """klass_out (uint8 array): receives which path answered each Cron -- 0 UTC machine, 1 UTC machine on a zone
    period's clock, 2 general machine, 3 general machine after the period machine rejected the Cron"""
    t = t or table()
    n = masks.shape[1]
    lib().shim_set_klass_out(None if klass_out is None else klass_out.ctypes.data)
    earliest = np.ascontiguousarray(earliest, dtype=np.int64)
    cv = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)  # noqa: E731
    zid, policy, active = cv(zone_id, np.int32), cv(policy, np.uint8), cv(active, np.uint32)
    flags, deadline = cv(flags, np.uint8), cv(deadline, np.int64)
    last, nxt = np.empty(n, np.int64), np.empty(n, np.int64)
    missed, status, decision = np.empty(n, np.uint32), np.empty(n, np.uint8), np.empty(n, np.uint8)
    lib().shim_unmet(t, _planes(masks), _p(zid), earliest.ctypes.data, now, now_nsec, _p(policy), _p(active),
                     _p(flags), _p(deadline), n, last.ctypes.data, nxt.ctypes.data, missed.ctypes.data,
                     status.ctypes.data, decision.ctypes.data)
    return dict(last=last, next=nxt, missed=missed, status=status, decision=decision)


def unmet_two_phase(masks, earliest, now, easy_rounds, now_nsec=0, policy=None, active=None, flags=None, deadline=None,
                    zone_id=None, t=None):
    """the kernels' easy/hard split: `easy_rounds` closed-form rounds, then the general machine for what is left;
    also returns rounds (easy rounds used) and deferred (1 = answered by the general machine)"""
    t = t or table()
    n = masks.shape[1]
    earliest = np.ascontiguousarray(earliest, dtype=np.int64)
    cv = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)  # noqa: E731
    zid, policy, active = cv(zone_id, np.int32), cv(policy, np.uint8), cv(active, np.uint32)
    flags, deadline = cv(flags, np.uint8), cv(deadline, np.int64)
    last, nxt = np.empty(n, np.int64), np.empty(n, np.int64)
    missed, status, decision = np.empty(n, np.uint32), np.empty(n, np.uint8), np.empty(n, np.uint8)
    rounds, deferred = np.empty(n, np.uint8), np.empty(n, np.uint8)
    lib().shim_unmet_two_phase(t, _planes(masks), _p(zid), earliest.ctypes.data, now, now_nsec, _p(policy), _p(active),
                               _p(flags), _p(deadline), n, easy_rounds, last.ctypes.data, nxt.ctypes.data,
                               missed.ctypes.data, status.ctypes.data, decision.ctypes.data, rounds.ctypes.data,
                               deferred.ctypes.data)
    return dict(last=last, next=nxt, missed=missed, status=status, decision=decision, rounds=rounds, deferred=deferred)


def parse_batch(texts, with_seconds=False):
    """the device parser's per-spec source (parse_kernels.cuh parse_one) on the host: masks (6, n), status, tz_off, tz_len"""
    blobs = [t if isinstance(t, bytes) else t.encode("utf-8", errors="surrogateescape") for t in texts]
    n = len(blobs)
    off = np.zeros(n + 1, dtype=np.uint64)
    off[1:] = np.cumsum([len(b) for b in blobs])
    masks = np.zeros((6, n), dtype=np.uint64)
    status = np.empty(n, np.uint8)
    tz_off, tz_len = np.empty(n, np.uint32), np.empty(n, np.uint32)
    lib().shim_parse_batch(b"".join(blobs), off.ctypes.data, n, int(with_seconds), _planes(masks), status.ctypes.data,
                           tz_off.ctypes.data, tz_len.ctypes.data)
    return masks, status, tz_off, tz_len
