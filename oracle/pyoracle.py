"""ctypes binding of the CPU ORACLE (oracle/liboracle.so) -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
The product package (cron_operator_b200) never does.  PARITY UNPINNED beyond the reference's K1-K4 vectors
(see oracle/cron_oracle.h).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
ZERO = -(2**63)  # Go's time.Time{} at this repo's boundaries
ZONEINFO_DIR = os.environ.get("ZONEINFO", "/usr/share/zoneinfo")

ST_OK, ST_UNSCHEDULABLE, ST_EARLIEST_AFTER_NOW, ST_NONTERMINATING = 0, 1, 2, 3
(DEC_GATE_DELETED, DEC_GATE_SUSPENDED, DEC_GATE_DEADLINE, DEC_ERR_UNSCHEDULABLE, DEC_REQUEUE_ONLY,
 DEC_SKIP_FORBID, DEC_REPLACE_CREATE, DEC_CREATE) = range(1, 9)


class Spec(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("second", "minute", "hour", "dom", "month", "dow")]

    def tuple(self):
        return (self.second, self.minute, self.hour, self.dom, self.month, self.dow)


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    src = [os.path.join(_HERE, f) for f in ("cron_oracle.c", "cron_oracle.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_load_location.restype = C.c_void_p
        L.orc_load_location.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
        L.orc_loc_free.argtypes = [C.c_void_p]
        L.orc_loc_lookup.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
        L.orc_go_date.restype = C.c_int64
        L.orc_go_date.argtypes = [C.c_int64] * 6 + [C.c_void_p]
        L.orc_fields.argtypes = [C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]
        for f in (L.orc_parse_standard, L.orc_parse_with_seconds):
            f.restype = C.c_int
            f.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(Spec), C.POINTER(C.c_int), C.POINTER(C.c_int64),
                          C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        L.orc_spec_next.restype = C.c_int64
        L.orc_spec_next.argtypes = [C.POINTER(Spec), C.c_void_p, C.c_int64]
        L.orc_last_steps.restype = C.c_int64
        L.orc_get_next_schedule.restype = C.c_int
        L.orc_get_next_schedule.argtypes = [C.POINTER(Spec), C.c_void_p, C.c_int64, C.c_int64, C.POINTER(C.c_int64),
                                            C.POINTER(C.c_int64), C.POINTER(C.c_uint64)]
        L.orc_reconcile_decide.restype = C.c_int
        L.orc_reconcile_decide.argtypes = [C.POINTER(Spec), C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_uint8,
                                           C.c_uint32, C.c_uint8, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                           C.POINTER(C.c_uint64), C.POINTER(C.c_uint8)]
        L.orc_history_prune.argtypes = [C.c_void_p, C.c_size_t, C.c_int64, C.c_void_p, C.c_void_p]
        L.orc_batch_next_n.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_int]
        L.orc_batch_unmet.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orc_batch_unmet_list.restype = C.c_uint64
        L.orc_batch_unmet_list.argtypes = [C.c_void_p] * 6 + [C.c_size_t, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t]
        L.orc_batch_history.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
        _LIB = L
    return _LIB


class ParseError(ValueError):
    pass


_LOCS: dict[str, int] = {}


def load_location(name: str, zoneinfo_dir: str = ZONEINFO_DIR) -> int:
    """time.LoadLocation; returns an opaque handle (cached for the life of the process)."""
    key = zoneinfo_dir + "\0" + name
    if key not in _LOCS:
        err = C.create_string_buffer(256)
        h = lib().orc_load_location(zoneinfo_dir.encode(), name.encode(), err, 256)
        if not h:
            raise ParseError(err.value.decode())
        _LOCS[key] = h
    return _LOCS[key]


def lookup(loc: int | None, sec: int):
    off, st, en = C.c_int32(), C.c_int64(), C.c_int64()
    lib().orc_loc_lookup(loc, sec, C.byref(off), C.byref(st), C.byref(en))
    return off.value, st.value, en.value


def go_date(y, mo, d, h, mi, s, loc=None) -> int:
    return lib().orc_go_date(y, mo, d, h, mi, s, loc)


def fields(loc, t):
    out = (C.c_int64 * 7)()
    lib().orc_fields(loc, t, out)
    return tuple(out)


def parse_standard(s: str | bytes, with_seconds: bool = False):
    """cronv3.ParseStandard -> (Spec, kind, every_sec, tz_name)."""
    b = s.encode() if isinstance(s, str) else s
    spec, kind, every = Spec(), C.c_int(), C.c_int64()
    tz, err = C.create_string_buffer(128), C.create_string_buffer(512)
    fn = lib().orc_parse_with_seconds if with_seconds else lib().orc_parse_standard
    rc = fn(b, len(b), C.byref(spec), C.byref(kind), C.byref(every), tz, 128, err, 512)
    if rc != 0:
        raise ParseError(err.value.decode(errors="replace"))
    return spec, kind.value, every.value, tz.value.decode()


def spec_from_masks(m) -> Spec:
    return Spec(*[int(x) for x in m])


def spec_next(spec: Spec, loc, t: int) -> int:
    return lib().orc_spec_next(C.byref(spec), loc, t)


def last_steps() -> int:
    return lib().orc_last_steps()


def get_next_schedule(spec: Spec, loc, earliest: int, now: int):
    last, nxt, missed = C.c_int64(), C.c_int64(), C.c_uint64()
    st = lib().orc_get_next_schedule(C.byref(spec), loc, earliest, now, C.byref(last), C.byref(nxt), C.byref(missed))
    return st, last.value, nxt.value, missed.value


def reconcile_decide(spec, loc, earliest, now, now_nsec=0, policy=0, active=0, flags=0, deadline=ZERO):
    last, nxt, missed, st = C.c_int64(), C.c_int64(), C.c_uint64(), C.c_uint8()
    d = lib().orc_reconcile_decide(C.byref(spec), loc, earliest, now, now_nsec, policy, active, flags, deadline,
                                   C.byref(last), C.byref(nxt), C.byref(missed), C.byref(st))
    return d, st.value, last.value, nxt.value, missed.value


def history_prune(keys, limit: int | None):
    k = np.ascontiguousarray(keys, dtype=np.int64)
    perm = np.empty(len(k), dtype=np.uint32)
    keep = np.empty(len(k), dtype=np.uint8)
    lib().orc_history_prune(k.ctypes.data, len(k), -1 if limit is None else limit, perm.ctypes.data, keep.ctypes.data)
    return perm, keep


def _ptr(a):
    return None if a is None else a.ctypes.data


def _plane_ptrs(masks):
    """masks: (6, n) uint64 C-contiguous -> C array of 6 plane pointers"""
    assert masks.dtype == np.uint64 and masks.shape[0] == 6 and masks.flags.c_contiguous
    arr = (C.c_void_p * 6)(*[masks[i].ctypes.data for i in range(6)])
    return arr


def _loc_table(zone_names, zoneinfo_dir=ZONEINFO_DIR):
    if zone_names is None:
        return None
    hs = [load_location(z, zoneinfo_dir) for z in zone_names]
    return (C.c_void_p * len(hs))(*hs)


def cpu_count() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return os.cpu_count() or 1


def batch_next_n(masks, from_, k=1, zone_id=None, zone_names=None, threads=None):
    n = masks.shape[1]
    from_ = np.ascontiguousarray(from_, dtype=np.int64)
    out = np.empty((n, k), dtype=np.int64)
    planes = _plane_ptrs(masks)
    locs = _loc_table(zone_names)
    zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
    lib().orc_batch_next_n(planes, _ptr(zid), locs, from_.ctypes.data, n, k, out.ctypes.data, threads or cpu_count())
    return out


def batch_unmet(masks, earliest, now, now_nsec=0, policy=None, active=None, flags=None, deadline=None, zone_id=None,
                zone_names=None, threads=None):
    n = masks.shape[1]
    earliest = np.ascontiguousarray(earliest, dtype=np.int64)
    last = np.empty(n, np.int64)
    nxt = np.empty(n, np.int64)
    missed = np.empty(n, np.uint32)
    status = np.empty(n, np.uint8)
    decision = np.empty(n, np.uint8)
    planes = _plane_ptrs(masks)
    locs = _loc_table(zone_names)
    zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
    cv = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)  # noqa: E731
    policy, active, flags, deadline = cv(policy, np.uint8), cv(active, np.uint32), cv(flags, np.uint8), cv(deadline, np.int64)
    lib().orc_batch_unmet(planes, _ptr(zid), locs, earliest.ctypes.data, now, now_nsec, _ptr(policy), _ptr(active),
                          _ptr(flags), _ptr(deadline), n, last.ctypes.data, nxt.ctypes.data, missed.ctypes.data,
                          status.ctypes.data, decision.ctypes.data, threads or cpu_count())
    return dict(last=last, next=nxt, missed=missed, status=status, decision=decision)


def batch_history(seg_off, key, limit, threads=None):
    seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
    key = np.ascontiguousarray(key, dtype=np.int64)
    limit = None if limit is None else np.ascontiguousarray(limit, dtype=np.int64)
    m = len(key)
    perm = np.empty(m, np.uint32)
    keep = np.empty(m, np.uint8)
    lib().orc_batch_history(seg_off.ctypes.data, key.ctypes.data, _ptr(limit), len(seg_off) - 1, perm.ctypes.data,
                            keep.ctypes.data, threads or cpu_count())
    return perm, keep


def batch_unmet_list(masks, earliest, missed, decision, cap, zone_id=None, zone_names=None):
    n = masks.shape[1]
    earliest = np.ascontiguousarray(earliest, dtype=np.int64)
    missed = np.ascontiguousarray(missed, dtype=np.uint32)
    decision = np.ascontiguousarray(decision, dtype=np.uint8)
    zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
    planes, locs = _plane_ptrs(masks), _loc_table(zone_names)
    off = np.empty(n + 1, np.uint64)
    total = int(np.minimum(missed, cap)[decision >= DEC_REQUEUE_ONLY].sum())
    lst = np.empty(total, np.int64)
    got = lib().orc_batch_unmet_list(planes, _ptr(zid), locs, earliest.ctypes.data, missed.ctypes.data,
                                     decision.ctypes.data, n, cap, off.ctypes.data, lst.ctypes.data, total)
    assert got == total
    return off, lst
