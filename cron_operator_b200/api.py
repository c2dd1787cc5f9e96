"""Host-side mirror of the reference interface for the schedule path, bound to libb200cron.so over its C ABI.

Reference shapes mirrored here (names follow the Go code so that parity tests read like the reference's own):
  cronv3.ParseStandard(spec) -> Schedule            internal/controller/cron_controller.go:392
  Schedule.Next(t)                                  :405, :409, :436
  getNextSchedule(cron, now) -> (lastMissed, next)  :389-437
  Reconcile's gates + Allow/Forbid/Replace decision :163-220
  sortByCreationTimestamp / syncCronHistory prune   internal/controller/cron_util.go:117-129, cron_controller.go:307-346
plus the batched calls the path exists for: NextN / UnmetSince / HistoryPrune over arrays of specs.

There is no CPU execution path: every compute call goes to the CUDA library and raises when it (or a GPU) is missing.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

from . import build as _build

ZERO = -(2**63)  # Go's time.Time{} at this boundary (B200CRON_ZERO_TIME)
OK, EINVAL, ECUDA, ERANGE, EPARSE, ENOMEM = 0, -1, -2, -3, -4, -5
ST_OK, ST_UNSCHEDULABLE, ST_EARLIEST_AFTER_NOW, ST_NONTERMINATING = 0, 1, 2, 3
(DEC_GATE_DELETED, DEC_GATE_SUSPENDED, DEC_GATE_DEADLINE, DEC_ERR_UNSCHEDULABLE, DEC_REQUEUE_ONLY, DEC_SKIP_FORBID,
 DEC_REPLACE_CREATE, DEC_CREATE) = range(1, 9)
POLICY = {"Allow": 0, "Forbid": 1, "Replace": 2, "": 0}  # api/v1alpha1/cron_types.go:127-139
FLAG_SUSPEND, FLAG_DELETED = 1, 2
KIND_SPEC, KIND_EVERY = 0, 1
EVERY_TAG = 1 << 62  # B200CRON_EVERY_TAG: month plane of a ConstantDelaySchedule, its delay (seconds) in the dom plane


class B200CronError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b200cron error {code}: {msg}")
        self.code = code


class ParseError(ValueError):
    """cronv3.ParseStandard returned an error (the reference reports it as 'unparsable cron %q: %w')."""


class _Spec(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("second", "minute", "hour", "dom", "month", "dow")]


_LIB = None


def lib_path() -> str:
    return _build.LIB


def lib(path: str | None = None):
    """Loads csrc/libb200cron.so (building it with nvcc when sources are newer).  Fails loudly when unavailable.
    `path`: load that build instead (first call only; the A/B scripts under profiles/ pass their tuning builds)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    L = C.CDLL(path or _build.build())
    vp, sz, i32, i64 = C.c_void_p, C.c_size_t, C.c_int, C.c_int64
    for f in (L.b200cron_parse_standard, L.b200cron_parse_with_seconds):
        f.restype = i32
        f.argtypes = [C.c_char_p, sz, C.POINTER(_Spec), C.POINTER(i32), C.POINTER(i64), C.c_char_p, sz, C.c_char_p, sz]
    L.b200cron_ctx_create.restype = i32
    L.b200cron_ctx_create.argtypes = [C.POINTER(i32), i32, C.c_char_p, i32, C.POINTER(vp)]
    L.b200cron_ctx_destroy.argtypes = [vp]
    L.b200cron_last_error.restype = C.c_char_p
    L.b200cron_last_error.argtypes = [vp]
    L.b200cron_device_count.argtypes = [vp]
    L.b200cron_zone_id.argtypes = [vp, C.c_char_p]
    L.b200cron_zone_count.argtypes = [vp]
    L.b200cron_zone_name.restype = C.c_char_p
    L.b200cron_zone_name.argtypes = [vp, i32]
    L.b200cron_zone_periods.argtypes = [vp, i32, C.POINTER(vp), C.POINTER(vp)]
    L.b200cron_host_alloc.restype = vp
    L.b200cron_host_alloc.argtypes = [sz]
    L.b200cron_host_free.argtypes = [vp]
    L.b200cron_next_n.restype = i32
    L.b200cron_next_n.argtypes = [vp, vp, vp, vp, sz, i32, vp]
    L.b200cron_next_n_dev.restype = i32
    L.b200cron_next_n_dev.argtypes = [vp, i32, vp, vp, vp, vp, sz, i32, vp]
    L.b200cron_unmet_since.restype = i32
    L.b200cron_unmet_since.argtypes = [vp, vp, vp, vp, i64, C.c_int32, vp, vp, vp, vp, sz, vp, vp, vp, vp, vp]
    L.b200cron_unmet_since_dev.restype = i32
    L.b200cron_unmet_since_dev.argtypes = [vp, i32, vp, vp, vp, vp, i64, C.c_int32, vp, vp, vp, vp, sz, vp, vp, vp, vp, vp]
    L.b200cron_unmet_list.restype = i32
    L.b200cron_unmet_list.argtypes = [vp, vp, vp, vp, vp, vp, sz, C.c_uint32, vp, vp, sz]
    L.b200cron_unmet_list_offsets_dev.restype = i32
    L.b200cron_unmet_list_offsets_dev.argtypes = [vp, i32, vp, vp, vp, sz, C.c_uint32, vp]
    L.b200cron_unmet_list_offsets_static.restype = i32
    L.b200cron_unmet_list_offsets_static.argtypes = [vp, i32, vp, vp, vp, sz, C.c_uint32, vp]
    L.b200cron_unmet_list_emit_dev.restype = i32
    L.b200cron_unmet_list_emit_dev.argtypes = [vp, i32, vp, vp, vp, vp, sz, vp, vp]
    L.b200cron_history_prune.restype = i32
    L.b200cron_history_prune.argtypes = [vp, vp, vp, vp, sz, vp, vp]
    L.b200cron_history_prune_dev.restype = i32
    L.b200cron_history_prune_dev.argtypes = [vp, i32, vp, vp, vp, vp, sz, sz, vp, vp]
    L.b200cron_synth_specs_dev.restype = i32
    L.b200cron_synth_specs_dev.argtypes = [vp, i32, vp, C.c_uint64, C.c_uint64, sz, i32, C.c_double, C.c_double, vp, vp]
    L.b200cron_synth_times_dev.restype = i32
    L.b200cron_synth_times_dev.argtypes = [vp, i32, vp, C.c_uint64, C.c_uint64, sz, i64, i64, C.c_uint64, vp]
    L.b200cron_synth_zones_dev.restype = i32
    L.b200cron_synth_zones_dev.argtypes = [vp, i32, vp, C.c_uint64, C.c_uint64, sz, vp, sz, C.c_double, vp]
    L.b200cron_parse_batch.restype = i32
    L.b200cron_parse_batch.argtypes = [vp, vp, vp, sz, i32, vp, vp, vp, vp]
    L.b200cron_parse_batch_dev.restype = i32
    L.b200cron_parse_batch_dev.argtypes = [vp, i32, vp, vp, vp, sz, i32, vp, vp, vp, vp]
    L.b200cron_specset_create.restype = i32
    L.b200cron_specset_create.argtypes = [vp, vp, vp, sz, C.POINTER(vp)]
    L.b200cron_specset_update.restype = i32
    L.b200cron_specset_update.argtypes = [vp, sz, sz, vp, vp]
    L.b200cron_specset_destroy.argtypes = [vp]
    L.b200cron_specset_size.restype = sz
    L.b200cron_specset_size.argtypes = [vp]
    L.b200cron_next_n_set.restype = i32
    L.b200cron_next_n_set.argtypes = [vp, vp, vp, i32, vp]
    L.b200cron_unmet_since_set.restype = i32
    L.b200cron_unmet_since_set.argtypes = [vp, vp, vp, i64, C.c_int32, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    L.b200cron_ctx_reserve_unmet.restype = i32
    L.b200cron_ctx_reserve_unmet.argtypes = [vp, sz]
    L.b200cron_debug_last_deferred.restype = i32
    L.b200cron_debug_last_deferred.argtypes = [vp, i32, vp, vp, sz, C.POINTER(sz)]
    L.b200cron_sync_check.restype = i32
    L.b200cron_sync_check.argtypes = [vp, i32, vp]
    L.b200cron_ctx_set_overlap.restype = i32
    L.b200cron_ctx_set_overlap.argtypes = [vp, i32]
    L.b200cron_launch_count.restype = C.c_uint64
    L.b200cron_launch_count.argtypes = [vp]
    _LIB = L
    return L


# --------------------------------------------------------------------------------------------------------------
# parsing (host)
# --------------------------------------------------------------------------------------------------------------
@dataclass(frozen=True)
class SpecSchedule:
    """robfig/cron v3 spec.go SpecSchedule; Location is carried as a zone name ('' == time.Local)."""
    Second: int
    Minute: int
    Hour: int
    Dom: int
    Month: int
    Dow: int
    Location: str = ""
    Every: int = 0  # > 0: ConstantDelaySchedule{Delay: Every seconds} from "@every <dur>"

    def masks(self):
        """the six planes as the batched calls take them (a ConstantDelaySchedule in its B200CRON_EVERY_TAG encoding)"""
        if self.Every:
            return (0, 0, 0, self.Every, EVERY_TAG, 0)
        return (self.Second, self.Minute, self.Hour, self.Dom, self.Month, self.Dow)


def _parse(text, with_seconds) -> SpecSchedule:
    b = text.encode() if isinstance(text, str) else bytes(text)
    spec, kind, every = _Spec(), C.c_int(), C.c_int64()
    tz, err = C.create_string_buffer(128), C.create_string_buffer(512)
    fn = lib().b200cron_parse_with_seconds if with_seconds else lib().b200cron_parse_standard
    rc = fn(b, len(b), C.byref(spec), C.byref(kind), C.byref(every), tz, 128, err, 512)
    if rc != 0:
        raise ParseError(err.value.decode(errors="replace"))
    if kind.value == KIND_EVERY:
        assert (spec.month, spec.dom) == (EVERY_TAG, every.value)
        return SpecSchedule(0, 0, 0, 0, 0, 0, tz.value.decode(), every.value)
    return SpecSchedule(spec.second, spec.minute, spec.hour, spec.dom, spec.month, spec.dow, tz.value.decode(), 0)


def ParseStandard(spec: str) -> SpecSchedule:
    """cronv3.ParseStandard (cron_controller.go:392)."""
    return _parse(spec, False)


def ParseWithSeconds(spec: str) -> SpecSchedule:
    """extension: upstream NewParser(Second|Minute|Hour|Dom|Month|Dow|Descriptor)."""
    return _parse(spec, True)


def pack_specs(specs) -> np.ndarray:
    """list of SpecSchedule -> (6, n) uint64 SoA planes"""
    return np.ascontiguousarray(np.array([s.masks() for s in specs], dtype=np.uint64).T.reshape(6, len(specs)))


# --------------------------------------------------------------------------------------------------------------
# context
# --------------------------------------------------------------------------------------------------------------
def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return a.data_ptr()  # torch tensor


def _planes(masks, five_field: bool = False):
    """six plane pointers; five_field=True passes NULL for the seconds plane (ParseStandard schedules fire at second 0)"""
    if isinstance(masks, np.ndarray):
        assert masks.dtype == np.uint64 and masks.ndim == 2 and masks.shape[0] == 6 and masks.flags.c_contiguous
        ptrs = [masks[i].ctypes.data for i in range(6)]
        if five_field:
            ptrs[0] = None
        return (C.c_void_p * 6)(*ptrs), masks.shape[1]
    # torch tensor (6, n) int64/uint64 on the device, or a sequence of 6 tensors
    planes = [masks[i] for i in range(6)]
    for p in planes:
        assert p.is_contiguous() and p.element_size() == 8
    ptrs = [p.data_ptr() for p in planes]
    if five_field:
        ptrs[0] = None
    return (C.c_void_p * 6)(*ptrs), planes[0].numel()


class Context:
    """b200cron_ctx: streams, device staging and the IANA zone table on each device."""

    def __init__(self, devices=None, zoneinfo_dir: str | None = None, horizon_year: int = 0, local_zone: str = "UTC"):
        self._h = C.c_void_p()
        L = lib()
        if devices is None:
            arr, n = None, 0
        else:
            arr, n = (C.c_int * len(devices))(*devices), len(devices)
        rc = L.b200cron_ctx_create(arr, n, zoneinfo_dir.encode() if zoneinfo_dir else None, horizon_year, C.byref(self._h))
        if rc != 0:
            raise B200CronError(rc, L.b200cron_last_error(None).decode())
        self.local_zone = local_zone  # stands in for the process's time.Local (chart value useHostTimezone)

    def close(self):
        if self._h:
            lib().b200cron_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            raise B200CronError(rc, lib().b200cron_last_error(self._h).decode())

    # ---- zones
    def zone_id(self, name: str) -> int:
        z = lib().b200cron_zone_id(self._h, name.encode())
        if z < 0:
            raise ParseError(f"provided bad location {name}: unknown time zone {name}")
        return z

    def location_id(self, sched: SpecSchedule) -> int:
        """the zone a schedule is evaluated in: its TZ= location, else time.Local (robfig spec.go Next)"""
        name = sched.Location
        if name in ("", "Local"):
            name = self.local_zone
        return self.zone_id(name)

    @property
    def zone_names(self):
        L = lib()
        return [L.b200cron_zone_name(self._h, i).decode() for i in range(L.b200cron_zone_count(self._h))]

    def zone_periods(self, zid: int):
        w, o = C.c_void_p(), C.c_void_p()
        n = lib().b200cron_zone_periods(self._h, zid, C.byref(w), C.byref(o))
        if n < 0:
            raise B200CronError(n, "bad zone id")
        when = np.ctypeslib.as_array(C.cast(w, C.POINTER(C.c_int64)), (n,)).copy()
        off = np.ctypeslib.as_array(C.cast(o, C.POINTER(C.c_int32)), (n,)).copy()
        return when, off

    @property
    def device_count(self):
        return lib().b200cron_device_count(self._h)

    @property
    def launch_count(self):
        return int(lib().b200cron_launch_count(self._h))

    # ---- host-pointer calls (numpy in, numpy out; copies inside the library)
    def next_n(self, masks: np.ndarray, from_, k: int = 1, zone_id=None, out=None, five_field: bool = False) -> np.ndarray:
        planes, n = _planes(masks, five_field)
        from_ = np.ascontiguousarray(from_, dtype=np.int64)
        zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
        if out is None:
            out = np.empty((n, k), dtype=np.int64)
        self._check(lib().b200cron_next_n(self._h, planes, _ptr(zid), from_.ctypes.data, n, k, out.ctypes.data))
        return out

    def unmet_since(self, masks: np.ndarray, earliest, now: int, now_nsec: int = 0, policy=None, active=None,
                    flags=None, deadline=None, zone_id=None, out=None, five_field: bool = False) -> dict:
        planes, n = _planes(masks, five_field)
        cv = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)  # noqa: E731
        earliest = cv(earliest, np.int64)
        zid, policy, active = cv(zone_id, np.int32), cv(policy, np.uint8), cv(active, np.uint32)
        flags, deadline = cv(flags, np.uint8), cv(deadline, np.int64)
        if out is None:
            out = dict(last=np.empty(n, np.int64), next=np.empty(n, np.int64), missed=np.empty(n, np.uint32),
                       status=np.empty(n, np.uint8), decision=np.empty(n, np.uint8))
        self._check(lib().b200cron_unmet_since(self._h, planes, _ptr(zid), earliest.ctypes.data, now, now_nsec,
                                               _ptr(policy), _ptr(active), _ptr(flags), _ptr(deadline), n,
                                               out["last"].ctypes.data, out["next"].ctypes.data,
                                               out["missed"].ctypes.data, out["status"].ctypes.data,
                                               out["decision"].ctypes.data))
        return out

    def unmet_list(self, masks: np.ndarray, earliest, missed, decision, cap: int = 100, zone_id=None):
        """the first min(missed, cap) missed fire times per Cron: (list_off[n+1], list) -- cap 100 mirrors the
        TooManyMissedTimes threshold (cron_controller.go:431)"""
        planes, n = _planes(masks)
        earliest = np.ascontiguousarray(earliest, dtype=np.int64)
        missed = np.ascontiguousarray(missed, dtype=np.uint32)
        decision = np.ascontiguousarray(decision, dtype=np.uint8)
        zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
        off = np.empty(n + 1, np.uint64)
        total = int(np.minimum(missed, cap)[decision >= DEC_REQUEUE_ONLY].sum())
        lst = np.empty(total, np.int64)
        self._check(lib().b200cron_unmet_list(self._h, planes, _ptr(zid), earliest.ctypes.data, missed.ctypes.data,
                                              decision.ctypes.data, n, cap, off.ctypes.data, lst.ctypes.data, total))
        return off, lst

    def unmet_list_offsets_dev(self, missed, decision, list_off, cap: int = 100, dev_index: int = 0, stream: int = 0):
        self._check(lib().b200cron_unmet_list_offsets_dev(self._h, dev_index, stream or None, _ptr(missed),
                                                          _ptr(decision), missed.numel(), cap, _ptr(list_off)))

    def unmet_list_offsets_static(self, missed, decision, list_off, cap: int = 100, dev_index: int = 0,
                                  stream: int = 0, n: int | None = None):
        """unmet_list_offsets_dev for hosts that must not allocate device memory at call time (a GPU already handed
        to the host's own pool allocator, or a loop that cannot take cudaMalloc's device-wide sync): the per-tile sums
        sit in a fixed 128 KiB array inside the library.  At most 16,777,216 Crons per call; calls that overlap on
        one GPU race on that array.  First call per device resolves the array's address -- do it before capturing."""
        self._check(lib().b200cron_unmet_list_offsets_static(self._h, dev_index, stream or None, _ptr(missed),
                                                             _ptr(decision), missed.numel() if n is None else n, cap,
                                                             _ptr(list_off)))

    def unmet_list_emit_dev(self, masks, earliest, list_off, out_list, zone_id=None, dev_index: int = 0, stream: int = 0):
        planes, n = _planes(masks)
        self._check(lib().b200cron_unmet_list_emit_dev(self._h, dev_index, stream or None, planes, _ptr(zone_id),
                                                       _ptr(earliest), n, _ptr(list_off), _ptr(out_list)))

    def history_prune(self, seg_off, key, limit=None):
        seg_off = np.ascontiguousarray(seg_off, dtype=np.uint64)
        key = np.ascontiguousarray(key, dtype=np.int64)
        limit = None if limit is None else np.ascontiguousarray(limit, dtype=np.int64)
        m = len(key)
        perm, keep = np.empty(m, np.uint32), np.empty(m, np.uint8)
        self._check(lib().b200cron_history_prune(self._h, seg_off.ctypes.data, key.ctypes.data, _ptr(limit),
                                                 len(seg_off) - 1, perm.ctypes.data, keep.ctypes.data))
        return perm, keep

    # ---- resident spec sets: the immutable masks stay in HBM, a tick moves only times in and results out
    def specset(self, masks: np.ndarray, zone_id=None, five_field: bool = False) -> "SpecSet":
        return SpecSet(self, masks, zone_id, five_field)

    # ---- device-pointer calls (torch tensors resident on device `dev_index`; enqueue only)
    def next_n_dev(self, masks, from_, out, k: int = 1, zone_id=None, dev_index: int = 0, stream: int = 0,
                   five_field: bool = False):
        planes, n = _planes(masks, five_field)
        self._check(lib().b200cron_next_n_dev(self._h, dev_index, stream or None, planes, _ptr(zone_id), _ptr(from_),
                                              n, k, _ptr(out)))

    def unmet_since_dev(self, masks, earliest, now, out: dict, now_nsec=0, policy=None, active=None, flags=None,
                        deadline=None, zone_id=None, dev_index: int = 0, stream: int = 0, five_field: bool = False):
        planes, n = _planes(masks, five_field)
        self._check(lib().b200cron_unmet_since_dev(self._h, dev_index, stream or None, planes, _ptr(zone_id),
                                                   _ptr(earliest), now, now_nsec, _ptr(policy), _ptr(active),
                                                   _ptr(flags), _ptr(deadline), n, _ptr(out["last"]),
                                                   _ptr(out["next"]), _ptr(out["missed"]), _ptr(out["status"]),
                                                   _ptr(out["decision"])))

    def history_prune_dev(self, seg_off, key, limit, perm, keep, dev_index: int = 0, stream: int = 0):
        nseg = seg_off.numel() - 1
        self._check(lib().b200cron_history_prune_dev(self._h, dev_index, stream or None, _ptr(seg_off), _ptr(key),
                                                     _ptr(limit), nseg, key.numel(), _ptr(perm), _ptr(keep)))

    # ---- batched ParseStandard on the device
    def parse_batch(self, specs, with_seconds: bool = False):
        """cronv3.ParseStandard over a list of schedule strings on the GPU.  Returns (masks (6, n) uint64, status (n,)
        uint8: 0 ok (an "@every" with plain h/m/s terms comes back in the EVERY_TAG encoding) / 1 unparsable / 2 an
        "@every" duration only the host parser handles, tz names list (None when the spec has no TZ= prefix))."""
        enc = [s.encode() if isinstance(s, str) else bytes(s) for s in specs]
        n = len(enc)
        off = np.zeros(n + 1, dtype=np.uint64)
        off[1:] = np.cumsum([len(b) for b in enc])
        blob = b"".join(enc)
        text = np.frombuffer(blob, dtype=np.uint8) if blob else np.zeros(1, np.uint8)
        masks = np.zeros((6, n), dtype=np.uint64)
        status = np.zeros(n, np.uint8)
        tzo, tzl = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        planes, _ = _planes(masks)
        self._check(lib().b200cron_parse_batch(self._h, text.ctypes.data, off.ctypes.data, n, int(with_seconds), planes,
                                               status.ctypes.data, tzo.ctypes.data, tzl.ctypes.data))
        tz = []
        for i in range(n):
            if tzl[i] == 0 and tzo[i] == 0:
                tz.append(None)
            elif tzo[i] == 0xFFFFFFFF:
                tz.append("UTC")
            else:
                tz.append(enc[i][int(tzo[i]):int(tzo[i]) + int(tzl[i])].decode(errors="replace"))
        return masks, status, tz

    def ParseStandardBatch(self, specs):
        """list of schedule strings -> list of SpecSchedule (ParseError raised for the first unparsable one, with
        upstream's message from the host parser; @every and zone-name resolution are finished on the host)"""
        masks, status, tz = self.parse_batch(specs)
        out = []
        for i, text in enumerate(specs):
            if status[i] != 0:
                out.append(ParseStandard(text))  # raises for status 1, returns the ConstantDelay form for status 2
                continue
            if tz[i] not in (None, "Local"):
                self.zone_id(tz[i])  # time.LoadLocation error for unknown names
            if int(masks[4, i]) == EVERY_TAG:  # "@every 1h30m" parsed on the device
                out.append(SpecSchedule(0, 0, 0, 0, 0, 0, tz[i] or "", int(masks[3, i])))
            else:
                out.append(SpecSchedule(*[int(masks[p, i]) for p in range(6)], tz[i] or "", 0))
        return out

    # ---- device-side synthetic workloads (bit-identical to synth.py)
    def synth_specs_dev(self, masks_out, n: int, seed: int = 0xC0FFEE, start: int = 0, with_seconds: bool = False,
                        dense_minutes: float = 0.0, plant_unschedulable: float = 0.001, dev_index: int = 0, stream: int = 0):
        from . import synth
        th = np.concatenate([synth.form_thresholds(0.40), synth.form_thresholds(0.60)]).astype(np.float64)
        planes, _ = _planes(masks_out)
        self._check(lib().b200cron_synth_specs_dev(self._h, dev_index, stream or None, seed, start, n, int(with_seconds),
                                                   dense_minutes, plant_unschedulable, th.ctypes.data, planes))

    def synth_times_dev(self, out, n: int, lo: int, hi: int, seed: int = 0xC0FFEE, start: int = 0, stream_id: int = 200,
                        dev_index: int = 0, stream: int = 0):
        self._check(lib().b200cron_synth_times_dev(self._h, dev_index, stream or None, seed, start, n, lo, hi, stream_id,
                                                   _ptr(out)))

    def synth_zones_dev(self, out, n: int, pool, seed: int = 0xC0FFEE, start: int = 0, p_utc: float = 0.5,
                        dev_index: int = 0, stream: int = 0):
        self._check(lib().b200cron_synth_zones_dev(self._h, dev_index, stream or None, seed, start, n, _ptr(pool),
                                                   pool.numel(), p_utc, _ptr(out)))

    def set_overlap(self, on: bool = True):
        """b200cron_ctx_set_overlap: back-to-back k == 1 next_n_dev launches on one stream may overlap (the caller
        promises they are independent batches -- see include/b200cron.h)"""
        self._check(lib().b200cron_ctx_set_overlap(self._h, 1 if on else 0))

    def reserve_unmet(self, max_n: int):
        """allocate the unmet scratch for batches of up to max_n Crons per device now (no allocation inside a tick)"""
        self._check(lib().b200cron_ctx_reserve_unmet(self._h, max_n))

    def last_deferred(self, cap: int = 1 << 24, dev_index: int = 0, stream: int = 0):
        """test support: (indices, count) of the Crons the most recent unmet launch handed to the general walk"""
        cnt = C.c_size_t()
        buf = np.empty(cap, np.uint32)
        self._check(lib().b200cron_debug_last_deferred(self._h, dev_index, stream or None, buf.ctypes.data, cap, C.byref(cnt)))
        return buf[:min(cap, cnt.value)].copy(), int(cnt.value)

    def sync_check(self, dev_index: int = 0, stream: int = 0):
        self._check(lib().b200cron_sync_check(self._h, dev_index, stream or None))

    # ---- the reference's scalar shapes, served by the batched kernels (n = 1: launch latency >> one CPU Next)
    def Next(self, sched: SpecSchedule, t: int) -> int:
        """Schedule.Next(t) (cron_controller.go:405,436), either kind of schedule"""
        return int(self.NextN([sched], [t], 1)[0, 0])

    def NextN(self, specs, from_, k: int) -> np.ndarray:
        zid = np.array([self.location_id(s) for s in specs], dtype=np.int32)
        return self.next_n(pack_specs(specs), np.asarray(from_, dtype=np.int64), k, zid)

    def getNextSchedule(self, schedule: str, earliest: int, now: int):
        """getNextSchedule (cron_controller.go:389-437): returns (lastMissed, next); raises like the reference errors"""
        try:
            s = ParseStandard(schedule)
        except ParseError as e:
            raise ValueError(f'unparsable cron "{schedule}": {e}') from e
        r = self.UnmetSince([s], [earliest], now)
        if r["status"][0] in (ST_UNSCHEDULABLE, ST_NONTERMINATING):
            raise ValueError(f'unschedulable cron "{schedule}"')
        return int(r["last"][0]), int(r["next"][0])

    def UnmetSince(self, specs, earliest, now: int, policy=None, active=None, flags=None, deadline=None,
                   now_nsec: int = 0) -> dict:
        zid = np.array([self.location_id(s) for s in specs], dtype=np.int32)
        pol = None if policy is None else np.array([POLICY[p] if isinstance(p, str) else p for p in policy], np.uint8)
        return self.unmet_since(pack_specs(specs), np.asarray(earliest, dtype=np.int64), now, now_nsec, pol, active,
                                flags, deadline, zid)

    def sortByCreationTimestamp(self, creation_timestamps):
        """cron_util.go:117-129: returns the stable ascending order as indices"""
        ts = np.asarray(creation_timestamps, dtype=np.int64)
        perm, _ = self.history_prune(np.array([0, len(ts)], dtype=np.uint64), ts, None)
        return perm

    def syncCronHistory(self, creation_timestamps, history_limit):
        """cron_controller.go:307-346 index math: (indices to delete, indices kept in history order)"""
        ts = np.asarray(creation_timestamps, dtype=np.int64)
        lim = np.array([-1 if history_limit is None else history_limit], dtype=np.int64)
        perm, keep = self.history_prune(np.array([0, len(ts)], dtype=np.uint64), ts, lim)
        return perm[keep == 0], perm[keep == 1]


class SpecSet:
    """b200cron_specset: n schedules resident on the context's devices (SpecSchedule is immutable after ParseStandard,
    cron_controller.go:392); next_n / unmet_since over them move only the per-tick arrays"""

    def __init__(self, ctx: Context, masks: np.ndarray, zone_id=None, five_field: bool = False):
        self.ctx, self._h = ctx, C.c_void_p()
        planes, n = _planes(masks, five_field)
        zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
        ctx._check(lib().b200cron_specset_create(ctx._h, planes, _ptr(zid), n, C.byref(self._h)))
        self.n, self.five_field, self.zoned = n, five_field, zone_id is not None

    def close(self):
        if self._h:
            lib().b200cron_specset_destroy(self._h)
            self._h = C.c_void_p()

    def __len__(self):
        return int(lib().b200cron_specset_size(self._h))

    def update(self, first: int, masks: np.ndarray, zone_id=None):
        planes, count = _planes(masks, self.five_field)
        zid = None if zone_id is None else np.ascontiguousarray(zone_id, dtype=np.int32)
        self.ctx._check(lib().b200cron_specset_update(self._h, first, count, planes, _ptr(zid)))

    def next_n(self, from_, k: int = 1, out=None) -> np.ndarray:
        from_ = np.ascontiguousarray(from_, dtype=np.int64)
        if out is None:
            out = np.empty((self.n, k), dtype=np.int64)
        self.ctx._check(lib().b200cron_next_n_set(self.ctx._h, self._h, from_.ctypes.data, k, out.ctypes.data))
        return out

    def unmet_since(self, earliest, now: int, now_nsec: int = 0, policy=None, active=None, flags=None, deadline=None,
                    out=None) -> dict:
        cv = lambda a, dt: None if a is None else np.ascontiguousarray(a, dtype=dt)  # noqa: E731
        earliest = cv(earliest, np.int64)
        policy, active, flags, deadline = cv(policy, np.uint8), cv(active, np.uint32), cv(flags, np.uint8), cv(deadline, np.int64)
        n = self.n
        if out is None:
            out = dict(last=np.empty(n, np.int64), next=np.empty(n, np.int64), missed=np.empty(n, np.uint32),
                       status=np.empty(n, np.uint8), decision=np.empty(n, np.uint8))
        self.ctx._check(lib().b200cron_unmet_since_set(self.ctx._h, self._h, earliest.ctypes.data, now, now_nsec,
                                                       _ptr(policy), _ptr(active), _ptr(flags), _ptr(deadline),
                                                       out["last"].ctypes.data, out["next"].ctypes.data,
                                                       out["missed"].ctypes.data, out["status"].ctypes.data,
                                                       out["decision"].ctypes.data))
        return out


def getDefaultJobName(cron_name: str, schedule_time: int) -> str:
    """cron_util.go:69-71: fmt.Sprintf("%s-%d", cron.Name, scheduleTime.Unix()) -- the workload is named from nextRun
    (cron_controller.go:222), which makes duplicate creates collide on AlreadyExists"""
    return f"{cron_name}-{int(schedule_time)}"


def actions_from(result: dict, names, now: int):
    """Ready-to-apply records for the step right after the path: per Cron the decision,
    RequeueAfter = next - now (cron_controller.go:192), the workload name for creates (:222, cron_util.go:69-71) and
    whether Status.LastScheduleTime is set to now (:237)."""
    out = []
    for i, name in enumerate(names):
        d = int(result["decision"][i])
        nxt = int(result["next"][i])
        creates = d in (DEC_CREATE, DEC_REPLACE_CREATE)
        out.append(dict(
            cron=name, decision=d,
            requeue_after=(nxt - now) if d >= DEC_REQUEUE_ONLY and nxt != ZERO else None,
            delete_active=d == DEC_REPLACE_CREATE,
            create=getDefaultJobName(name, nxt) if creates else None,
            set_last_schedule_time=now if creates else None,
            too_many_missed=int(result["missed"][i]) > 100,  # TooManyMissedTimes event, :431-434
        ))
    return out


class PinnedArray:
    """numpy view over cudaHostAlloc'd memory (b200cron_host_alloc) for the host-pointer calls"""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(shape) if isinstance(shape, (tuple, list)) else (shape,)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        self._p = lib().b200cron_host_alloc(max(nbytes, 1))
        if not self._p:
            raise MemoryError("cudaHostAlloc failed")
        buf = (C.c_char * max(nbytes, 1)).from_address(self._p)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self._p:
            self.array = None
            lib().b200cron_host_free(self._p)
            self._p = None

    def __del__(self):  # pragma: no cover -- views handed out via .array must not outlive the PinnedArray
        try:
            self.free()
        except Exception:
            pass
