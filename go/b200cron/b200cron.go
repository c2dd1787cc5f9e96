// Package b200cron is the cgo shim a cron-operator maintainer would add to call libb200cron.so
// (include/b200cron.h) from internal/controller.  It keeps the robfig `cron.Schedule` shape
// (Next(time.Time) time.Time) and adds the batched calls the GPU path exists for.
//
// UNVERIFIED SOURCE: the image this repository is built in has no Go toolchain (`go version`: not
// found), so this file has never been compiled.  The C ABI it binds is exercised by the C++/Python
// hosts and tests of this repository.
package b200cron

/*
#cgo CFLAGS: -I${SRCDIR}/../../include
#cgo LDFLAGS: -L${SRCDIR}/../../cron_operator_b200/csrc -lb200cron
#include <stdlib.h>
#include "b200cron.h"
*/
import "C"

import (
	"errors"
	"fmt"
	"math"
	"os"
	"runtime"
	"strings"
	"time"
	"unsafe"
)

// ZeroTime is how the ABI spells Go's time.Time{}.
const ZeroTime = math.MinInt64

// Spec mirrors robfig's SpecSchedule; Zone is the id of its Location (0 = UTC).
type Spec struct {
	Second, Minute, Hour, Dom, Month, Dow uint64
	Zone                                  int32
	Every                                 int64 // > 0: ConstantDelaySchedule (seconds)
}

// Engine owns a b200cron_ctx.
type Engine struct {
	ctx   *C.b200cron_ctx
	local int32 // zone id standing in for time.Local
}

// NewEngine loads the zone table on the given CUDA devices (nil = current device).
func NewEngine(devices []int, zoneinfoDir string) (*Engine, error) {
	var ctx *C.b200cron_ctx
	var devp *C.int
	if len(devices) > 0 {
		d := make([]C.int, len(devices))
		for i, v := range devices {
			d[i] = C.int(v)
		}
		devp = &d[0]
	}
	var dir *C.char
	if zoneinfoDir != "" {
		dir = C.CString(zoneinfoDir)
		defer C.free(unsafe.Pointer(dir))
	}
	if rc := C.b200cron_ctx_create(devp, C.int(len(devices)), dir, 0, &ctx); rc != 0 {
		return nil, fmt.Errorf("b200cron: %s", C.GoString(C.b200cron_last_error(nil)))
	}
	e := &Engine{ctx: ctx}
	// time.Local.String() is "Local" whenever Go loaded the zone from /etc/localtime (which is how the chart's
	// useHostTimezone mounts it), so the IANA name has to be found the way Go's own zoneinfo_unix.go finds the data:
	// $TZ first, then /etc/localtime.  A host zone that is not UTC and cannot be named is an error, never UTC.
	name := localZoneName()
	if id := e.ZoneID(name); name != "" && id >= 0 {
		e.local = id
	} else if _, off := time.Now().Zone(); off != 0 || !isUTCAllYear() {
		e.Close()
		return nil, fmt.Errorf("b200cron: cannot map time.Local to an IANA zone (TZ=%q, /etc/localtime -> %q)",
			os.Getenv("TZ"), name)
	}
	runtime.SetFinalizer(e, func(e *Engine) { e.Close() })
	return e, nil
}

// localZoneName returns the IANA name behind time.Local: $TZ (":Name" or "Name"), else the target of the
// /etc/localtime symlink below a zoneinfo directory; "" when neither names a zone (a copied file, for instance).
func localZoneName() string {
	if tz, ok := os.LookupEnv("TZ"); ok {
		tz = strings.TrimPrefix(tz, ":")
		if tz == "" {
			return "UTC" // Go: TZ="" means UTC
		}
		if i := strings.Index(tz, "zoneinfo/"); strings.HasPrefix(tz, "/") && i >= 0 {
			return tz[i+len("zoneinfo/"):]
		}
		return tz
	}
	if target, err := os.Readlink("/etc/localtime"); err == nil {
		if i := strings.Index(target, "zoneinfo/"); i >= 0 {
			return target[i+len("zoneinfo/"):]
		}
	}
	return ""
}

// isUTCAllYear reports whether time.Local has offset 0 at both solstices (an unnamed zone that is UTC in effect).
func isUTCAllYear() bool {
	y := time.Now().Year()
	for _, m := range []time.Month{time.January, time.July} {
		if _, off := time.Date(y, m, 1, 0, 0, 0, 0, time.Local).Zone(); off != 0 {
			return false
		}
	}
	return true
}

func (e *Engine) Close() {
	if e.ctx != nil {
		C.b200cron_ctx_destroy(e.ctx)
		e.ctx = nil
	}
}

func (e *Engine) err(rc C.int) error {
	if rc == 0 {
		return nil
	}
	return errors.New("b200cron: " + C.GoString(C.b200cron_last_error(e.ctx)))
}

// ZoneID maps an IANA name to the id used in Spec.Zone (-1: unknown, as time.LoadLocation would fail).
func (e *Engine) ZoneID(name string) int32 {
	cs := C.CString(name)
	defer C.free(unsafe.Pointer(cs))
	return int32(C.b200cron_zone_id(e.ctx, cs))
}

// ParseStandard replaces cronv3.ParseStandard (internal/controller/cron_controller.go:392).
func (e *Engine) ParseStandard(spec string) (*Schedule, error) {
	var out C.b200cron_spec
	var kind C.int
	var every C.int64_t
	tz := make([]byte, 128)
	msg := make([]byte, 512)
	cs := C.CString(spec)
	defer C.free(unsafe.Pointer(cs))
	rc := C.b200cron_parse_standard(cs, C.size_t(len(spec)), &out, &kind, &every,
		(*C.char)(unsafe.Pointer(&tz[0])), C.size_t(len(tz)), (*C.char)(unsafe.Pointer(&msg[0])), C.size_t(len(msg)))
	if rc != 0 {
		return nil, errors.New(C.GoString((*C.char)(unsafe.Pointer(&msg[0]))))
	}
	s := Spec{Second: uint64(out.second), Minute: uint64(out.minute), Hour: uint64(out.hour),
		Dom: uint64(out.dom), Month: uint64(out.month), Dow: uint64(out.dow), Zone: e.local}
	if kind == C.B200CRON_KIND_EVERY {
		s.Every = int64(every)
	}
	if name := C.GoString((*C.char)(unsafe.Pointer(&tz[0]))); name != "" && name != "Local" {
		id := e.ZoneID(name)
		if id < 0 {
			return nil, fmt.Errorf("provided bad location %s: unknown time zone %s", name, name)
		}
		s.Zone = id
	}
	return &Schedule{e: e, Spec: s}, nil
}

// Schedule satisfies robfig's cron.Schedule.
type Schedule struct {
	e *Engine
	Spec
}

// Next replaces (*SpecSchedule).Next (cron_controller.go:405,409,436).  A single evaluation pays a
// kernel launch; batch with Engine.NextN whenever more than a handful of Crons are due.
func (s *Schedule) Next(t time.Time) time.Time {
	if s.Every > 0 {
		return t.Add(time.Duration(s.Every)*time.Second - time.Duration(t.Nanosecond()))
	}
	out, err := s.e.NextN([]Spec{s.Spec}, []int64{t.Unix()}, 1)
	if err != nil || out[0][0] == ZeroTime {
		return time.Time{}
	}
	return time.Unix(out[0][0], 0).In(t.Location())
}

func planes(specs []Spec) (p [6][]uint64, zone []int32) {
	n := len(specs)
	for i := range p {
		p[i] = make([]uint64, n)
	}
	zone = make([]int32, n)
	for i, s := range specs {
		p[0][i], p[1][i], p[2][i], p[3][i], p[4][i], p[5][i] = s.Second, s.Minute, s.Hour, s.Dom, s.Month, s.Dow
		zone[i] = s.Zone
	}
	return
}

// NextN returns, per spec, the k chained fire times after from[i] (ZeroTime once none is found).
func (e *Engine) NextN(specs []Spec, from []int64, k int) ([][]int64, error) {
	n := len(specs)
	if n == 0 {
		return nil, nil
	}
	p, zone := planes(specs)
	flat := make([]int64, n*k)
	ptrs := (*[6]*C.uint64_t)(C.malloc(C.size_t(6 * unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(ptrs))
	var pin runtime.Pinner
	defer pin.Unpin()
	for i := range p {
		pin.Pin(&p[i][0])
		ptrs[i] = (*C.uint64_t)(unsafe.Pointer(&p[i][0]))
	}
	rc := C.b200cron_next_n(e.ctx, (**C.uint64_t)(unsafe.Pointer(ptrs)), (*C.int32_t)(unsafe.Pointer(&zone[0])),
		(*C.int64_t)(unsafe.Pointer(&from[0])), C.size_t(n), C.int(k), (*C.int64_t)(unsafe.Pointer(&flat[0])))
	if err := e.err(rc); err != nil {
		return nil, err
	}
	out := make([][]int64, n)
	for i := range out {
		out[i] = flat[i*k : (i+1)*k]
	}
	return out, nil
}

// Result is getNextSchedule's outcome plus Reconcile's decision for one Cron.
type Result struct {
	LastMissed, Next int64 // ZeroTime when unset
	Missed           uint32
	Status, Decision uint8
}

// UnmetSince replaces getNextSchedule (cron_controller.go:389-437) and the gates + concurrency
// decision of Reconcile (:163-220) for a batch of Crons.  policy: 0 Allow, 1 Forbid, 2 Replace;
// flags: 1 suspend, 2 deleted; deadline: ZeroTime for nil.
func (e *Engine) UnmetSince(specs []Spec, earliest []int64, now time.Time, policy []uint8, active []uint32,
	flags []uint8, deadline []int64) ([]Result, error) {
	n := len(specs)
	if n == 0 {
		return nil, nil
	}
	p, zone := planes(specs)
	last, next := make([]int64, n), make([]int64, n)
	missed := make([]uint32, n)
	status, decision := make([]uint8, n), make([]uint8, n)
	ptrs := (*[6]*C.uint64_t)(C.malloc(C.size_t(6 * unsafe.Sizeof(uintptr(0)))))
	defer C.free(unsafe.Pointer(ptrs))
	var pin runtime.Pinner
	defer pin.Unpin()
	for i := range p {
		pin.Pin(&p[i][0])
		ptrs[i] = (*C.uint64_t)(unsafe.Pointer(&p[i][0]))
	}
	opt8 := func(s []uint8) *C.uint8_t {
		if len(s) == 0 {
			return nil
		}
		return (*C.uint8_t)(unsafe.Pointer(&s[0]))
	}
	var act *C.uint32_t
	if len(active) > 0 {
		act = (*C.uint32_t)(unsafe.Pointer(&active[0]))
	}
	var dl *C.int64_t
	if len(deadline) > 0 {
		dl = (*C.int64_t)(unsafe.Pointer(&deadline[0]))
	}
	rc := C.b200cron_unmet_since(e.ctx, (**C.uint64_t)(unsafe.Pointer(ptrs)), (*C.int32_t)(unsafe.Pointer(&zone[0])),
		(*C.int64_t)(unsafe.Pointer(&earliest[0])), C.int64_t(now.Unix()), C.int32_t(now.Nanosecond()),
		opt8(policy), act, opt8(flags), dl, C.size_t(n),
		(*C.int64_t)(unsafe.Pointer(&last[0])), (*C.int64_t)(unsafe.Pointer(&next[0])),
		(*C.uint32_t)(unsafe.Pointer(&missed[0])), (*C.uint8_t)(unsafe.Pointer(&status[0])),
		(*C.uint8_t)(unsafe.Pointer(&decision[0])))
	if err := e.err(rc); err != nil {
		return nil, err
	}
	out := make([]Result, n)
	for i := range out {
		out[i] = Result{last[i], next[i], missed[i], status[i], decision[i]}
	}
	return out, nil
}

// HistoryPrune replaces sortByCreationTimestamp (cron_util.go:117-129) + the prune rule of
// syncCronHistory (cron_controller.go:313-323): per segment, the stable ascending order and keep flags.
func (e *Engine) HistoryPrune(segOff []uint64, created []int64, limit []int64) (perm []uint32, keep []uint8, err error) {
	m := len(created)
	perm, keep = make([]uint32, m+1), make([]uint8, m+1)
	var kp *C.int64_t
	if m > 0 {
		kp = (*C.int64_t)(unsafe.Pointer(&created[0]))
	}
	var lp *C.int64_t
	if len(limit) > 0 {
		lp = (*C.int64_t)(unsafe.Pointer(&limit[0]))
	}
	rc := C.b200cron_history_prune(e.ctx, (*C.uint64_t)(unsafe.Pointer(&segOff[0])), kp, lp, C.size_t(len(segOff)-1),
		(*C.uint32_t)(unsafe.Pointer(&perm[0])), (*C.uint8_t)(unsafe.Pointer(&keep[0])))
	return perm[:m], keep[:m], e.err(rc)
}
