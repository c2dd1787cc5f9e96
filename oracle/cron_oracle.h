/*
 * cron_oracle.h -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * A plain-C restatement of the reference's schedule-evaluation path:
 *   - internal/controller/cron_controller.go:389-437  (getNextSchedule)
 *   - internal/controller/cron_controller.go:163-220  (gates + concurrency decision)
 *   - internal/controller/cron_controller.go:307-346, cron_util.go:117-129 (history sort + prune)
 * and of the third-party arithmetic those lines call, which is NOT under /root/reference:
 *   - github.com/robfig/cron/v3 v3.0.1 (go.mod:10): parser.go, spec.go, constantdelay.go
 *   - Go stdlib `time` (go 1.25.5, go.mod:3): Date, AddDate, Add, Truncate, field getters,
 *     Location.lookup, tzset (TZif footer rule), LoadLocationFromTZData.
 *
 * PARITY STATUS: the reference's own known-answer tests K1-K4 (cron_controller_test.go:170-222,
 * cron_util_test.go:233-250) are reproduced (tests/test_oracle_golden.py).  Everything else --
 * time zones, DST edges, DOM/DOW OR rule, star-bit clearing, seconds field, 5-year limit, prune
 * indices, tie stability -- is PARITY UNPINNED: no Go toolchain exists on this box, so the
 * authority is the published upstream algorithm as restated here.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * link or call this library.  The product (libb200cron.so) never does.
 */
#ifndef CRON_ORACLE_H
#define CRON_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Go's zero time.Time{} is represented as INT64_MIN at every boundary of this repo. */
#define ORC_ZERO INT64_MIN

/* robfig/cron v3 spec.go: SpecSchedule{Second, Minute, Hour, Dom, Month, Dow uint64}; bit 63 = starBit */
typedef struct {
  uint64_t second, minute, hour, dom, month, dow;
} orc_spec;

enum { ORC_KIND_SPEC = 0, ORC_KIND_EVERY = 1 };

/* A ConstantDelaySchedule (constantdelay.go, what ParseStandard returns for "@every <dur>") travels through the
 * batched entry points in the same six planes: month == ORC_EVERY_TAG (a bit no parser sets) marks it and dom holds
 * Delay in whole seconds; the other four planes are ignored.  The parsers fill *out this way for @every. */
#define ORC_EVERY_TAG (1ULL << 62)

typedef struct orc_loc orc_loc; /* restatement of Go's time.Location */

/* time.LoadLocation(name) reading TZif files below `zoneinfo_dir` (Go: zoneinfo_read.go). */
orc_loc* orc_load_location(const char* zoneinfo_dir, const char* name, char* err, size_t errcap);
void orc_loc_free(orc_loc* loc);
/* (*Location).lookup(sec): offset east of UTC in force at `sec`, and the [start,end) it is valid for. */
void orc_loc_lookup(const orc_loc* loc, int64_t sec, int32_t* offset, int64_t* start, int64_t* end);
/* time.Date(y, mo, d, h, mi, s, 0, loc).Unix(); loc == NULL means UTC. */
int64_t orc_go_date(int64_t y, int64_t mo, int64_t d, int64_t h, int64_t mi, int64_t s, const orc_loc* loc);
/* civil fields of unix second `t` in loc: out[0..6] = year, month, day, hour, minute, second, weekday */
void orc_fields(const orc_loc* loc, int64_t t, int64_t out[7]);

/* cronv3.ParseStandard(spec) (parser.go).  Returns 0 on success.  `tz` receives the TZ=/CRON_TZ= name
 * ("" when absent, i.e. Location == time.Local).  For "@every <dur>" *kind = ORC_KIND_EVERY and
 * *every_sec is ConstantDelaySchedule.Delay in seconds. */
int orc_parse_standard(const char* s, size_t len, orc_spec* out, int* kind, int64_t* every_sec,
                       char* tz, size_t tzcap, char* err, size_t errcap);
/* extension following upstream NewParser(Second|Minute|Hour|Dom|Month|Dow|Descriptor): 6 fields */
int orc_parse_with_seconds(const char* s, size_t len, orc_spec* out, int* kind, int64_t* every_sec,
                           char* tz, size_t tzcap, char* err, size_t errcap);

/* (*SpecSchedule).Next(t) (spec.go), literal step walk; loc == NULL means UTC. ORC_ZERO = time.Time{} */
int64_t orc_spec_next(const orc_spec* s, const orc_loc* loc, int64_t t);
/* number of loop-body iterations the last orc_spec_next on this thread executed */
int64_t orc_last_steps(void);
/* Schedule.Next(t) for either kind: SpecSchedule (orc_spec_next) or ConstantDelaySchedule
 * (constantdelay.go Next: t.Add(Delay - t.Nanosecond()); whole seconds here, Delay clamped to >= 1s as Every() does) */
int64_t orc_sched_next(const orc_spec* s, const orc_loc* loc, int64_t t);

/* getNextSchedule (cron_controller.go:389-437) on an already-parsed spec.
 * status: 0 ok, 1 "unschedulable cron" error, 2 earliest.After(now) early return, 3 upstream's loop would never
 * terminate (time went backwards across a DST overlap; see cron_oracle.c). */
enum { ORC_ST_OK = 0, ORC_ST_UNSCHEDULABLE = 1, ORC_ST_EARLIEST_AFTER_NOW = 2, ORC_ST_NONTERMINATING = 3 };
int orc_get_next_schedule(const orc_spec* s, const orc_loc* loc, int64_t earliest, int64_t now,
                          int64_t* last_missed, int64_t* next, uint64_t* missed);

/* Reconcile gates + concurrency decision (cron_controller.go:163-220). */
enum {
  ORC_DEC_GATE_DELETED = 1,
  ORC_DEC_GATE_SUSPENDED = 2,
  ORC_DEC_GATE_DEADLINE = 3,
  ORC_DEC_ERR_UNSCHEDULABLE = 4,
  ORC_DEC_REQUEUE_ONLY = 5,
  ORC_DEC_SKIP_FORBID = 6,
  ORC_DEC_REPLACE_CREATE = 7,
  ORC_DEC_CREATE = 8
};
enum { ORC_POLICY_ALLOW = 0, ORC_POLICY_FORBID = 1, ORC_POLICY_REPLACE = 2 };
enum { ORC_FLAG_SUSPEND = 1, ORC_FLAG_DELETED = 2 };
/* deadline == ORC_ZERO means Spec.Deadline == nil */
int orc_reconcile_decide(const orc_spec* s, const orc_loc* loc, int64_t earliest, int64_t now, int32_t now_nsec,
                         uint8_t policy, uint32_t active, uint8_t flags, int64_t deadline,
                         int64_t* last_missed, int64_t* next, uint64_t* missed, uint8_t* status);

/* sortByCreationTimestamp (cron_util.go:117-129) + prune indices of syncCronHistory
 * (cron_controller.go:313-323) for one segment of `n` records.  perm[i] = original index of the i-th
 * record in ascending stable order; keep[i] = 1 when sorted position i is kept (i >= n - limit).
 * limit < 0 means HistoryLimit == nil (keep all). */
void orc_history_prune(const int64_t* key, size_t n, int64_t limit, uint32_t* perm, uint8_t* keep);

/* ---- batched drivers used by tests and by the timed CPU baseline (pthread, contiguous shards) ---- */
/* zone table: locs[zone_id[i]]; zone_id == NULL or locs == NULL means UTC for all */
void orc_batch_next_n(const uint64_t* const masks[6], const int32_t* zone_id, const orc_loc* const* locs,
                      const int64_t* from, size_t n, int k, int64_t* out, int threads);
void orc_batch_unmet(const uint64_t* const masks[6], const int32_t* zone_id, const orc_loc* const* locs,
                     const int64_t* earliest, int64_t now, int32_t now_nsec, const uint8_t* policy,
                     const uint32_t* active, const uint8_t* flags, const int64_t* deadline, size_t n,
                     int64_t* last, int64_t* next, uint32_t* missed, uint8_t* status, uint8_t* decision,
                     int threads);
uint64_t orc_batch_unmet_list(const uint64_t* const masks[6], const int32_t* zone_id, const orc_loc* const* locs,
                              const int64_t* earliest, const uint32_t* missed, const uint8_t* decision, size_t n,
                              uint32_t cap, uint64_t* list_off, int64_t* list, size_t list_cap);
void orc_batch_history(const uint64_t* seg_off, const int64_t* key, const int64_t* limit, size_t nseg,
                       uint32_t* perm, uint8_t* keep, int threads);

#ifdef __cplusplus
}
#endif
#endif
