/*
 * b200cron.h -- C ABI of libb200cron.so: batched cron-schedule evaluation on NVIDIA B200 (sm_100a).
 *
 * Drop-in boundary for ONE path of AliyunContainerService/cron-operator: the schedule arithmetic that
 * internal/controller/cron_controller.go reaches through `cron.Schedule` (robfig/cron/v3 v3.0.1, go.mod:10).
 * Each entry point cites the reference interface it replaces.  A cgo shim (go/b200cron) binds
 * exactly these symbols; so do the C++/Python hosts in this repo.
 *
 * Conventions
 *  - All times are int64 Unix seconds.  Go's zero time.Time{} is B200CRON_ZERO_TIME (INT64_MIN).
 *  - Schedules are robfig SpecSchedule bit masks (bit i = value i allowed, bit 63 = starBit) in SoA form: six
 *    planes second, minute, hour, dom, month, dow, each `n` uint64.
 *  - zone_id[i] is the location the schedule is evaluated in: the TZ=/CRON_TZ= location when the spec has one, else
 *    the id of the process's time.Local (0 = UTC).  NULL means UTC for every spec.  Ids come from b200cron_zone_id.
 *  - Functions return 0 on success or a negative B200CRON_E* code; b200cron_last_error(ctx) has the text.
 *  - The caller owns every buffer; the library owns only the context (streams, device staging, zone table).
 *  - Host-pointer entry points copy H2D/D2H internally (pinned host memory recommended: b200cron_host_alloc) and
 *    shard `n` contiguously over the context's devices.  `_dev` entry points take device pointers that live on
 *    device `dev_index` of the context and only enqueue work on `stream` (NULL = the context's stream).
 *  - Thread-safe per context (one call at a time per context; calls set the CUDA device explicitly).
 *  - There is no CPU fallback: without a usable GPU b200cron_ctx_create fails.
 *  - Supported times: years 1..9999 for UTC; zoned evaluations must start below (zone horizon - 6 years), the
 *    horizon being the year up to which DST footer rules are expanded (default 2200).  Out-of-range elements get
 *    B200CRON_ZERO_TIME and the call returns B200CRON_ERANGE.
 */
#ifndef B200CRON_H
#define B200CRON_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200CRON_ZERO_TIME INT64_MIN

enum {
  B200CRON_OK = 0,
  B200CRON_EINVAL = -1,  /* bad argument */
  B200CRON_ECUDA = -2,   /* CUDA runtime error (text in last_error) */
  B200CRON_ERANGE = -3,  /* some element outside the supported time range or with an unknown zone id */
  B200CRON_EPARSE = -4,  /* b200cron_parse_*: the schedule text is not parsable */
  B200CRON_ENOMEM = -5
};

/* per-element status of b200cron_unmet_since (getNextSchedule outcome, cron_controller.go:389-437) */
enum {
  B200CRON_ST_OK = 0,
  B200CRON_ST_UNSCHEDULABLE = 1,     /* Next() hit the zero time: "unschedulable cron" (:410-412) */
  B200CRON_ST_EARLIEST_AFTER_NOW = 2, /* earliestTime.After(now): only `next` is set (:404-406) */
  B200CRON_ST_NONTERMINATING = 3      /* upstream's loop :409 would cycle forever (chained Next() went backwards across a
                                         DST overlap); reported with decision ERR_UNSCHEDULABLE instead of hanging */
};
/* per-element decision of b200cron_unmet_since (Reconcile, cron_controller.go:163-238) */
enum {
  B200CRON_DEC_GATE_DELETED = 1,      /* :163-166 DeletionTimestamp set */
  B200CRON_DEC_GATE_SUSPENDED = 2,    /* :169-173 */
  B200CRON_DEC_GATE_DEADLINE = 3,     /* :176-180 now.After(deadline) */
  B200CRON_DEC_ERR_UNSCHEDULABLE = 4, /* :185-190 error swallowed, no requeue */
  B200CRON_DEC_REQUEUE_ONLY = 5,      /* :196-199 nothing missed: RequeueAfter = next - now */
  B200CRON_DEC_SKIP_FORBID = 6,       /* :204-207 */
  B200CRON_DEC_REPLACE_CREATE = 7,    /* :210-220 delete every active workload, then create */
  B200CRON_DEC_CREATE = 8             /* :222-238 create the workload named <cron>-<next> */
};
enum { B200CRON_POLICY_ALLOW = 0, B200CRON_POLICY_FORBID = 1, B200CRON_POLICY_REPLACE = 2 }; /* api/v1alpha1/cron_types.go:127-139 */
enum { B200CRON_FLAG_SUSPEND = 1, B200CRON_FLAG_DELETED = 2 };

enum { B200CRON_KIND_SPEC = 0, B200CRON_KIND_EVERY = 1 };

/* "@every <dur>" parses to robfig's ConstantDelaySchedule (constantdelay.go: Next(t) = t + Delay on whole seconds), which
 * the reference reaches through the same cron.Schedule interface (cron_controller.go:392,405,409,436).  In a batch it
 * travels in the same six planes: month == B200CRON_EVERY_TAG (a bit no parser sets) marks it and dom holds Delay in
 * whole seconds (values below 1 are read as 1, as upstream's Every() clamps); the other four planes and the zone id
 * are ignored.  b200cron_parse_* fill *out this way for KIND_EVERY, so next_n / unmet_since / unmet_list take mixed
 * batches of both kinds. */
#define B200CRON_EVERY_TAG (1ULL << 62)

/* robfig/cron v3 spec.go SpecSchedule (the Location is carried separately as a zone name / id) */
typedef struct b200cron_spec {
  uint64_t second, minute, hour, dom, month, dow;
} b200cron_spec;

typedef struct b200cron_ctx b200cron_ctx;

/* ---- host-only: parsing ------------------------------------------------------------------------------------ */

/* Replaces cronv3.ParseStandard(spec) [robfig/cron v3.0.1 parser.go], called at cron_controller.go:392.
 * 5 fields (minute hour dom month dow), descriptors (@yearly ... @hourly, @every <dur>), optional TZ=/CRON_TZ=
 * prefix.  On success fills *out, *kind (B200CRON_KIND_*), *every_sec (ConstantDelaySchedule.Delay, seconds) and
 * `tz` with the prefix's location name ("" when absent, i.e. Location == time.Local).  On failure returns
 * B200CRON_EPARSE and writes upstream's error text to `err` (the reference wraps it as "unparsable cron %q: %w",
 * cron_controller.go:394).  Pure host code: needs no context and no GPU. */
int b200cron_parse_standard(const char* spec, size_t len, b200cron_spec* out, int* kind, int64_t* every_sec,
                            char* tz, size_t tzcap, char* err, size_t errcap);
/* Extension (not reachable from the reference, whose parser is 5-field): same grammar with a leading seconds
 * field, i.e. upstream's cron.NewParser(Second|Minute|Hour|Dom|Month|Dow|Descriptor). */
int b200cron_parse_with_seconds(const char* spec, size_t len, b200cron_spec* out, int* kind, int64_t* every_sec,
                                char* tz, size_t tzcap, char* err, size_t errcap);

/* Batched form of b200cron_parse_standard / _with_seconds, run on the GPU (one thread per spec): the reference
 * re-parses every Cron's schedule on every reconcile (cron_controller.go:392).  Spec i is the bytes
 * text[text_off[i] .. text_off[i+1]).  Outputs per spec: the six masks (SoA planes, zero on error), status
 * (0 parsed -- "@every" with a duration made of whole h/m/s terms, e.g. "@every 1h30m", comes back in the
 * B200CRON_EVERY_TAG encoding; 1 not parsable -- call b200cron_parse_standard on it for upstream's message; 2 an
 * "@every" whose duration has a sign, a fraction, a sub-second unit or more than 9 digits in a term: left to the
 * host parser, which is time.ParseDuration to the letter), and the span of the TZ=/CRON_TZ= location name inside the
 * spec (tz_len 0 = no prefix; tz_off 0xFFFFFFFF = an empty name, i.e. UTC) for the caller to turn into a zone id
 * with b200cron_zone_id.  Same accept/reject decisions as the host parser.  The host-pointer form runs on the
 * context's first device; the _dev form takes device pointers.  Declared below the context functions' types. */
int b200cron_parse_batch(b200cron_ctx* ctx, const char* text, const uint64_t* text_off, size_t n, int with_seconds,
                         uint64_t* const masks_out[6], uint8_t* status, uint32_t* tz_off, uint32_t* tz_len);
int b200cron_parse_batch_dev(b200cron_ctx* ctx, int dev_index, void* stream, const char* text,
                             const uint64_t* text_off, size_t n, int with_seconds, uint64_t* const masks_out[6],
                             uint8_t* status, uint32_t* tz_off, uint32_t* tz_len);

/* ---- context ----------------------------------------------------------------------------------------------- */

/* Creates a context on `ndev` CUDA devices (ids in `devices`; NULL/0 = current device only) and loads every TZif
 * zone below `zoneinfo_dir` (NULL = $ZONEINFO or /usr/share/zoneinfo) -- the batched stand-in for
 * time.LoadLocation, which the reference reaches through ParseStandard's TZ= handling and through time.Local
 * (charts/cron-operator values.yaml:70-71 `useHostTimezone`).  horizon_year <= 0 selects the default (2200). */
int b200cron_ctx_create(const int* devices, int ndev, const char* zoneinfo_dir, int horizon_year, b200cron_ctx** out);
void b200cron_ctx_destroy(b200cron_ctx* ctx);
/* text of the calling thread's most recent failed call into this library (kept per thread: Reconcile runs on up to
 * MaxConcurrentReconciles OS threads at once); ctx is unused and may be NULL */
const char* b200cron_last_error(const b200cron_ctx* ctx);
int b200cron_device_count(const b200cron_ctx* ctx);
/* zone name -> id usable in zone_id[] (0 for "" and "UTC"); -1 when the name is unknown (LoadLocation error) */
int b200cron_zone_id(const b200cron_ctx* ctx, const char* name);
int b200cron_zone_count(const b200cron_ctx* ctx);
const char* b200cron_zone_name(const b200cron_ctx* ctx, int id);
/* read-only view of one zone's flattened periods: (*when)[i] <= t < (*when)[i+1] has UTC offset (*off)[i] */
int b200cron_zone_periods(const b200cron_ctx* ctx, int id, const int64_t** when, const int32_t** off);

/* pinned host memory helpers (cudaHostAlloc / cudaFreeHost); optional but recommended for the host-pointer calls */
void* b200cron_host_alloc(size_t bytes);
void b200cron_host_free(void* p);

/* ---- Next / NextN ------------------------------------------------------------------------------------------ */

/* Replaces (*SpecSchedule).Next [robfig/cron v3.0.1 spec.go] as called at cron_controller.go:405,409,436, batched:
 * out[i*k + j] = the (j+1)-th fire time after from[i] (each the Next() of the previous one), or
 * B200CRON_ZERO_TIME from the first call that finds no fire time within upstream's 5-year search.
 * masks[p] points to plane p (n uint64 each).  masks[0] (seconds) may be NULL: every spec then fires at second 0, which
 * is what ParseStandard's 5-field grammar always produces -- 8 bytes per spec less to move (b200cron_unmet_since and
 * the list calls accept the same).  When every buffer of the
 * host-pointer form is pinned host memory the kernels work on it in place over PCIe (no staging copies). */
int b200cron_next_n(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id, const int64_t* from,
                    size_t n, int k, int64_t* out);
int b200cron_next_n_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint64_t* const masks[6],
                        const int32_t* zone_id, const int64_t* from, size_t n, int k, int64_t* out);

/* Batches of specs are independent of each other (the reference evaluates every Cron in its own Reconcile call,
 * cron_controller.go:90), so back-to-back k == 1 launches need not wait for each other.  With on != 0 a
 * b200cron_next_n_dev launch (k == 1) may begin while the previous k == 1 launch on the same stream is still
 * draining its last wave of thread blocks (programmatic dependent launch).  The caller promises that such a launch
 * does not read what the previous one writes (e.g. does not pass its `out` as the next `from`) and that whatever
 * produced its inputs on that stream was followed by some other operation (an event record, a copy, any other
 * kernel).  Everything else -- k > 1, other entry points, other work queued on the stream -- still runs in stream
 * order after both have finished.  Off by default. */
int b200cron_ctx_set_overlap(b200cron_ctx* ctx, int on);

/* ---- unmet-schedule enumeration + concurrency decision -------------------------------------------------------- */

/* Replaces getNextSchedule (cron_controller.go:389-437) and the gates + Allow/Forbid/Replace decision of Reconcile
 * (:163-220) for n Crons at once.
 *   earliest[i] = Status.LastScheduleTime ?: CreationTimestamp (:397-402); now/now_nsec = time.Now() split into
 *   whole seconds and nanoseconds; policy[i] B200CRON_POLICY_*; active[i] = len(activeWorkloads) (:136-151);
 *   flags[i] B200CRON_FLAG_*; deadline[i] = Spec.Deadline or B200CRON_ZERO_TIME for nil.  policy, active, flags and
 *   deadline may be NULL (Allow, 0, 0, nil).
 * Outputs per Cron: last (lastMissed, ZERO when none), next (Next(now)), missed (missedTimes, saturating at
 * UINT32_MAX; the caller emits TooManyMissedTimes when > 100, :431-434), status, decision.  RequeueAfter is
 * next - now (:192).  Gated Crons get ZERO times and missed 0.
 * A call without zone ids is one kernel and uses no scratch.  With zone ids the Crons evaluated in UTC are answered in
 * place, the zoned ones through an index list grouped by zone, and whatever the closed form cannot follow through a
 * second list by the general walk: the lists live in device memory owned by the context (12 bytes per Cron of the
 * largest batch seen, three such buffers per device, allocated when a larger batch first arrives -- or up front by
 * b200cron_ctx_reserve_unmet; calls on different streams are kept apart by events). */
int b200cron_unmet_since(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id,
                         const int64_t* earliest, int64_t now, int32_t now_nsec, const uint8_t* policy,
                         const uint32_t* active, const uint8_t* flags, const int64_t* deadline, size_t n,
                         int64_t* last, int64_t* next, uint32_t* missed, uint8_t* status, uint8_t* decision);
/* Optional: allocate that scratch now, for batches of up to max_n Crons per device, so that no later unmet_since call
 * allocates or frees device memory in the middle of a tick (no counterpart in the reference). */
int b200cron_ctx_reserve_unmet(b200cron_ctx* ctx, size_t max_n);
int b200cron_unmet_since_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint64_t* const masks[6],
                             const int32_t* zone_id, const int64_t* earliest, int64_t now, int32_t now_nsec,
                             const uint8_t* policy, const uint32_t* active, const uint8_t* flags,
                             const int64_t* deadline, size_t n, int64_t* last, int64_t* next, uint32_t* missed,
                             uint8_t* status, uint8_t* decision);

/* Optional enumeration of the missed runs themselves.  The reference only counts them (cron_controller.go:409-430,
 * keeping the last one); callers that want to start every missed run need the times.  Given the `missed` and
 * `decision` arrays a previous b200cron_unmet_since call produced for the same Crons, writes
 *   list_off[i] (n+1 entries) = exclusive prefix sum of min(missed[i], cap) over Crons whose decision is
 *                                REQUEUE_ONLY / SKIP_FORBID / REPLACE_CREATE / CREATE (0 for gated or failed ones), and
 *   list[list_off[i] .. list_off[i+1]) = that Cron's first min(missed, cap) missed fire times in order.
 * cap <= 1,048,576 (larger values are rejected with B200CRON_EINVAL).  list_cap is the capacity of `list` in entries; when it is too small the call returns B200CRON_ENOMEM after
 * filling list_off (list_off[n] is the capacity needed).  The host-pointer variant runs on the context's first
 * device. */
int b200cron_unmet_list(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id,
                        const int64_t* earliest, const uint32_t* missed, const uint8_t* decision, size_t n,
                        uint32_t cap, uint64_t* list_off, int64_t* list, size_t list_cap);
/* Device-pointer variant in two steps so that the caller can size `list` in between: _offsets_dev fills list_off
 * (the context's scratch buffer holds the per-tile sums; it grows with cudaMalloc on first use or when n grows),
 * the caller reads list_off[n] after syncing the stream, then _emit_dev writes the fire times. */
int b200cron_unmet_list_offsets_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint32_t* missed,
                                    const uint8_t* decision, size_t n, uint32_t cap, uint64_t* list_off);
/* No counterpart in the reference (which never touches a GPU): b200cron_unmet_list_offsets_dev for hosts that
 * cannot let the library allocate device memory while it is being called -- for example a service that gave the whole
 * GPU to its own pool allocator at start-up, where a later cudaMalloc fails or fragments the pool, or a reconcile
 * loop that cannot afford the device-wide synchronisation a cudaMalloc/cudaFree causes while other streams are busy.
 * Same inputs, same list_off output, but the per-tile sums live in a fixed 128 KiB device array that is part of the
 * library image (every GPU that loads the library holds it until the process exits), so the call never allocates
 * or frees anything.
 *  - Limit: n <= 16,777,216 Crons per call; larger n is rejected with B200CRON_EINVAL (split the batch).
 *  - The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that first
 *    call before any stream capture begins.
 *  - Every call on a device shares the one array and finds the previous call's contents in it; each call overwrites
 *    the entries it reads, so nothing needs clearing, but two calls that overlap on the same GPU (different
 *    streams, threads or contexts) race on it -- run them one after the other. */
int b200cron_unmet_list_offsets_static(b200cron_ctx* ctx, int dev_index, void* stream, const uint32_t* missed,
                                       const uint8_t* decision, size_t n, uint32_t cap, uint64_t* list_off);
int b200cron_unmet_list_emit_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint64_t* const masks[6],
                                 const int32_t* zone_id, const int64_t* earliest, size_t n, const uint64_t* list_off,
                                 int64_t* list);

/* ---- resident spec sets --------------------------------------------------------------------------------------- */

/* Extension with no counterpart in the reference: a SpecSchedule is immutable once ParseStandard has built it
 * (cron_controller.go:392) and a Cron's schedule changes only when its CR does (api/v1alpha1/cron_types.go:77), so a
 * host that ticks the same Crons again and again need not ship 48 bytes of masks per Cron every time.  A spec set
 * keeps the six mask planes (masks[0] may be NULL: second 0) and the zone ids (NULL: all UTC) resident in device
 * memory, sharded contiguously over the context's devices like the host-pointer calls; the _set calls then move only
 * the per-tick arrays (8 bytes in and 8k bytes out per spec for next_n_set).  n is fixed at creation;
 * b200cron_specset_update overwrites specs [first, first + count) (masks[p] / zone_id point at the first replaced
 * spec; give exactly the planes the set was created with) and returns when the new values are in place.  A set
 * belongs to the context that created it and must be destroyed before it.  Host buffers of the _set calls follow the
 * same rule as everywhere: all pinned -> used in place over PCIe, otherwise staged copies. */
typedef struct b200cron_specset b200cron_specset;
int b200cron_specset_create(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id, size_t n,
                            b200cron_specset** out);
int b200cron_specset_update(b200cron_specset* set, size_t first, size_t count, const uint64_t* const masks[6],
                            const int32_t* zone_id);
void b200cron_specset_destroy(b200cron_specset* set);
size_t b200cron_specset_size(const b200cron_specset* set);
/* b200cron_next_n / b200cron_unmet_since over the set's specs, in the set's order (same outputs, same errors) */
int b200cron_next_n_set(b200cron_ctx* ctx, const b200cron_specset* set, const int64_t* from, int k, int64_t* out);
int b200cron_unmet_since_set(b200cron_ctx* ctx, const b200cron_specset* set, const int64_t* earliest, int64_t now,
                             int32_t now_nsec, const uint8_t* policy, const uint32_t* active, const uint8_t* flags,
                             const int64_t* deadline, int64_t* last, int64_t* next, uint32_t* missed, uint8_t* status,
                             uint8_t* decision);

/* ---- finished-history ordering + pruning ---------------------------------------------------------------------- */

/* Replaces sortByCreationTimestamp (cron_util.go:117-129) and the prune index rule of syncCronHistory
 * (cron_controller.go:313-323) for nseg Crons at once.  Segment s owns records [seg_off[s], seg_off[s+1]) of
 * key[] (creation timestamps; any int64 key works).  limit[s] = Spec.HistoryLimit, negative = nil (keep all);
 * limit == NULL keeps all.  Outputs, per record position p inside its segment's sorted order:
 *   perm[seg_off[s] + p] = index inside the segment of the record that sorts p-th (stable, ascending);
 *   keep[seg_off[s] + p] = 0 when that record is deleted (p < len - limit), 1 when it stays in Status.History.
 * Segments are expected to be short (the reference keeps HistoryLimit records per Cron): up to 2,048 records a
 * segment is ranked in shared memory, up to 65,536 by a quadratic pass over global memory, and a longer one is
 * rejected (B200CRON_EINVAL; the _dev form reports it at the next b200cron_sync_check and leaves that segment in
 * input order, everything kept). */
int b200cron_history_prune(b200cron_ctx* ctx, const uint64_t* seg_off, const int64_t* key, const int64_t* limit,
                           size_t nseg, uint32_t* perm, uint8_t* keep);
int b200cron_history_prune_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint64_t* seg_off,
                               const int64_t* key, const int64_t* limit, size_t nseg, size_t nrec, uint32_t* perm,
                               uint8_t* keep);

/* ---- synthetic workloads (benchmark / test support; no counterpart in the reference) ---------------------------
 * Device-side generators of the BASELINE.json workloads, bit-identical to cron_operator_b200/synth.py: every draw is
 * a counter-based hash of (seed, index, stream), so shards can be generated where they are consumed and any index
 * can be regenerated on the host for sampled parity.  `thresholds` = the ten doubles of synth.form_thresholds(0.40)
 * followed by synth.form_thresholds(0.60); outputs are device pointers on device dev_index. */
int b200cron_synth_specs_dev(b200cron_ctx* ctx, int dev_index, void* stream, uint64_t seed, uint64_t start, size_t n,
                             int with_seconds, double dense_minutes, double plant_unschedulable,
                             const double thresholds[10], uint64_t* const masks_out[6]);
int b200cron_synth_times_dev(b200cron_ctx* ctx, int dev_index, void* stream, uint64_t seed, uint64_t start, size_t n,
                             int64_t lo, int64_t hi, uint64_t stream_id, int64_t* out);
int b200cron_synth_zones_dev(b200cron_ctx* ctx, int dev_index, void* stream, uint64_t seed, uint64_t start, size_t n,
                             const int32_t* pool, size_t pool_len, double p_utc, int32_t* out);

/* Test support: the Crons the most recent unmet_since launch on device dev_index could not answer with the closed form
 * and handed to the general walk (indices relative to that launch's arrays; up to `cap` of them are copied to
 * idx_host, *count receives how many there were).  Waits for `stream` first.  Lets a parity test aim its oracle
 * samples at exactly the Crons that took the rare path. */
int b200cron_debug_last_deferred(b200cron_ctx* ctx, int dev_index, void* stream, uint32_t* idx_host, size_t cap,
                                 size_t* count);

/* `_dev` calls only enqueue work: this waits for `stream` (NULL = the context's stream) on device dev_index and
 * reports what the kernels flagged since the last check (B200CRON_ERANGE for out-of-range times / zone ids). */
int b200cron_sync_check(b200cron_ctx* ctx, int dev_index, void* stream);

/* number of kernels this context has launched so far (bench.py's gpu_launches) */
uint64_t b200cron_launch_count(const b200cron_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* B200CRON_H */
