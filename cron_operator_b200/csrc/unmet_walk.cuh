// unmet_walk.cuh -- per-spec missed-run enumeration + concurrency decision.
//
// Replaces getNextSchedule (internal/controller/cron_controller.go:389-437) and the gates + Allow/Forbid/Replace
// decision of Reconcile (:163-220) for one Cron.  The reference chains sched.Next() once per missed run (43,200
// calls for "* * * * *" over a 30-day backlog); here the chain is only walked call by call where a zone transition
// can perturb it, and stretches inside one constant-offset period are counted with popcounts (chain_fast_forward).
#pragma once
#include "cron_walk.cuh"
#include <limits.h>

namespace cronwalk {

enum : uint8_t { kStOk = 0, kStUnschedulable = 1, kStEarliestAfterNow = 2, kStNonTerminating = 3 };
enum : uint8_t {
  kDecGateDeleted = 1,       // cron_controller.go:163-166
  kDecGateSuspended = 2,     // :169-173
  kDecGateDeadline = 3,      // :176-180
  kDecErrUnschedulable = 4,  // :185-190  ("unschedulable cron" error: no requeue)
  kDecRequeueOnly = 5,       // :196-199  (no missed run)
  kDecSkipForbid = 6,        // :204-207
  kDecReplaceCreate = 7,     // :210-220 then :222-238
  kDecCreate = 8             // :222-238
};
enum : uint8_t { kPolicyAllow = 0, kPolicyForbid = 1, kPolicyReplace = 2 };
enum : uint8_t { kFlagSuspend = 1, kFlagDeleted = 2 };

struct UnmetIn {
  int64_t earliest;  // Status.LastScheduleTime ?: CreationTimestamp  (:397-402)
  int64_t now;       // floor(now) in unix seconds
  int32_t now_nsec;  // sub-second part of now; only the deadline gate can observe it
  uint8_t policy, flags;
  uint32_t active;
  int64_t deadline;  // kZero == Spec.Deadline is nil
};
struct UnmetOut {
  int64_t last, next;
  uint32_t missed;  // saturates at 0xffffffff
  uint8_t status, decision;
};

CW_HD uint32_t sat_add(uint32_t a, uint64_t b) {
  const uint64_t r = (uint64_t)a + b;
  return r > 0xffffffffULL ? 0xffffffffu : (uint32_t)r;
}

// number of matching times of day with second-of-day <= sod (sod in [0, 86399])
CW_HD uint32_t count_tod_le(const Spec& s, int32_t sod) {
  const int32_t h = sod / 3600, mi = (sod % 3600) / 60, sc = sod % 60;
  const uint32_t pm = (uint32_t)popc64(s.minute), psec = (uint32_t)popc64(s.second);
  uint32_t c = (uint32_t)popc32(s.hour & ((1u << h) - 1u)) * pm * psec;
  if ((s.hour >> h) & 1u) {
    c += (uint32_t)popc64(s.minute & ((1ULL << mi) - 1ULL)) * psec;
    if ((s.minute >> mi) & 1ULL) c += (uint32_t)popc64(s.second & ((2ULL << sc) - 1ULL));
  }
  return c;
}
// largest matching second-of-day <= sod, or -1
CW_HD int32_t last_tod_le(const Spec& s, int32_t sod) {
  const int32_t h = sod / 3600, mi = (sod % 3600) / 60, sc = sod % 60;
  const int32_t top_s = 63 - clz64(s.second), top_m = 63 - clz64(s.minute);
  if ((s.hour >> h) & 1u) {
    if ((s.minute >> mi) & 1ULL) {
      const uint64_t m = s.second & ((2ULL << sc) - 1ULL);
      if (m) return h * 3600 + mi * 60 + (63 - clz64(m));
    }
    const uint64_t m = s.minute & ((1ULL << mi) - 1ULL);
    if (m) return h * 3600 + (63 - clz64(m)) * 60 + top_s;
  }
  const uint32_t hm = s.hour & ((1u << h) - 1u);
  if (hm) return (31 - clz32(hm)) * 3600 + top_m * 60 + top_s;
  return -1;
}

// Fast-forward the chain of Next() calls from the fire time held in `w` towards `limit` (absolute, inclusive) as far
// as that is provably identical to upstream's one-call-per-fire chain; advances *missed and leaves `w` on the last
// fire counted (unchanged when nothing could be proven).  The caller continues with a literal Next() from there.
//
// Why it is exact.  From a fire time f, the chained Next(f) starts at f+1s in a matching month, day and hour.
//  (1) While fires remain before the next local midnight and before the end of the zone period, the walk only runs
//      the Add-based second/minute/hour loops: they visit absolute instants in order and test local fields, which
//      advance uniformly inside one period, so the chain enumerates the remaining matching times of day in order.
//      The one Date() call that can appear is the hour reset when f is hh:59:59 (f+1s is an exact hour start);
//      specs that can fire at :59:59 therefore also need every instant of the stretch to be Walk::natural.
//  (2) After the last fire of a day the walk wraps, always aligned, to the next local midnight (strictly inside the
//      period) and crosses non-matching days with AddDate steps at hour 0 (no DST fix) and months with the Day()==1
//      wrap; while all those midnights are natural in this period the chain visits exactly the matching days, and
//      each such day's fires are enumerated by (1).  The 5-year limit can only bite between two matching days: the
//      fast-forward stops before any day whose year exceeds (last fire year + 5) and lets the literal call decide.
// kUtcOnly: the walk is bound to UTC (one endless period, offset 0, everything natural): the zone logic compiles away.
template <bool kUtcOnly = false>
CW_HD void chain_fast_forward(const Spec& s, Walk& w, int64_t limit, uint32_t* missed) {
  // A Date() call inside a day can only be the hour reset, taken when the call starts on an exact hour start whose
  // hour does not match: the spec must be able to fire at :59:59 AND exclude some hour.
  const bool edge59 = !kUtcOnly && ((s.second >> 59) & 1ULL) && ((s.minute >> 59) & 1ULL) && (s.hour != 0xFFFFFFu);
  // A fire produced while a loop stepped across a transition can sit on fields upstream never re-checked (e.g. the
  // hour loop going 23h -> 01h over a midnight gap lands on a day that does not match).  The next call re-checks
  // them from scratch, so nothing may be assumed about such a fire: only fast-forward from a fully matching one.
  if (!(((s.month >> w.M) & 1u) && day_matches(s, w.D, w.dn) && ((s.hour >> w.hour()) & 1u) &&
        ((s.minute >> w.minute()) & 1ULL) && ((s.second >> w.second()) & 1ULL)))
    return;
  // ---- (1) rest of the current local day, up to the period edge ----
  const int64_t day_end = w.t + (86400 - w.sod);  // absolute instant of the next local midnight at this offset
  {
    int64_t lim_abs = limit < day_end - 1 ? limit : day_end - 1;
    if (!kUtcOnly && w.pe - 1 < lim_abs) lim_abs = w.pe - 1;
    bool hour_only = false;
    if (edge59 && !(w.natural(w.t + w.off) && w.natural(lim_abs + w.off))) {
      // Some hour start of the stretch may not resolve naturally.  No chained call starts on an hour start before the
      // current local hour is over, so the rest of THIS hour is still a plain enumeration; the literal call that
      // follows crosses the hour boundary the way upstream does (one literal call per hour instead of one per fire).
      const int64_t hour_end = w.t + (3599 - w.sod % 3600);
      if (hour_end < lim_abs) lim_abs = hour_end;
      hour_only = true;
    }
    const int32_t lim_sod = w.sod + (int32_t)(lim_abs - w.t);
    const int32_t last = last_tod_le(s, lim_sod);
    if (last > w.sod) {
      *missed = sat_add(*missed, count_tod_le(s, lim_sod) - count_tod_le(s, w.sod));
      w.t += last - w.sod;
      w.sod = last;
    }
    if (hour_only) return;
  }
  if (limit < day_end || (!kUtcOnly && day_end >= w.pe)) return;

  // ---- (2) whole following days whose midnights are natural in this period ----
  const int64_t kFar = (int64_t)1 << 46;  // ~2.2 million years: outside every supported input
  int32_t dhi = INT32_MAX / 2, dlo = INT32_MIN / 2;              // days d with natural(d * 86400): dlo <= d <= dhi
  if (!kUtcOnly) {
    const int64_t xmax = w.pe == kOmega ? kFar : w.pe - 1 + (w.off < 0 ? w.off : 0);
    const int64_t xmin = w.ps == kAlpha ? -kFar : w.ps + (w.off > 0 ? w.off : 0);
    int64_t dhi64 = xmax / 86400; if (xmax % 86400 < 0) --dhi64;   // floor
    int64_t dlo64 = xmin / 86400; if (xmin % 86400 > 0) ++dlo64;   // ceil
    dhi = (int32_t)dhi64; dlo = (int32_t)dlo64;
  }
  int32_t cur = w.dn + 1;
  if (cur < dlo || cur > dhi) return;

  const bool lim_inside = kUtcOnly || limit < w.pe;
  int32_t Ld = 0, Lsod = 0;
  if (lim_inside) {
    const int64_t L = limit + (kUtcOnly ? 0 : w.off);
    if (kUtcOnly) {  // |L| < 2^38 inside the supported range: floor-divide by 86400 = 128 * 675 in 32 bits
      const int32_t hi = (int32_t)(L >> 7);
      int32_t q = hi / 675, r = hi - q * 675;
      if (r < 0) { r += 675; q -= 1; }
      Ld = q; Lsod = r * 128 + (int32_t)(L & 127);
    } else {
      int64_t q = L / 86400; if (L % 86400 < 0) --q;
      Ld = (int32_t)q; Lsod = (int32_t)(L - q * 86400);
    }
  }
  int32_t full_last = dhi - 1;  // last day whose closing midnight is still natural
  if (lim_inside && Ld - 1 < full_last) full_last = Ld - 1;
  const bool partial = lim_inside && Ld <= dhi && (!edge59 || w.natural(limit + w.off));

  const uint32_t per_day = (uint32_t)popc32(s.hour) * (uint32_t)popc64(s.minute) * (uint32_t)popc64(s.second);
  int32_t Y = w.Y, M = w.M, D = w.D + 1;
  if (D > days_in_month(Y, M)) { D = 1; if (++M == 13) { M = 1; ++Y; } }
  int32_t fire_year = w.Y;
  uint64_t total = 0;
  int32_t last_dn = INT32_MIN, last_sod = 0;
  bool stopped = false;
  while (cur <= full_last) {
    const int32_t dim = days_in_month(Y, M);
    const int32_t dend = (full_last - cur) < (dim - D) ? D + (full_last - cur) : dim;  // last day of month in range
    if ((s.month >> M) & 1u) {
      const uint32_t range = (uint32_t)(((2ULL << dend) - 1ULL) & ~((1ULL << D) - 1ULL));
      const uint32_t dm = month_daymask(s, Y, M, cur - (D - 1)) & range;
      if (dm) {
        if (Y > fire_year + 5) { stopped = true; break; }
        total += (uint64_t)popc32(dm) * per_day;
        last_dn = cur - D + (31 - clz32(dm));
        fire_year = Y;
      }
    }
    cur += dend - D + 1;
    D = dend + 1;
    if (D > dim) { D = 1; if (++M == 13) { M = 1; ++Y; } }
  }
  if (last_dn != INT32_MIN) last_sod = last_tod_le(s, 86399);
  if (!stopped && partial && cur == Ld) {  // the day holding `limit`: fires with time of day <= Lsod
    if (((s.month >> M) & 1u) && day_matches(s, D, Ld) && Y <= fire_year + 5) {
      const uint32_t c = count_tod_le(s, Lsod);
      if (c) { total += c; last_dn = Ld; last_sod = last_tod_le(s, Lsod); }
    }
  }
  if (last_dn == INT32_MIN) return;
  *missed = sat_add(*missed, total);
  w.dn = last_dn;
  w.sod = last_sod;
  w.t = (int64_t)last_dn * 86400 + last_sod - (kUtcOnly ? 0 : w.off);
  civil_from_days(last_dn, w.Y, w.M, w.D);
  w.today_ok = true;  // last_dn came out of the month's day mask
  w.dm_ok = false;
}

// getNextSchedule + gates + decision for one Cron, as a small state machine with ONE Next() call site per step:
//   chain phase  -- :409  for t := sched.Next(earliest); !t.After(now); t = sched.Next(t)
//   final phase  -- :405 / :436  sched.Next(now)
// The kernel drives all 32 lanes of a warp through unmet_step() in lock step (`while (__any_sync(..., !done))`), which
// makes the lanes reconverge after every call; a plain per-lane loop with early exits left them diverged for good
// (5 of 32 lanes active) and an inlined Next() per call site made the kernel instruction-fetch bound.
//
// Upstream bug guard: across a DST overlap a chained Next() can return a time at or before its argument (e.g.
// Pacific/Chatham: a fire found at 02:59 on the repeated hour makes the next call reset to 02:00 and come back to
// the first 03:00), and then upstream's loop above cycles forever.  Next() is a pure function of t, so the loop
// is in a cycle exactly when it revisits a fire time: after every backward step the largest origin is remembered
// and a fire equal to it ends the call with kStNonTerminating instead of hanging.  No other input is affected.
struct UnmetState {
  Walk w;
  UnmetOut o;
  int64_t arg, anchor, prev;
  bool resident, final_phase, after_now, done;
};

CW_HD void unmet_finish(const UnmetIn& in, UnmetState& st) {
  st.done = true;
  if (st.after_now) {
    st.o.status = kStEarliestAfterNow;
    st.o.decision = kDecRequeueOnly;
  } else if (st.o.last == kZero) st.o.decision = kDecRequeueOnly;                            // :196-199
  else if (in.policy == kPolicyForbid && in.active > 0) st.o.decision = kDecSkipForbid;      // :204-207
  else if (in.policy == kPolicyReplace) st.o.decision = kDecReplaceCreate;                   // :210-220
  else st.o.decision = kDecCreate;
}

// Reconcile's gates (:163-180); returns the gate's decision code, or 0 when the Cron goes on to getNextSchedule
CW_HD uint8_t unmet_gate(const UnmetIn& in) {
  if (in.flags & kFlagDeleted) return kDecGateDeleted;
  if (in.flags & kFlagSuspend) return kDecGateSuspended;
  // now.After(deadline): now carries nanoseconds, the deadline is a whole second
  if (in.deadline != kZero && (in.now > in.deadline || (in.now == in.deadline && in.now_nsec > 0))) return kDecGateDeadline;
  return 0;
}

CW_HD void unmet_begin(const ZoneTable& zt, int32_t zone, const UnmetIn& in, UnmetState& st) {
  st.o.last = kZero; st.o.next = kZero; st.o.missed = 0; st.o.status = kStOk; st.o.decision = 0;
  st.done = true;
  st.w.bind_utc();
  st.o.decision = unmet_gate(in);
  if (st.o.decision) return;
  if (zone != 0) st.w.bind_zone(zt, zone);
  st.after_now = in.earliest > in.now;  // :404-406
  st.final_phase = st.after_now;
  st.arg = st.after_now ? in.now : in.earliest;
  st.resident = false;
  st.anchor = kZero;
  st.prev = kZero;
  st.done = false;
}

CW_HD void unmet_step(const Spec& s, const UnmetIn& in, UnmetState& st) {
  const int64_t t = next_from_t<true>(s, st.w, st.arg, st.resident);
  if (st.final_phase) {
    st.o.next = t;
    unmet_finish(in, st);
    return;
  }
  bool fail = false;
  uint8_t fail_status = kStUnschedulable;
  if (st.prev != kZero) {  // t is the Next() of the fire `prev`
    if (t != kZero && t <= st.prev && st.prev > st.anchor) st.anchor = st.prev;
    if (st.anchor != kZero && t == st.anchor) { fail = true; fail_status = kStNonTerminating; }
  }
  if (!fail && t == kZero) fail = true;  // :410-412
  if (fail) {
    st.o.last = kZero; st.o.missed = 0; st.o.status = fail_status; st.o.decision = kDecErrUnschedulable;
    st.done = true;
    return;
  }
  if (t > in.now) {  // the chain is over: one more call, from now
    st.final_phase = true; st.arg = in.now; st.resident = false;
    return;
  }
  st.o.missed = sat_add(st.o.missed, 1);
  if (st.anchor == kZero || st.w.t > st.anchor) chain_fast_forward(s, st.w, in.now, &st.o.missed);
  st.o.last = st.w.t;
  st.prev = st.w.t;
  st.arg = st.prev;
  st.resident = true;
}


// unmet_step for the "easy" kernel: same bookkeeping, but only the closed form may answer.  Whenever the step
// emulation would be needed, or a chained call steps backwards (the cycle guard's business), the Cron is handed to
// the general kernel (*defer) and recomputed there from scratch.
CW_HD void unmet_step_easy(const Spec& s, const UnmetIn& in, UnmetState& st, bool* defer) {
  int64_t t;
  if (!next_try(s, st.w, st.arg, st.resident, &t)) { *defer = true; st.done = true; return; }
  if (st.final_phase) {
    st.o.next = t;
    unmet_finish(in, st);
    return;
  }
  if (st.prev != kZero && t != kZero && t <= st.prev) { *defer = true; st.done = true; return; }
  if (t == kZero) {  // :410-412
    st.o.last = kZero; st.o.missed = 0; st.o.status = kStUnschedulable; st.o.decision = kDecErrUnschedulable;
    st.done = true;
    return;
  }
  if (t > in.now) {  // the chain is over: one more call, from now
    st.final_phase = true; st.arg = in.now; st.resident = false;
    return;
  }
  st.o.missed = sat_add(st.o.missed, 1);
  chain_fast_forward(s, st.w, in.now, &st.o.missed);
  st.o.last = st.w.t;
  st.prev = st.w.t;
  st.arg = st.prev;
  st.resident = true;
}

// getNextSchedule + gates + decision for a ConstantDelaySchedule ("@every <dur>", constantdelay.go), which the
// reference reaches through the same cron.Schedule interface (cron_controller.go:392,405,409,436).  Next(t) = t + delay
// on whole seconds, so the loop at :409 visits earliest + delay, earliest + 2 delay, ... <= now: a division.
CW_HD UnmetOut unmet_decide_every(int64_t delay, const UnmetIn& in) {
  UnmetOut o;
  o.last = kZero; o.next = kZero; o.missed = 0; o.status = kStOk;
  o.decision = unmet_gate(in);
  if (o.decision) return o;
  o.next = in.now + delay;  // :405 / :436
  if (in.earliest > in.now) {
    o.status = kStEarliestAfterNow;
    o.decision = kDecRequeueOnly;
    return o;
  }
  const uint64_t k = (uint64_t)(in.now - in.earliest) / (uint64_t)delay;
  o.missed = k > 0xffffffffULL ? 0xffffffffu : (uint32_t)k;
  if (k) o.last = in.earliest + (int64_t)(k * (uint64_t)delay);
  if (o.last == kZero) o.decision = kDecRequeueOnly;
  else if (in.policy == kPolicyForbid && in.active > 0) o.decision = kDecSkipForbid;
  else if (in.policy == kPolicyReplace) o.decision = kDecReplaceCreate;
  else o.decision = kDecCreate;
  return o;
}

// unmet_step for a Cron evaluated in UTC (zone id 0, or a launch without zone ids).  One endless period, every Date()
// natural: the closed form always answers, chained calls never step backwards, and the call at :436 is redundant --
// when the chain ends with cur <= now < t = Next(cur), Next(now) is that same t: in UTC Next(x) is the smallest matching
// instant after x inside x's five-year window, no instant of (cur, t) matches, and t lies inside now's window because
// it lies inside cur's (cur.Year() <= now.Year()).  So a Cron costs one Next() plus, when something was missed, one
// fast-forward and one more Next().
CW_HD void unmet_step_utc(const Spec& s, const UnmetIn& in, UnmetState& st) {
  const int64_t t = next_from_t<false, true>(s, st.w, st.arg, st.resident);
  if (st.final_phase) {  // earliest.After(now): Next(now) is the whole answer (:404-406)
    st.o.next = t;
    unmet_finish(in, st);
    return;
  }
  if (t == kZero) {  // :410-412
    st.o.last = kZero; st.o.missed = 0; st.o.status = kStUnschedulable; st.o.decision = kDecErrUnschedulable;
    st.done = true;
    return;
  }
  if (t > in.now) {
    st.o.next = t;
    unmet_finish(in, st);
    return;
  }
  st.o.missed = sat_add(st.o.missed, 1);
  chain_fast_forward<true>(s, st.w, in.now, &st.o.missed);
  st.o.last = st.w.t;
  st.arg = st.w.t;
  st.resident = true;
}

// ---- the UTC machine on one zone period's local clock (cron_walk.cuh: period_clock) ---------------------------------
// unmet_class_a: the Cron's window runs from min(earliest, now) to now
CW_HD bool unmet_class_a(const ZoneTable& zt, int32_t zone, int64_t earliest, int64_t now, int64_t t_min, int64_t t_max,
                         int64_t zoned_max, int32_t* shift, int64_t* nhi) {
  return period_clock(zt, zone, earliest < now ? earliest : now, now, t_min, t_max, zoned_max, shift, nhi);
}
// gates on the real clock, then `in` moves onto the local clock and the UTC machine starts
CW_HD void unmet_begin_shifted(const ZoneTable& zt, UnmetIn& in, int32_t shift, UnmetState& st) {
  st.o.last = kZero; st.o.next = kZero; st.o.missed = 0; st.o.status = kStOk;
  st.done = true;
  st.w.bind_utc();
  st.o.decision = unmet_gate(in);
  if (st.o.decision) return;
  in.earliest += shift; in.now += shift; in.deadline = kZero; in.flags = 0;
  unmet_begin(zt, 0, in, st);
}
// after the machine has run: true = some Next() left the period (or found nothing in five years, i.e. searched beyond
// it) and the answer must come from the general path; false = st.o holds the answer, moved back to the real clock
CW_HD bool unmet_finish_shifted(int32_t shift, int64_t nhi, UnmetState& st) {
  if (shift == 0 && nhi == kOmega) return false;  // a Cron evaluated in UTC: one endless period
  if (st.o.status == kStUnschedulable || (st.o.next != kZero && st.o.next >= nhi)) return true;
  if (st.o.last != kZero) st.o.last -= shift;
  if (st.o.next != kZero) st.o.next -= shift;
  return false;
}

// the kernels' class logic in one place (host build / tests): UTC machine for class A, general machine otherwise
CW_HD UnmetOut unmet_decide_classed(const Spec& s, const ZoneTable& zt, int32_t zone, const UnmetIn& in0, int64_t t_min,
                                    int64_t t_max, int64_t zoned_max, int* klass) {
  int32_t shift = 0;
  int64_t nhi = kOmega;
  *klass = 2;
  if (!s.every && (zone == 0 || unmet_class_a(zt, zone, in0.earliest, in0.now, t_min, t_max, zoned_max, &shift, &nhi))) {
    UnmetIn in = in0;
    UnmetState st;
    unmet_begin_shifted(zt, in, shift, st);
    while (!st.done) unmet_step_utc(s, in, st);
    if (!unmet_finish_shifted(shift, nhi, st)) { *klass = zone == 0 ? 0 : 1; return st.o; }
    *klass = 3;  // rejected: joins class B
  }
  if (s.every) return unmet_decide_every((int64_t)s.second, in0);
  UnmetState st;
  unmet_begin(zt, zone, in0, st);
  while (!st.done) unmet_step(s, in0, st);
  return st.o;
}

CW_HD UnmetOut unmet_decide_utc(const Spec& s, const ZoneTable& zt, const UnmetIn& in) {
  UnmetState st;
  unmet_begin(zt, 0, in, st);
  while (!st.done) unmet_step_utc(s, in, st);
  return st.o;
}

CW_HD UnmetOut unmet_decide(const Spec& s, const ZoneTable& zt, int32_t zone, const UnmetIn& in) {
  if (s.every) return unmet_decide_every((int64_t)s.second, in);
  UnmetState st;
  unmet_begin(zt, zone, in, st);
  while (!st.done) unmet_step(s, in, st);
  return st.o;
}

}  // namespace cronwalk
