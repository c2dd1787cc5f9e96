// cron_walk.cuh -- per-spec schedule walk shared by every kernel in this library.
//
// What it computes: (*SpecSchedule).Next(t) of robfig/cron v3.0.1 (spec.go; called from the reference at
// internal/controller/cron_controller.go:405,409,436) with Go `time` semantics (Date's two-lookup zone resolution,
// absolute Add, AddDate through Date, the day loop's "hour != 0" DST fix, the 5-year limit).
//
// How: the reference walks month/day/hour/minute/second loops one unit at a time, re-deriving civil fields from
// absolute time at every step (58 steps for "* * * * *", 206 for an unschedulable spec).  Here the walk keeps its
// civil fields resident in registers (local day number, second of day, Y/M/D, current zone period) and every loop
// *jumps* with a bit scan over the mask when the whole jump provably lies inside one constant-offset period of the
// zone -- in which case the unit steps it replaces are uniform and the end state is identical.  When a jump would
// touch a zone transition (or a Date() target that is not "natural"), the walk falls back to one literal upstream
// step and tries again, so DST gaps/overlaps, midnight-gap zones and 30-minute shifts reproduce upstream's
// behaviour step for step.  UTC is the degenerate zone with one infinite period: every jump is taken.
//
// One lane walks one spec ("warp-cooperative" in this library means the loads and the output staging, see
// next_kernels.cu); nothing here is a contraction, so no tensor cores.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CW_HD __host__ __device__ __forceinline__
#else
#define CW_HD inline
#endif
#if defined(__CUDA_ARCH__)
#define CW_LDG(p) __ldg(p)  // zone table reads go through the read-only path
#else
#define CW_LDG(p) (*(p))
#endif

#ifndef CW_TRACE
#define CW_TRACE(tag, w)
#endif
#ifndef CW_COUNT
#define CW_COUNT(what)  // host-side instrumentation hook (profiles/); compiles to nothing in the library
#endif

namespace cronwalk {

constexpr int64_t kZero = INT64_MIN;  // Go's time.Time{} at every boundary of this library
constexpr int64_t kAlpha = INT64_MIN;
constexpr int64_t kOmega = INT64_MAX;
constexpr int32_t kStuck = 400;  // iterations of one loop entry no terminating upstream walk can reach
constexpr int64_t kYearBase = -2208988800LL;  // 1900-01-01T00:00:00Z: origin of the zone table's year index
constexpr uint32_t kYearStep128 = 246375u;    // bucket width of 365 days, in units of 128 seconds

// Flattened zone table in HBM (L2-resident: ~0.5 MB for all IANA zones).  Zone z owns entries
// [zstart[z], zstart[z+1]); entry i is in force for when[i] <= sec < when[i+1] (when of the first entry is kAlpha,
// the last period is open-ended).  Footer rules are pre-expanded to explicit entries up to `horizon`.
struct ZoneTable {
  const int64_t* when;
  const int32_t* off;
  const int32_t* fpre;  // per zone: running count of transitions whose gap/overlap contains a local midnight
  const int32_t* zstart;
  const uint16_t* yidx;  // year index: row z = period in force at the start of every 365-day bucket since 1900 (zone_table.h)
  int32_t ystride;       // entries per row (buckets + 1)
  int32_t nzones;
  int64_t horizon;  // zoned inputs must be < horizon - 6 years (checked by the caller of the kernels)
};

// A ConstantDelaySchedule (robfig constantdelay.go; what ParseStandard returns for "@every <dur>", reached by the
// reference at cron_controller.go:392,405,409,436) travels in the same six planes: month == kEveryTag (a bit no
// parser sets) marks it, dom holds Delay in whole seconds, the other planes are ignored.
constexpr uint64_t kEveryTag = 1ULL << 62;

struct Spec {  // sanitised masks: only the bits upstream can test are kept
  uint64_t second, minute;  // every: `second` holds the delay in seconds (>= 1) instead of a mask
  uint32_t hour, dom, month, dow;
  uint32_t month_eff;  // month mask minus the months that can never hold a matching day (closed form only)
  bool star;   // Dom&starBit || Dow&starBit  -> AND rule, else OR rule (spec.go dayMatches)
  bool never;  // some field can never match: upstream walks to the 5-year limit and returns the zero time
  bool every;  // ConstantDelaySchedule: Next(t) = t + delay; `never` is set too, so no walk ever starts on it
  int32_t tod1;     // first matching second of day (first hour, first minute, first second); valid when !never
  int32_t first_ms; // first matching minute * 60 + first matching second
  int32_t first_s;  // first matching second
};

CW_HD int ctz32(uint32_t x);
CW_HD int ctz64(uint64_t x);

CW_HD Spec make_spec(uint64_t second, uint64_t minute, uint64_t hour, uint64_t dom, uint64_t month, uint64_t dow) {
  Spec s;
  s.second = second & ((1ULL << 60) - 1);
  s.minute = minute & ((1ULL << 60) - 1);
  s.hour = (uint32_t)(hour & 0xFFFFFFu);
  s.dom = (uint32_t)(dom & 0xFFFFFFFEu);
  s.month = (uint32_t)(month & 0x1FFEu);
  s.dow = (uint32_t)(dow & 0x7Fu);
  s.star = ((dom | dow) >> 63) != 0;
  const bool no_day = s.star ? (s.dom == 0 || s.dow == 0) : (s.dom == 0 && s.dow == 0);
  // Months in which no day can match: when the day-of-month mask decides alone (AND rule, or OR rule with an empty
  // day-of-week mask), a 30-day month needs a dom bit <= 30 and February one <= 29.  With a non-empty day-of-week
  // mask under the OR rule every month matches (each weekday occurs in any 28 days).  The step emulation keeps the
  // original mask: it walks such months day by day, as upstream does.
  s.month_eff = s.month;
  if (s.star || s.dow == 0) {
    if ((s.dom & 0x7FFFFFFEu) == 0) s.month_eff &= ~((1u << 4) | (1u << 6) | (1u << 9) | (1u << 11));
    if ((s.dom & 0x3FFFFFFEu) == 0) s.month_eff &= ~(1u << 2);
  }
  // first set positions with a sentinel bit above the valid range: an empty mask shows up as the sentinel
  const int32_t fs = ctz64(s.second | (1ULL << 60)), fm = ctz64(s.minute | (1ULL << 60)), fh = ctz32(s.hour | (1u << 24));
  s.never = (fs == 60) | (fm == 60) | (fh == 24) | (s.month == 0) | no_day;
  s.first_s = fs;
  s.first_ms = fm * 60 + fs;
  s.tod1 = fh * 3600 + s.first_ms;
  s.every = month == kEveryTag;
  if (s.every) {  // constantdelay.go Every() never builds a delay below one second; clamp a hand-built plane likewise
    s.never = true;
    s.second = (int64_t)dom < 1 ? 1ULL : dom;
  }
  return s;
}

// ---- small integer helpers --------------------------------------------------------------------------------
CW_HD int ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __ffs((int)x) - 1;
#else
  return __builtin_ctz(x);
#endif
}
CW_HD int ctz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  const uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
  return lo ? __ffs((int)lo) - 1 : 32 + __ffs((int)hi) - 1;
#else
  return __builtin_ctzll(x);
#endif
}
CW_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
CW_HD int popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __popcll(x);
#else
  return __builtin_popcountll(x);
#endif
}
CW_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}
CW_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return x ? __builtin_clzll(x) : 64;
#endif
}

CW_HD int32_t days_from_civil(int32_t y, int32_t m, int32_t d) {  // proleptic Gregorian, 1970-01-01 = 0
  y -= m <= 2;
  const int32_t era = (y >= 0 ? y : y - 399) / 400;
  const uint32_t yoe = (uint32_t)(y - era * 400);
  const uint32_t doy = (153u * (uint32_t)(m > 2 ? m - 3 : m + 9) + 2u) / 5u + (uint32_t)d - 1u;
  const uint32_t doe = yoe * 365u + yoe / 4u - yoe / 100u + doy;
  return era * 146097 + (int32_t)doe - 719468;
}
CW_HD void civil_from_days(int32_t z, int32_t& y, int32_t& m, int32_t& d) {
  // Neri & Schneider, "Euclidean affine functions and their application to calendar algorithms" (2022), 32-bit
  // variant on days since 0000-03-01; exact for every day number of the supported range (years 1..9999, checked
  // exhaustively against the textbook algorithm in tests/test_walk_host.py through the host build).
  const uint32_t n = (uint32_t)(z + 719468);
  const uint32_t n1 = 4u * n + 3u;
  const uint32_t c = n1 / 146097u;
  const uint32_t nc = (n1 - c * 146097u) / 4u;
  const uint32_t n2 = 4u * nc + 3u;
  const uint64_t p2 = 2939745ULL * n2;
  const uint32_t zc = (uint32_t)(p2 >> 32);
  const uint32_t ny = (uint32_t)p2 / 2939745u / 4u;
  const uint32_t n3 = 2141u * ny + 197913u;
  const uint32_t mm = n3 >> 16;
  const uint32_t dd = (n3 & 0xffffu) / 2141u;
  const bool j = ny >= 306u;
  y = (int32_t)(100u * c + zc) + (j ? 1 : 0);
  m = (int32_t)(j ? mm - 12u : mm);
  d = (int32_t)dd + 1;
}
CW_HD bool is_leap(int32_t y) {  // y % 100 == 0 <=> y % 25 == 0 given y % 4 == 0; y % 400 == 0 <=> additionally y % 16 == 0
  return ((y & 3) == 0) && ((y % 25 != 0) || ((y & 15) == 0));
}
// days before month m (1..13; 13 = next January) in a non-leap year
CW_HD int32_t days_before_month(int32_t m) {
  // 30*(m-1) plus a 3-bit correction per month packed into one constant (no division)
  return 30 * m - 31 + (int32_t)((0x35B23489088ULL >> (3 * m)) & 7ULL);
}
CW_HD int32_t days_in_month(int32_t y, int32_t m) {
  // 31 for 1,3,5,7,8,10,12; 30 for 4,6,9,11; February by leap rule
  return (m == 2) ? (is_leap(y) ? 29 : 28) : 30 + (int32_t)((0x15AAu >> m) & 1u);
}
CW_HD int32_t weekday_of(int32_t dn) {  // Sunday = 0; 1970-01-01 was a Thursday
  // 1073741823 = 7 * 153391689: keeps the dividend non-negative for every supported day number (unsigned mod 7)
  return (int32_t)(((uint32_t)(dn + 4 + 1073741823)) % 7u);
}

// ---- the walk ------------------------------------------------------------------------------------------------
struct Walk {
  // zone
  const int64_t* zw;   // this zone's `when` entries
  const int32_t* zo;   // this zone's offsets
  const int32_t* zf;   // this zone's midnight-transition prefix counts
  const uint16_t* zy;  // this zone's row of the year index
  int32_t zn;          // entry count (1 for UTC / fixed zones)
  int32_t ynb;         // buckets in the row
  // cached period holding the last looked-up instant
  int32_t pidx;
  int32_t off;
  int64_t ps, pe;
  int64_t nlo, nhi;  // natural(x) <=> nlo <= x < nhi; recomputed whenever the cached period changes
  // current instant and its civil fields in the zone
  int64_t t;
  int32_t dn;   // local day number
  int32_t sod;  // local second of day
  int32_t Y, M, D;
  bool today_ok;  // the resident day is known to satisfy the month and day masks (set when a fire was found on it)
  // chained calls (NextN): day mask of the resident month, and the day's last matching second (86400 = not set)
  uint32_t dm_month;
  bool dm_ok;
  int32_t tod_last;

  CW_HD void set_natural_range() {
    // x in [ps, pe) and x - off in [ps, pe)  <=>  ps + max(off, 0) <= x < pe + min(off, 0)
    nlo = (ps == kAlpha) ? kAlpha : ps + (off > 0 ? off : 0);
    nhi = (pe == kOmega) ? kOmega : pe + (off < 0 ? off : 0);
  }
  CW_HD void bind_utc() {
    zw = nullptr; zo = nullptr; zf = nullptr; zy = nullptr; ynb = 0; zn = 1; pidx = 0; off = 0; ps = kAlpha; pe = kOmega; nlo = kAlpha; nhi = kOmega; today_ok = false; dm_ok = false; dm_month = 0; tod_last = 86400;
  }
  CW_HD void bind_zone(const ZoneTable& zt, int32_t zone) {
    const int32_t a = CW_LDG(zt.zstart + zone), b = CW_LDG(zt.zstart + zone + 1);
    zw = zt.when + a; zo = zt.off + a; zf = zt.fpre + a; zn = b - a;
    zy = zt.yidx + (size_t)zone * (size_t)zt.ystride; ynb = zt.ystride - 1;
    pidx = -1; off = 0; ps = 1; pe = 0; nlo = 1; nhi = 0;  // empty cache
    today_ok = false; dm_ok = false; dm_month = 0; tod_last = 86400;
    if (zn == 1) { pidx = 0; off = CW_LDG(zo); ps = kAlpha; pe = kOmega; nlo = kAlpha; nhi = kOmega; }
  }

  // (*Location).lookup(x).offset; moves the cached period to the one containing x
  CW_HD int32_t lookup(int64_t x) {
    if (x >= ps && x < pe) return off;
    int32_t lo = 0, hi = zn;  // largest lo with when[lo] <= x  (when[0] == kAlpha)
    if (pidx >= 0) {          // transitions are usually crossed one at a time: try the neighbours first
      if (x >= pe) { lo = pidx + 1; if (lo + 1 >= zn || x < CW_LDG(zw + lo + 1)) hi = lo + 1; }
      else { hi = pidx; if (hi - 1 >= 0 && x >= CW_LDG(zw + hi - 1)) lo = hi - 1; }
    }
    if (hi - lo > 1 && x >= kYearBase) {
      // year index: the period in force at the start of x's 365-day bucket and of the next bucket bracket the answer
      const uint32_t b = (uint32_t)((uint64_t)(x - kYearBase) >> 7) / kYearStep128;
      if (b < (uint32_t)ynb) {
        const int32_t l2 = (int32_t)CW_LDG(zy + b), h2 = (int32_t)CW_LDG(zy + b + 1) + 1;
        lo = lo > l2 ? lo : l2;
        hi = hi < h2 ? hi : h2;
      }
    }
    while (hi - lo > 1) {
      const int32_t mid = (int32_t)((uint32_t)(lo + hi) >> 1);
      if (x < CW_LDG(zw + mid)) hi = mid; else lo = mid;
    }
    pidx = lo;
    off = CW_LDG(zo + lo);
    ps = CW_LDG(zw + lo);
    pe = (lo + 1 < zn) ? CW_LDG(zw + lo + 1) : kOmega;
    set_natural_range();
    return off;
  }

  // field getters on the resident state
  CW_HD int32_t hour() const { return sod / 3600; }
  CW_HD int32_t minute() const { return (sod % 3600) / 60; }
  CW_HD int32_t second() const { return sod % 60; }

  // set the current instant and re-derive every field (what Go does on each t.Month()/t.Day()/... call)
  CW_HD void load(int64_t tt) {
    t = tt;
    const int64_t l = tt + lookup(tt);
    // 86400 = 128 * 675 and |l| < 2^38 inside the supported range (years 1..9999): floor-divide in 32 bits
    const int32_t hi = (int32_t)(l >> 7);
    int32_t q = hi / 675;
    int32_t r = hi - q * 675;
    if (r < 0) { r += 675; q -= 1; }
    dn = q;
    sod = r * 128 + (int32_t)(l & 127);
    civil_from_days(dn, Y, M, D);
    today_ok = false;
    dm_ok = false;
  }

  // load() for a walk bound to UTC: no period lookup
  CW_HD void load_utc(int64_t tt) {
    t = tt;
    const int32_t hi = (int32_t)(tt >> 7);
    int32_t q = hi / 675;
    int32_t r = hi - q * 675;
    if (r < 0) { r += 675; q -= 1; }
    dn = q;
    sod = r * 128 + (int32_t)(tt & 127);
    civil_from_days(dn, Y, M, D);
    today_ok = false;
    dm_ok = false;
  }

  // time.Date(y, mo, d, h, mi, s, 0, loc): as-if-UTC seconds, then the two-lookup offset resolution
  CW_HD int64_t go_date(int32_t y, int32_t mo, int32_t d, int32_t sod_) {
    int32_t m0 = mo - 1;
    if (m0 >= 12) { y += m0 / 12; m0 %= 12; }
    const int64_t x = (int64_t)(days_from_civil(y, m0 + 1, 1) + (d - 1)) * 86400 + sod_;
    int32_t o = lookup(x);
    if (o != 0) o = lookup(x - o);  // Go: keep o when utc stays in [start,end) -- same value either way
    return x - o;
  }
  // Is local wall time x (as-if-UTC seconds) resolved by Date() to x - off with the current period untouched?
  CW_HD bool natural(int64_t x) const { return x >= nlo && x < nhi; }

  // one literal Add(delta) step, 0 < delta <= 3600: inside the period the fields advance with it; across a period
  // edge they are re-derived from the new offset
  CW_HD void step(int32_t delta) {
    const int64_t tn = t + delta;
    if (tn < pe) {
      t = tn;
      sod += delta;
      if (sod >= 86400) { sod -= 86400; advance_days(1); }
    } else {
      load(tn);
    }
  }
  // May an Add-based loop whose field repeats every `modulus` seconds (60 for the second loop, 3600 for the minute
  // loop) jump from t to tk >= pe?  Yes when tk lies in the very next period and both offsets are congruent modulo
  // `modulus`: the field then reads the same at every instant of the jump whichever period it falls in.
  CW_HD bool congruent_across(int64_t tk, int32_t modulus) const {
    if (pidx < 0 || pidx + 1 >= zn) return false;
    if (pidx + 2 < zn && tk >= CW_LDG(zw + pidx + 2)) return false;
    const int32_t o2 = CW_LDG(zo + pidx + 1);
    return (off % modulus == 0) && (o2 % modulus == 0);
  }

  // move the resident fields forward by k whole days inside the same period (k >= 1, small)
  CW_HD void advance_days(int32_t k) {
    today_ok = false;
    dm_ok = false;
    dn += k;
    D += k;
    int32_t dim = days_in_month(Y, M);
    while (D > dim) {
      D -= dim;
      if (++M == 13) { M = 1; ++Y; }
      dim = days_in_month(Y, M);
    }
  }
};

// 31-bit mask of the days of month (Y, M) on which dayMatches() holds; bit d = day d
CW_HD uint32_t month_daymask(const Spec& s, int32_t Y, int32_t M, int32_t dn_of_first) {
  const int32_t dim = days_in_month(Y, M);
  const uint32_t valid = (uint32_t)(((1ULL << (dim + 1)) - 1ULL) & ~1ULL);
  const int32_t wd1 = weekday_of(dn_of_first);
  // replicate the 7-bit dow mask so that bit j = dow bit (j mod 7), j < 42; day d has weekday (wd1 + d - 1) mod 7
  const uint64_t rep = (uint64_t)s.dow * 0x0000000810204081ULL;  // 1 + 2^7 + 2^14 + 2^21 + 2^28 + 2^35
  const uint32_t dowm = (uint32_t)((rep >> wd1) << 1);
  const uint32_t m = s.star ? (s.dom & dowm) : (s.dom | dowm);
  return m & valid;
}

CW_HD bool day_matches(const Spec& s, int32_t D, int32_t dn) {
  const bool dom = (s.dom >> D) & 1u;
  const bool dow = (s.dow >> weekday_of(dn)) & 1u;
  return s.star ? (dom && dow) : (dom || dow);
}

// Continue an upstream Next() walk from the resident state `w` (which already holds t = input + 1s).
// Returns the fire time (w then holds its fields) or kZero.
CW_HD int64_t walk_next(const Spec& s, Walk& w) {
  if (s.never) return kZero;
  bool added = false;
  const int32_t year_limit = w.Y + 5;

  for (;;) {  // WRAP:
    if (w.Y > year_limit) return kZero;

    // ---------------- month loop ----------------
    bool wrapped = false;
    int32_t guard = 0;
    while (((s.month >> w.M) & 1u) == 0) {
      CW_TRACE("month", w);
      CW_COUNT(month_iter);
      if (++guard > kStuck) return kZero;
      if (!added) {
        added = true;
        w.load(w.go_date(w.Y, w.M, 1, 0));
      }
      // jump: D == 1 means AddDate(0,1,0) lands on the 1st of each following month with h:m:s unchanged
      const uint32_t later = s.month & ~((2u << w.M) - 1u);
      const int32_t tgtM = later ? ctz32(later) : 13;  // 13 = January of next year (-> WRAP)
      bool jumped = false;
      if (w.D == 1) {
        const int32_t dn_first = days_from_civil(w.Y, w.M + 1, 1);  // first step
        const int32_t dn_last = (tgtM == 13) ? days_from_civil(w.Y + 1, 1, 1) : days_from_civil(w.Y, tgtM, 1);
        if (w.natural((int64_t)dn_first * 86400 + w.sod) && w.natural((int64_t)dn_last * 86400 + w.sod)) {
          w.t = (int64_t)dn_last * 86400 + w.sod - w.off;
          w.dn = dn_last;
          if (tgtM == 13) { w.M = 1; w.Y += 1; } else { w.M = tgtM; }
          jumped = true;
        }
      }
      if (!jumped) {
        w.load(w.go_date(w.Y, w.M + 1, w.D, w.sod));  // literal AddDate(0,1,0)
      }
      if (w.M == 1) { wrapped = true; break; }
    }
    if (wrapped) continue;

    // ---------------- day loop ----------------
    // Upstream bug guard: in zones that skipped a whole calendar day (Kiritimati 1994-12-31, Apia 2011-12-30, ...)
    // AddDate(0,0,1) from the day before resolves back to the same instant and upstream's day loop never returns.
    // A terminating day loop wraps at Day()==1 within 31 (+ a few DST-fix) steps; beyond kStuck we report the zero
    // time ("unschedulable"), as the oracle does.
    guard = 0;
    while (!day_matches(s, w.D, w.dn)) {
      CW_TRACE("day", w);
      CW_COUNT(day_iter);
      if (++guard > kStuck) return kZero;
      if (!added) {
        added = true;
        w.load(w.go_date(w.Y, w.M, w.D, 0));
      }
      bool jumped = false;
      if (w.sod < 3600) {  // hour == 0: the "hour != 0" fix cannot trigger on natural steps
        const uint32_t dm = month_daymask(s, w.Y, w.M, w.dn - (w.D - 1));
        const uint32_t later = dm & ~((2u << w.D) - 1u);
        const int32_t dim = days_in_month(w.Y, w.M);
        const int32_t k = (later ? ctz32(later) : dim + 1) - w.D;  // dim+1 = the 1st of next month (-> WRAP)
        const int64_t x1 = (int64_t)(w.dn + 1) * 86400 + w.sod;
        const int64_t xk = (int64_t)(w.dn + k) * 86400 + w.sod;
        if (w.natural(x1) && w.natural(xk)) {
          w.t = xk - w.off;
          w.advance_days(k);
          jumped = true;
        }
      }
      if (!jumped) {
        w.load(w.go_date(w.Y, w.M, w.D + 1, w.sod));  // literal AddDate(0,0,1)
        const int32_t h = w.hour();
        if (h != 0) {  // "Notice if the hour is no longer midnight due to DST."
          if (h > 12) w.load(w.t + (int64_t)(24 - h) * 3600);
          else w.load(w.t - (int64_t)h * 3600);
        }
      }
      if (w.D == 1) { wrapped = true; break; }
    }
    if (wrapped) continue;

    // ---------------- hour loop ----------------
    guard = 0;
    while (((s.hour >> w.hour()) & 1u) == 0) {
      CW_TRACE("hour", w);
      CW_COUNT(hour_iter);
      if (++guard > kStuck) return kZero;
      if (!added) {
        added = true;
        w.load(w.go_date(w.Y, w.M, w.D, w.hour() * 3600));
      }
      const int32_t h = w.hour();  // after the reset: a Date() inside a gap can move the hour
      const uint32_t later = s.hour & ~((2u << h) - 1u);
      const int32_t k = (later ? ctz32(later) : 24) - h;  // 24 = hour 0 of the next day (-> WRAP)
      const int64_t tk = w.t + (int64_t)k * 3600;
      if (tk < w.pe) {  // k absolute one-hour steps inside one period: the local hour advances by one each
        w.t = tk;
        w.sod += k * 3600;
        if (w.sod >= 86400) { w.sod -= 86400; w.advance_days(1); }
      } else {
        w.step(3600);  // literal step towards / across the transition
      }
      if (w.hour() == 0) { wrapped = true; break; }
    }
    if (wrapped) continue;

    // ---------------- minute loop ----------------
    guard = 0;
    while (((s.minute >> w.minute()) & 1ULL) == 0) {
      CW_TRACE("minute", w);
      CW_COUNT(minute_iter);
      if (++guard > kStuck) return kZero;
      if (!added) {
        added = true;
        // Truncate(time.Minute) works on absolute time; the zero Time is a multiple of 60 s away from 1970
        int64_t r = w.t % 60;
        if (r < 0) r += 60;
        if (r != 0) w.load(w.t - r);
      }
      const int32_t mi = w.minute();
      const uint64_t later = s.minute & ~((2ULL << mi) - 1ULL);
      const int32_t k = (later ? ctz64(later) : 60) - mi;  // 60 = minute 0 of the next hour (-> WRAP)
      const int64_t tk = w.t + (int64_t)k * 60;
      if (tk < w.pe) {
        w.t = tk;
        w.sod += k * 60;
        if (w.sod >= 86400) { w.sod -= 86400; w.advance_days(1); }
      } else if (w.congruent_across(tk, 3600)) {
        w.load(tk);  // the minute field is blind to a whole-hour offset change: same stopping instant
      } else {
        w.step(60);
      }
      if (w.minute() == 0) { wrapped = true; break; }
    }
    if (wrapped) continue;

    // ---------------- second loop ----------------
    guard = 0;
    while (((s.second >> w.second()) & 1ULL) == 0) {
      CW_TRACE("second", w);
      CW_COUNT(second_iter);
      if (++guard > kStuck) return kZero;
      added = true;  // Truncate(time.Second) is the identity on whole seconds
      const int32_t sc = w.second();
      const uint64_t later = s.second & ~((2ULL << sc) - 1ULL);
      const int32_t k = (later ? ctz64(later) : 60) - sc;  // 60 = second 0 of the next minute (-> WRAP)
      const int64_t tk = w.t + k;
      if (tk < w.pe) {
        w.t = tk;
        w.sod += k;
        if (w.sod >= 86400) { w.sod -= 86400; w.advance_days(1); }
      } else if (w.congruent_across(tk, 60)) {
        w.load(tk);  // the second field is blind to a whole-minute offset change: same stopping instant
      } else {
        w.step(1);
      }
      if (w.second() == 0) { wrapped = true; break; }
    }
    if (wrapped) continue;

    return w.t;
  }
}


// last matching second of a day (for chained calls: past it, the day has no fire left)
CW_HD int32_t spec_tod_last(const Spec& s) {
  return (31 - clz32(s.hour)) * 3600 + (63 - clz64(s.minute)) * 60 + (63 - clz64(s.second));
}

// first matching time of day >= sod on the same day, or -1.  Branch-free: the three candidates (same minute, later
// minute of the same hour, later hour) are all computed and the earliest valid one is selected.
CW_HD int32_t next_tod_ge(const Spec& s, int32_t sod) {
  const int32_t h = sod / 3600, rem = sod - h * 3600, mi = rem / 60, sc = rem - mi * 60;
  const bool hourbit = (s.hour >> h) & 1u;
  const bool minbit = (s.minute >> mi) & 1ULL;
  const uint64_t sec_m = s.second >> sc;
  const uint64_t min_m = (s.minute >> mi) >> 1;
  const uint32_t hour_m = (s.hour >> h) >> 1;
  const int32_t c1 = sod + ctz64(sec_m | (1ULL << 62));
  const int32_t c2 = h * 3600 + (mi + 1 + ctz64(min_m | (1ULL << 62))) * 60 + s.first_s;
  const int32_t c3 = (h + 1 + ctz32(hour_m | (1u << 30))) * 3600 + s.first_ms;
  int32_t tod = hour_m ? c3 : -1;
  tod = (hourbit && min_m) ? c2 : tod;
  tod = (hourbit && minbit && sec_m) ? c1 : tod;
  return tod;
}


// First fire on the local day `w` is on, at or after local second-of-day `sod` (absolute instant `t`, inside the
// cached period), allowing the search to cross ONE zone transition.  On success returns true with *r / *rsod set and
// the period cache on the result.
//
// Without a transition in the way this is next_tod_ge().  Across a transition at instant T the Add-based loops keep
// stepping absolute time while the local fields jump by d = new offset - old offset.  When T falls on a whole local
// hour, d is a whole number of hours and the local day stays the same, every loop upstream can be in when it reaches
// T is aligned (minutes and seconds read the same on both sides, the minute loop wraps into a full re-check exactly at
// T, the hour loop simply continues with the shifted hour), so the walk equals: nothing matched in [sod, local(T)),
// then the first matching time of day >= local(T) + d in the new period.  Anything else (sub-hour shifts, midnight
// transitions, no fire left that day) is left to the step emulation.
template <bool kUtcOnly = false>
CW_HD bool find_tod_today(const Spec& s, Walk& w, int64_t t, int32_t sod, int64_t* r, int32_t* rsod) {
  const int32_t tod = next_tod_ge(s, sod);
  if (tod >= 0 && (kUtcOnly || t + (tod - sod) < w.pe)) {
    *r = t + (tod - sod);
    *rsod = tod;
    return true;
  }
  if (kUtcOnly) return false;  // one endless period: no fire left today
  if (w.pe - t >= 86400 - sod) return false;  // the period outlives the day: no fire left today (tod < 0)
  if (w.pidx < 0 || w.pidx + 1 >= w.zn) return false;
  const int32_t sod_t = sod + (int32_t)(w.pe - t);  // local time of day at which the transition strikes (< 86400)
  const int32_t o2 = CW_LDG(w.zo + w.pidx + 1);
  const int32_t d = o2 - w.off;
  const int32_t sod2 = sod_t + d;  // local time of day right after it
  if (sod_t % 3600 != 0 || d % 3600 != 0 || o2 % 60 != 0 || sod2 < 0 || sod2 >= 86400) return false;
  const int32_t tod2 = next_tod_ge(s, sod2);
  if (tod2 < 0) return false;
  const int64_t rr = w.pe + (tod2 - sod2);
  if (w.pidx + 2 < w.zn && rr >= CW_LDG(w.zw + w.pidx + 2)) return false;  // a second transition is in the way
  w.lookup(rr);
  *r = rr;
  *rsod = tod2;
  return true;
}

// Closed-form Next() from the resident state (t = input + 1s).  Returns true and leaves `w` on the result (or sets
// *out = kZero) when the answer is provably what upstream's walk produces; returns false -- with w's fields (not
// its period cache) untouched -- when only the step emulation (walk_next) can tell.
//
// Upstream's walk equals "the smallest instant >= t whose local fields match" whenever (1) every Date() call it
// makes resolves naturally and (2) its Add-based hour/minute/second loops never step across a zone transition.
// Date() targets are: the hour start of t (hour reset), the midnight of t's day or the 1st of t's month (day /
// month reset), every midnight up to the fire day, and the 1st of skipped months -- all local midnights except the
// first.  A local midnight is natural unless it lies in the gap/overlap of a transition; those transitions are
// counted in the zone table (fpre), so "none between here and the fire day" is two loads.  Add-based loops run on
// t's own day (until the fire, or until the Hour()==0 wrap at the next midnight) and on the fire day from its
// midnight: both stretches must sit inside one period.  Offsets must be whole minutes (Truncate is absolute).
//
// The code is straight-line on purpose (one lane per spec, 32 specs per warp): the only data-dependent control flow
// is the trip count of the month loop, which visits matching months only.
// kUtcOnly: the launch has no zone ids at all (every spec in UTC): the zone logic compiles away.
template <bool kUtcOnly = false>
CW_HD bool fast_next(const Spec& s, Walk& w, int64_t* out) {
  const int32_t sod = w.sod;
  const bool zoned = !kUtcOnly && w.zn > 1;
  // Zoned prerequisites, evaluated without an early exit inside the zoned branch: with a `return` in there the
  // compiler kept UTC and zoned lanes apart for the rest of the function (every mixed warp ran it twice).
  // natural() is trivially true for UTC (one infinite period, offset 0), so all lanes run the same code.
  const bool prereq = kUtcOnly || ((w.off % 60 == 0) && w.natural((int64_t)w.dn * 86400 + (sod / 3600) * 3600));  // hour reset Date(Y,M,D,h,0,0)
  if (!prereq) return false;
  // Chained-call shortcut: the previous fire was the last of its (matching) day, the resident month's day mask is
  // cached, and the next matching day lies in the same month.  Upstream wraps at the next local midnight
  // (Add-based, must happen inside this period), crosses the non-matching days with AddDate steps at hour 0 (all
  // those midnights natural in this period: the range is convex, so the first and the last suffice) and runs the
  // Add-based loops from the fire day's midnight to its first fire (inside this period).  Anything else falls
  // through to the general code below.
  if (w.today_ok && w.dm_ok && sod > w.tod_last) {
    const uint32_t rest = w.dm_month & ~((2u << w.D) - 1u);
    if (rest) {
      const int32_t Dn = ctz32(rest);
      const int32_t k = Dn - w.D;
      const int64_t x1 = (int64_t)(w.dn + 1) * 86400, xk = (int64_t)(w.dn + k) * 86400;
      const int64_t r = xk - w.off + s.tod1;
      if (kUtcOnly || (w.t + (86400 - sod) < w.pe && w.natural(x1) && w.natural(xk) && r < w.pe)) {
        w.t = r;
        w.dn += k; w.D = Dn; w.sod = s.tod1;
        *out = r;
        return true;
      }
    }
  }
  const bool month_ok = (s.month >> w.M) & 1u;
  const uint64_t rep = (uint64_t)s.dow * 0x0000000810204081ULL;  // dow pattern repeated every 7 bits
  bool leap = false;
  int32_t first_dn = w.dn - (w.D - 1);
  uint32_t dm = 0;
  bool have_dm = false;
  bool today = w.today_ok;  // chained call on a day already known to match: only the time of day is open
  if (!today) {
    // day mask of the current month, evaluated by every lane of a fresh call before any branch
    leap = is_leap(w.Y);
    const int32_t dim = (w.M == 2) ? (leap ? 29 : 28) : 30 + (int32_t)((0x15AAu >> w.M) & 1u);
    const uint32_t valid = (uint32_t)((2ULL << dim) - 2ULL);
    const uint32_t dowm = (uint32_t)((rep >> weekday_of(first_dn)) << 1);
    dm = month_ok ? ((s.star ? (s.dom & dowm) : (s.dom | dowm)) & valid) : 0u;
    have_dm = true;
    today = (dm >> w.D) & 1u;  // == month matches && dayMatches(t)
  }
  if (today) {
    if (sod <= w.tod_last) {  // (chained calls know the day's last fire: past it there is nothing to search)
      int64_t r;
      int32_t rsod;
      if (find_tod_today<kUtcOnly>(s, w, w.t, sod, &r, &rsod)) {
        w.t = r;
        w.sod = rsod;
        w.today_ok = true;
        if (have_dm) { w.dm_month = dm; w.dm_ok = true; }
        *out = r;
        return true;
      }
    }
    if (!kUtcOnly && w.t + (86400 - sod) >= w.pe) return false;  // the Hour()==0 wrap must happen inside this period
  }
  if (!have_dm) {
    leap = is_leap(w.Y);
    if (w.dm_ok) {
      dm = w.dm_month;  // cached by the call that brought the walk into this month
    } else {
      const int32_t dim = (w.M == 2) ? (leap ? 29 : 28) : 30 + (int32_t)((0x15AAu >> w.M) & 1u);
      const uint32_t valid = (uint32_t)((2ULL << dim) - 2ULL);
      const uint32_t dowm = (uint32_t)((rep >> weekday_of(first_dn)) << 1);
      dm = month_ok ? ((s.star ? (s.dom & dowm) : (s.dom | dowm)) & valid) : 0u;
    }
  }
  const uint32_t dm_cur = dm;  // full mask of the resident month

  // ---- calendar search for the next matching day after today, visiting matching months only ----
  const int32_t year_limit = w.Y + 5;
  int32_t Yc = w.Y, Mc = w.M;
  int32_t jan1 = first_dn - days_before_month(Mc) - ((leap && Mc > 2) ? 1 : 0);
  dm &= ~((2u << w.D) - 1u);  // current month: days after today
  CW_COUNT(cal_enter);
  while (dm == 0) {
    CW_COUNT(cal_iter);
    const uint32_t later = s.month_eff & ~((2u << Mc) - 1u);
    if (later) {
      Mc = ctz32(later);
    } else {
      jan1 += 365 + (leap ? 1 : 0);
      ++Yc;
      leap = is_leap(Yc);
      Mc = ctz32(s.month_eff | (1u << 13));  // 13: no month can ever match -> only the year counter advances
      if (Yc > year_limit) break;
    }
    if (Mc > 12) continue;
    first_dn = jan1 + days_before_month(Mc) + ((leap && Mc > 2) ? 1 : 0);
    const int32_t dim = (Mc == 2) ? (leap ? 29 : 28) : 30 + (int32_t)((0x15AAu >> Mc) & 1u);
    const uint32_t valid = (uint32_t)((2ULL << dim) - 2ULL);
    const uint32_t dowm = (uint32_t)((rep >> weekday_of(first_dn)) << 1);
    dm = (s.star ? (s.dom & dowm) : (s.dom | dowm)) & valid;
  }
  const bool found = dm != 0;
  // day mask of the month the result lands in (full mask when the loop advanced, else the resident month's)
  const uint32_t dm_found = (Yc == w.Y && Mc == w.M) ? dm_cur : dm;
  const int32_t Dc = found ? ctz32(dm) : 1;
  const int32_t found_dn = found ? first_dn + Dc - 1 : jan1;  // not found: Jan 1 of year_limit + 1 (the WRAP that fails)
  const int64_t x_mid = (int64_t)found_dn * 86400;            // local midnight of the fire day
  if (!zoned) {
    if (!found) { *out = kZero; return true; }
    w.t = x_mid + s.tod1 - (kUtcOnly ? 0 : w.off);
    w.dn = found_dn; w.sod = s.tod1; w.Y = Yc; w.M = Mc; w.D = Dc;
    w.today_ok = true;
    w.dm_month = dm_found; w.dm_ok = true;
    *out = w.t;
    return true;
  }
  // ---- zoned: are all the midnights natural, and is the fire day's stretch inside one period? ----
  const int32_t total_flagged = CW_LDG(w.zf + (w.zn - 1));
  int32_t i_lo = 0;
  if (total_flagged != 0) {
    // earliest Date() target: the 1st of t's month when the month is reset, else today's midnight
    const int64_t x_lo = (int64_t)(month_ok ? w.dn : (w.dn - (w.D - 1))) * 86400;
    w.lookup(x_lo - w.off - 2 * 86400);
    i_lo = w.pidx;
  }
  const int32_t o = w.lookup(x_mid);
  if (o != 0) {
    const int32_t o2 = w.lookup(x_mid - o);
    if (o2 != o) return false;
  }
  const int64_t u_mid = x_mid - o;  // Date(found day, 00:00:00); the cache now holds its period
  if (u_mid < w.ps || u_mid >= w.pe) return false;
  if (total_flagged != 0) {
    // transitions strictly after i_lo up to the one that ends the fire day's period
    const int32_t i_hi = (w.pidx + 1 < w.zn) ? w.pidx + 1 : w.pidx;
    if (CW_LDG(w.zf + i_hi) != CW_LDG(w.zf + i_lo)) return false;
  }
  if (!found) { *out = kZero; return true; }
  if (o % 60 != 0) return false;
  // the fire day: Add-based loops run from its midnight (hour 0, minute 0, second 0) to the first fire
  int64_t r = u_mid + s.tod1;
  int32_t rsod = s.tod1;
  if (r >= w.pe && !find_tod_today(s, w, u_mid, 0, &r, &rsod)) return false;  // a transition sits before the fire
  w.t = r;
  w.dn = found_dn; w.sod = rsod; w.Y = Yc; w.M = Mc; w.D = Dc;
  w.today_ok = true;
  w.dm_month = dm_found; w.dm_ok = true;
  *out = r;
  return true;
}

// Next(t) in two halves so that kernels can run the straight-line half for several specs in lock step and share one
// copy of the step emulation: next_try() = t.Add(1s), field decomposition and the closed form; it returns false when
// only walk_next() can decide (then call next_slow()).  `resident` = w already holds the fields of `t`.
template <bool kUtcOnly = false>
CW_HD bool next_try(const Spec& s, Walk& w, int64_t t, bool resident, int64_t* out) {
  CW_COUNT(next);
  if (s.never) { *out = s.every ? t + (int64_t)s.second : kZero; return true; }  // constantdelay.go Next: t.Add(Delay)
  const int64_t t1 = t + 1;
  if (resident && (kUtcOnly || t1 < w.pe) && w.sod + 1 < 86400) {
    w.t = t1;
    w.sod += 1;
  } else if (kUtcOnly) {
    w.load_utc(t1);
  } else {
    w.load(t1);
  }
  return fast_next<kUtcOnly>(s, w, out);
}
// The step emulation is rare and large (thousands of instructions).  On the device it lives out of line, once per
// kernel image, and takes/returns its state by value so that callers keep their Walk in registers on the fast path
// (an unmet kernel with it inlined at every call site was 256 KB of SASS and stalled on instruction fetch).
struct SlowResult {
  Walk w;
  int64_t r;
};
#if defined(__CUDA_ARCH__)
__device__ __noinline__
#else
inline
#endif
SlowResult walk_next_outlined(Spec s, Walk w) {
  SlowResult x;
  x.w = w;
  x.w.lookup(x.w.t);  // fast_next may have moved the period cache; the step emulation wants it on t
  x.r = walk_next(s, x.w);
  x.w.today_ok = false;  // the emulation can end on a day upstream never re-checked
  x.w.dm_ok = false;
  return x;
}
// kOutline = true: call the single out-of-line copy (unmet / list kernels: several phases share it and the kernel
// must stay inside the instruction cache); false: inline it (next_n kernels: one call site, and measured faster).
template <bool kOutline>
CW_HD int64_t next_slow(const Spec& s, Walk& w) {
  CW_COUNT(slow);
  if (kOutline) {
    const SlowResult x = walk_next_outlined(s, w);
    w = x.w;
    return x.r;
  }
  w.lookup(w.t);
  const int64_t r = walk_next(s, w);
  w.today_ok = false;
  w.dm_ok = false;
  return r;
}
template <bool kOutline, bool kUtcOnly = false>
CW_HD int64_t next_from_t(const Spec& s, Walk& w, int64_t t, bool resident) {
  int64_t r = kZero;
  // in UTC the closed form always answers (one endless period, every Date() natural): no step emulation in the kernel
  if (next_try<kUtcOnly>(s, w, t, resident, &r) || kUtcOnly) return r;
  return next_slow<kOutline>(s, w);
}
CW_HD int64_t next_from(const Spec& s, Walk& w, int64_t t, bool resident) { return next_from_t<false>(s, w, t, resident); }

// ---- one zone period's local clock --------------------------------------------------------------------------------
// Inside one constant-offset period local time is UTC moved by the offset.  When everything a computation can touch
// lies inside the natural window of ONE period -- from the earliest Date() target upstream can make (the midnight of
// the 1st of the starting month: at most 32 days back) to its last Next() result -- every Date() resolves naturally
// and every Add-based stretch stays in the period, so upstream's walk is the UTC walk on that clock.
//
// period_clock: does the period in force 32 days before `lo` still hold `hi`?  (t_min / t_max: supported absolute
// range; zoned_max: zoned evaluations must start below it.)  *shift = the period's offset, *nhi = end of its natural
// window on the local clock; the caller must still check that its results stay below *nhi.  Invalid ids and anything
// doubtful say no: the general path decides and reports.
CW_HD bool period_clock(const ZoneTable& zt, int32_t zone, int64_t lo, int64_t hi, int64_t t_min, int64_t t_max,
                        int64_t zoned_max, int32_t* shift, int64_t* nhi) {
  if (zone <= 0 || zone >= zt.nzones) return false;
  if (lo < t_min + 33 * 86400LL || hi > t_max) return false;
  Walk w;
  w.bind_zone(zt, zone);
  if (w.zn > 1 && hi >= zoned_max) return false;
  const int32_t off = w.lookup(lo - 32 * 86400LL);
  if (!(hi < w.pe) || off % 60 != 0) return false;
  *shift = off;
  *nhi = w.nhi;
  return true;
}

// period_clock for a chain of Next() calls starting at t, with the exact earliest Date() target instead of the 32-day
// margin: upstream's first call resets to the midnight of (t+1s)'s day when only the day mismatches, or of the 1st of
// its month when the month mismatches (month_mask = the spec's month plane); every later call starts on a fire, whose
// month and day match, so its targets are never earlier.  The period holding t+1s must hold that midnight naturally.
CW_HD bool period_clock_next(const ZoneTable& zt, int32_t zone, int64_t t, uint64_t month_mask, int64_t t_min, int64_t t_max,
                             int64_t zoned_max, int32_t* shift, int64_t* nhi) {
  if (zone <= 0 || zone >= zt.nzones) return false;
  if (t < t_min + 33 * 86400LL || t > t_max) return false;
  Walk w;
  w.bind_zone(zt, zone);
  if (w.zn > 1 && t >= zoned_max) return false;
  w.load(t + 1);
  if (w.off % 60 != 0) return false;
  const int64_t x_lo = (int64_t)(((month_mask >> w.M) & 1ULL) ? w.dn : w.dn - (w.D - 1)) * 86400;
  if (x_lo < w.nlo) return false;
  *shift = w.off;
  *nhi = w.nhi;
  return true;
}

// k chained Next() calls after `from` on a period's clock (shift, nhi from period_clock): true = out[0..k) holds the
// fire times; false = some call left the period (or found nothing within five years, i.e. searched beyond it) and the
// general walk must answer.
CW_HD bool next_n_on_period_clock(const Spec& s, int64_t from, int k, int32_t shift, int64_t nhi, int64_t* out) {
  Walk w;
  w.bind_utc();
  if (!s.never) w.tod_last = spec_tod_last(s);
  int64_t t = from + shift;
  for (int j = 0; j < k; ++j) {
    t = next_from_t<false, true>(s, w, t, j > 0);
    if (t == kZero || t >= nhi) return false;
    out[j] = t - shift;
  }
  return true;
}

}  // namespace cronwalk
