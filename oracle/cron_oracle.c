/*
 * cron_oracle.c -- CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).  See cron_oracle.h.
 *
 * Literal restatement: every loop of robfig/cron v3.0.1 spec.go Next() is executed step by step and every
 * field read re-derives civil fields from absolute seconds through Location.lookup, like Go does.  No bit
 * tricks, no closed forms: this file is the thing the CUDA kernels are checked against, so it favours
 * one-to-one correspondence with the upstream source over speed.  PARITY UNPINNED beyond K1-K4 (header).
 */
#include "cron_oracle.h"

#include <ctype.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define ALPHA INT64_MIN /* Go time/zoneinfo.go: alpha = -1<<63 */
#define OMEGA INT64_MAX /* omega = 1<<63 - 1 */
#define STAR_BIT (1ULL << 63) /* spec.go: starBit */

/* ------------------------------------------------------------------------------------------------
 * Proleptic Gregorian civil arithmetic (what Go's absDate / Date normalisation compute).
 * ---------------------------------------------------------------------------------------------- */
static int64_t fdiv(int64_t a, int64_t b) { /* floor division, b > 0 */
  int64_t q = a / b;
  if ((a % b) < 0) q--;
  return q;
}
static int64_t fmod_(int64_t a, int64_t b) { return a - fdiv(a, b) * b; }

static int64_t days_from_civil(int64_t y, int64_t m /*1..12*/, int64_t d) {
  y -= m <= 2;
  int64_t era = fdiv(y, 400);
  int64_t yoe = y - era * 400;
  int64_t doy = (153 * (m > 2 ? m - 3 : m + 9) + 2) / 5 + d - 1;
  int64_t doe = yoe * 365 + yoe / 4 - yoe / 100 + doy;
  return era * 146097 + doe - 719468;
}
static void civil_from_days(int64_t z, int64_t* y, int64_t* m, int64_t* d) {
  z += 719468;
  int64_t era = fdiv(z, 146097);
  int64_t doe = z - era * 146097;
  int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  int64_t yy = yoe + era * 400;
  int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  int64_t mp = (5 * doy + 2) / 153;
  *d = doy - (153 * mp + 2) / 5 + 1;
  *m = mp < 10 ? mp + 3 : mp - 9;
  *y = yy + (*m <= 2);
}
static int is_leap(int64_t y) { return (y % 4 == 0) && (y % 100 != 0 || y % 400 == 0); }
static int days_in(int64_t m, int64_t y) {
  static const int dm[13] = {0, 31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
  return (m == 2 && is_leap(y)) ? 29 : dm[m];
}
static const int days_before[13] = {0, 31, 59, 90, 120, 151, 181, 212, 243, 273, 304, 334, 365};

/* ------------------------------------------------------------------------------------------------
 * time.Location  (Go: zoneinfo.go, zoneinfo_read.go)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t offset;
  int is_dst;
} orc_zone;
typedef struct {
  int64_t when;
  uint8_t index;
} orc_trans;
struct orc_loc {
  orc_zone* zone;
  int nzone;
  orc_trans* tx;
  int ntx;
  char* extend;
};

/* ---- tzset: evaluation of the TZif footer POSIX-TZ string (zoneinfo.go: tzset and helpers) ---- */
typedef struct {
  int kind; /* 0 Julian (Jn), 1 DOY (n), 2 MonthWeekDay (Mm.w.d) */
  int day, week, mon;
  int time;
} tz_rule;

static int tzset_name(const char** ps) {
  const char* s = *ps;
  size_t len = strlen(s);
  if (len == 0) return 0;
  if (s[0] != '<') {
    for (size_t i = 0; i < len; i++) {
      char r = s[i];
      if ((r >= '0' && r <= '9') || r == ',' || r == '-' || r == '+') {
        if (i < 3) return 0;
        *ps = s + i;
        return 1;
      }
    }
    if (len < 3) return 0;
    *ps = s + len;
    return 1;
  }
  for (size_t i = 0; i < len; i++)
    if (s[i] == '>') {
      *ps = s + i + 1;
      return 1;
    }
  return 0;
}
static int tzset_num(const char** ps, int min, int max, int* out) {
  const char* s = *ps;
  size_t len = strlen(s);
  if (len == 0) return 0;
  int num = 0;
  for (size_t i = 0; i < len; i++) {
    char r = s[i];
    if (r < '0' || r > '9') {
      if (i == 0 || num < min) return 0;
      *out = num;
      *ps = s + i;
      return 1;
    }
    num = num * 10 + (r - '0');
    if (num > max) return 0;
  }
  if (num < min) return 0;
  *out = num;
  *ps = s + len;
  return 1;
}
static int tzset_offset(const char** ps, int* out) {
  const char* s = *ps;
  if (*s == 0) return 0;
  int neg = 0;
  if (*s == '+')
    s++;
  else if (*s == '-') {
    s++;
    neg = 1;
  }
  int hours;
  if (!tzset_num(&s, 0, 24 * 7, &hours)) return 0;
  int off = hours * 3600;
  if (*s != ':') {
    *out = neg ? -off : off;
    *ps = s;
    return 1;
  }
  s++;
  int mins;
  if (!tzset_num(&s, 0, 59, &mins)) return 0;
  off += mins * 60;
  if (*s != ':') {
    *out = neg ? -off : off;
    *ps = s;
    return 1;
  }
  s++;
  int secs;
  if (!tzset_num(&s, 0, 59, &secs)) return 0;
  off += secs;
  *out = neg ? -off : off;
  *ps = s;
  return 1;
}
static int tzset_rule(const char** ps, tz_rule* r) {
  const char* s = *ps;
  memset(r, 0, sizeof *r);
  if (*s == 0) return 0;
  if (*s == 'J') {
    s++;
    int jday;
    if (!tzset_num(&s, 1, 365, &jday)) return 0;
    r->kind = 0;
    r->day = jday;
  } else if (*s == 'M') {
    s++;
    int mon, week, day;
    if (!tzset_num(&s, 1, 12, &mon) || *s != '.') return 0;
    s++;
    if (!tzset_num(&s, 1, 5, &week) || *s != '.') return 0;
    s++;
    if (!tzset_num(&s, 0, 6, &day)) return 0;
    r->kind = 2;
    r->day = day;
    r->week = week;
    r->mon = mon;
  } else {
    int day;
    if (!tzset_num(&s, 0, 365, &day)) return 0;
    r->kind = 1;
    r->day = day;
  }
  if (*s != '/') {
    r->time = 2 * 3600; /* 2am is the default */
    *ps = s;
    return 1;
  }
  s++;
  int off;
  if (!tzset_offset(&s, &off)) return 0;
  r->time = off;
  *ps = s;
  return 1;
}
static int tzrule_time(int64_t year, const tz_rule* r, int off) {
  int s = 0;
  switch (r->kind) {
    case 0:
      s = (r->day - 1) * 86400;
      if (is_leap(year) && r->day >= 60) s += 86400;
      break;
    case 1:
      s = r->day * 86400;
      break;
    case 2: {
      /* Zeller's congruence, as upstream */
      int m1 = (r->mon + 9) % 12 + 1;
      int64_t yy0 = year;
      if (r->mon <= 2) yy0--;
      int64_t yy1 = yy0 / 100;
      int64_t yy2 = yy0 % 100;
      int64_t dow = ((26 * m1 - 2) / 10 + 1 + yy2 + yy2 / 4 + yy1 / 4 - 2 * yy1) % 7;
      if (dow < 0) dow += 7;
      int d = r->day - (int)dow;
      if (d < 0) d += 7;
      for (int i = 1; i < r->week; i++) {
        if (d + 7 >= days_in(r->mon, year)) break;
        d += 7;
      }
      d += days_before[r->mon - 1];
      if (is_leap(year) && r->mon > 2) d++;
      s = d * 86400;
      break;
    }
  }
  return s + r->time - off;
}
/* tzset(s, lastTxSec, sec) -> ok; outputs offset, start, end */
static int tzset_eval(const char* s, int64_t last_tx, int64_t sec, int32_t* offset, int64_t* start, int64_t* end) {
  int std_off, dst_off;
  if (!tzset_name(&s)) return 0;
  if (!tzset_offset(&s, &std_off)) return 0;
  std_off = -std_off;
  if (*s == 0 || *s == ',') {
    *offset = std_off;
    *start = last_tx;
    *end = OMEGA;
    return 1;
  }
  if (!tzset_name(&s)) return 0;
  if (*s == 0 || *s == ',') {
    dst_off = std_off + 3600;
  } else {
    if (!tzset_offset(&s, &dst_off)) return 0;
    dst_off = -dst_off;
  }
  if (*s == 0) s = ",M3.2.0,M11.1.0";
  if (*s != ',' && *s != ';') return 0;
  s++;
  tz_rule sr, er;
  if (!tzset_rule(&s, &sr) || *s != ',') return 0;
  s++;
  if (!tzset_rule(&s, &er) || *s != 0) return 0;

  /* year and seconds-into-year of `sec`, in UTC */
  int64_t days = fdiv(sec, 86400);
  int64_t y, m, d;
  civil_from_days(days, &y, &m, &d);
  int64_t yday0 = days - days_from_civil(y, 1, 1); /* 0-based */
  int64_t ysec = yday0 * 86400 + fmod_(sec, 86400);
  int64_t ystart = sec - ysec;

  int64_t start_sec = tzrule_time(y, &sr, std_off);
  int64_t end_sec = tzrule_time(y, &er, dst_off);
  if (end_sec < start_sec) {
    int64_t t = start_sec;
    start_sec = end_sec;
    end_sec = t;
    int o = std_off;
    std_off = dst_off;
    dst_off = o;
  }
  if (ysec < start_sec) {
    *offset = std_off;
    *start = ystart;
    *end = start_sec + ystart;
  } else if (ysec >= end_sec) {
    *offset = std_off;
    *start = end_sec + ystart;
    *end = ystart + 365 * 86400;
  } else {
    *offset = dst_off;
    *start = start_sec + ystart;
    *end = end_sec + ystart;
  }
  return 1;
}

/* lookupFirstZone (zoneinfo.go) */
static int first_zone_used(const orc_loc* l) {
  for (int i = 0; i < l->ntx; i++)
    if (l->tx[i].index == 0) return 1;
  return 0;
}
static int lookup_first_zone(const orc_loc* l) {
  if (!first_zone_used(l)) return 0;
  if (l->ntx > 0 && l->zone[l->tx[0].index].is_dst) {
    for (int zi = (int)l->tx[0].index - 1; zi >= 0; zi--)
      if (!l->zone[zi].is_dst) return zi;
  }
  for (int zi = 0; zi < l->nzone; zi++)
    if (!l->zone[zi].is_dst) return zi;
  return 0;
}

void orc_loc_lookup(const orc_loc* l, int64_t sec, int32_t* offset, int64_t* start, int64_t* end) {
  if (l == NULL || l->nzone == 0) { /* UTC */
    *offset = 0;
    *start = ALPHA;
    *end = OMEGA;
    return;
  }
  if (l->ntx == 0 || sec < l->tx[0].when) {
    const orc_zone* z = &l->zone[lookup_first_zone(l)];
    *offset = z->offset;
    *start = ALPHA;
    *end = l->ntx > 0 ? l->tx[0].when : OMEGA;
    return;
  }
  /* binary search for the entry with the largest when <= sec */
  const orc_trans* tx = l->tx;
  int64_t e = OMEGA;
  int lo = 0, hi = l->ntx;
  while (hi - lo > 1) {
    int m = (int)((unsigned)(lo + hi) >> 1);
    int64_t lim = tx[m].when;
    if (sec < lim) {
      e = lim;
      hi = m;
    } else {
      lo = m;
    }
  }
  const orc_zone* z = &l->zone[tx[lo].index];
  *offset = z->offset;
  *start = tx[lo].when;
  *end = e;
  if (lo == l->ntx - 1 && l->extend && l->extend[0]) {
    int32_t eo;
    int64_t es, ee;
    if (tzset_eval(l->extend, *start, sec, &eo, &es, &ee)) {
      *offset = eo;
      *start = es;
      *end = ee;
    }
  }
}

static uint32_t be32(const unsigned char* p) {
  return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3];
}
static uint64_t be64(const unsigned char* p) { return ((uint64_t)be32(p) << 32) | be32(p + 4); }

static void set_err(char* err, size_t cap, const char* fmt, const char* a) {
  if (err && cap) snprintf(err, cap, fmt, a);
}

/* LoadLocationFromTZData (zoneinfo_read.go) */
static orc_loc* loc_from_tzdata(const unsigned char* data, size_t len) {
  size_t pos = 0;
#define NEED(nbytes)                 \
  do {                               \
    if (pos + (nbytes) > len) return NULL; \
  } while (0)
  NEED(4);
  if (memcmp(data, "TZif", 4) != 0) return NULL;
  pos = 4;
  NEED(16);
  int version;
  switch (data[pos]) {
    case 0: version = 1; break;
    case '2': version = 2; break;
    case '3': version = 3; break;
    default: return NULL;
  }
  pos += 16;
  enum { NUTCLocal, NStdWall, NLeap, NTime, NZone, NChar };
  int64_t n[6];
  for (int i = 0; i < 6; i++) {
    NEED(4);
    n[i] = be32(data + pos);
    pos += 4;
  }
  int is64 = 0;
  if (version > 1) {
    size_t skip = n[NTime] * 4 + n[NTime] + n[NZone] * 6 + n[NChar] + n[NLeap] * 8 + n[NStdWall] + n[NUTCLocal];
    skip += 4 + 16;
    NEED(skip);
    pos += skip;
    is64 = 1;
    for (int i = 0; i < 6; i++) {
      NEED(4);
      n[i] = be32(data + pos);
      pos += 4;
    }
  }
  int size = is64 ? 8 : 4;
  NEED(n[NTime] * size);
  const unsigned char* txtimes = data + pos;
  pos += n[NTime] * size;
  NEED(n[NTime]);
  const unsigned char* txzones = data + pos;
  pos += n[NTime];
  NEED(n[NZone] * 6);
  const unsigned char* zonedata = data + pos;
  pos += n[NZone] * 6;
  NEED(n[NChar]);
  pos += n[NChar];
  NEED(n[NLeap] * (size + 4));
  pos += n[NLeap] * (size + 4);
  NEED(n[NStdWall]);
  pos += n[NStdWall];
  NEED(n[NUTCLocal]);
  pos += n[NUTCLocal];
#undef NEED
  char* extend = NULL;
  size_t rest = len - pos;
  if (rest > 2 && data[pos] == '\n' && data[len - 1] == '\n') {
    extend = (char*)malloc(rest - 1);
    memcpy(extend, data + pos + 1, rest - 2);
    extend[rest - 2] = 0;
  }
  if (n[NZone] == 0) {
    free(extend);
    return NULL;
  }
  orc_loc* l = (orc_loc*)calloc(1, sizeof *l);
  l->nzone = (int)n[NZone];
  l->zone = (orc_zone*)calloc(l->nzone, sizeof(orc_zone));
  for (int i = 0; i < l->nzone; i++) {
    const unsigned char* z = zonedata + 6 * i;
    l->zone[i].offset = (int32_t)be32(z);
    l->zone[i].is_dst = z[4] != 0;
    if ((int64_t)z[5] >= n[NChar]) {
      free(l->zone); free(l); free(extend);
      return NULL;
    }
  }
  l->ntx = (int)n[NTime];
  l->tx = (orc_trans*)calloc(l->ntx > 0 ? l->ntx : 1, sizeof(orc_trans));
  for (int i = 0; i < l->ntx; i++) {
    l->tx[i].when = is64 ? (int64_t)be64(txtimes + 8 * i) : (int64_t)(int32_t)be32(txtimes + 4 * i);
    if (txzones[i] >= l->nzone) {
      free(l->zone); free(l->tx); free(l); free(extend);
      return NULL;
    }
    l->tx[i].index = txzones[i];
  }
  if (l->ntx == 0) { /* fake transition covering all time (fixed locations such as Etc/GMT+5) */
    l->tx[0].when = ALPHA;
    l->tx[0].index = 0;
    l->ntx = 1;
  }
  l->extend = extend;
  return l;
}

orc_loc* orc_load_location(const char* dir, const char* name, char* err, size_t errcap) {
  /* time.LoadLocation: "" and "UTC" are UTC; reject "..", leading '/' or '\\' */
  if (name[0] == 0 || strcmp(name, "UTC") == 0) {
    orc_loc* l = (orc_loc*)calloc(1, sizeof *l); /* nzone == 0 <=> UTC */
    return l;
  }
  if (strstr(name, "..") || name[0] == '/' || name[0] == '\\') {
    set_err(err, errcap, "time: invalid location name%s", "");
    return NULL;
  }
  char path[1024];
  snprintf(path, sizeof path, "%s/%s", dir, name);
  FILE* f = fopen(path, "rb");
  if (!f) {
    set_err(err, errcap, "unknown time zone %s", name);
    return NULL;
  }
  unsigned char* buf = (unsigned char*)malloc(1 << 20);
  size_t len = fread(buf, 1, 1 << 20, f);
  fclose(f);
  orc_loc* l = loc_from_tzdata(buf, len);
  free(buf);
  if (!l) set_err(err, errcap, "malformed time zone information: %s", name);
  return l;
}
void orc_loc_free(orc_loc* l) {
  if (!l) return;
  free(l->zone);
  free(l->tx);
  free(l->extend);
  free(l);
}

/* ------------------------------------------------------------------------------------------------
 * time.Time operations used by spec.go
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int64_t year, month, day, hour, minute, second, weekday;
} civil;

static civil fields_of(const orc_loc* loc, int64_t t) {
  int32_t off;
  int64_t st, en;
  orc_loc_lookup(loc, t, &off, &st, &en);
  int64_t l = t + off;
  int64_t days = fdiv(l, 86400);
  int64_t sod = l - days * 86400;
  civil c;
  civil_from_days(days, &c.year, &c.month, &c.day);
  c.hour = sod / 3600;
  c.minute = (sod % 3600) / 60;
  c.second = sod % 60;
  c.weekday = fmod_(days + 4, 7); /* 1970-01-01 was a Thursday */
  return c;
}
void orc_fields(const orc_loc* loc, int64_t t, int64_t out[7]) {
  civil c = fields_of(loc, t);
  out[0] = c.year; out[1] = c.month; out[2] = c.day; out[3] = c.hour;
  out[4] = c.minute; out[5] = c.second; out[6] = c.weekday;
}

/* time.Date: normalise, compute as-if-UTC, then resolve the zone offset with the two-lookup rule */
int64_t orc_go_date(int64_t y, int64_t mo, int64_t d, int64_t h, int64_t mi, int64_t s, const orc_loc* loc) {
  int64_t m0 = mo - 1;
  y += fdiv(m0, 12);
  m0 = fmod_(m0, 12);
  int64_t unix_ = (days_from_civil(y, m0 + 1, 1) + (d - 1)) * 86400 + h * 3600 + mi * 60 + s;
  int32_t offset;
  int64_t start, end;
  orc_loc_lookup(loc, unix_, &offset, &start, &end);
  if (offset != 0) {
    int64_t utc = unix_ - offset;
    if (utc < start || utc >= end) orc_loc_lookup(loc, utc, &offset, &start, &end);
    unix_ -= offset;
  }
  return unix_;
}
static int64_t add_date(const orc_loc* loc, int64_t t, int years, int months, int days) {
  civil c = fields_of(loc, t);
  return orc_go_date(c.year + years, c.month + months, c.day + days, c.hour, c.minute, c.second, loc);
}

/* ------------------------------------------------------------------------------------------------
 * robfig/cron v3.0.1 spec.go
 * ---------------------------------------------------------------------------------------------- */
static __thread int64_t g_steps;
int64_t orc_last_steps(void) { return g_steps; }

static int day_matches(const orc_spec* s, const civil* c) {
  int dom_match = ((1ULL << (unsigned)c->day) & s->dom) > 0;
  int dow_match = ((1ULL << (unsigned)c->weekday) & s->dow) > 0;
  if ((s->dom & STAR_BIT) > 0 || (s->dow & STAR_BIT) > 0) return dom_match && dow_match;
  return dom_match || dow_match;
}

/* DELIBERATE DIVERGENCE FROM UPSTREAM (the only one): in zones that skipped a whole calendar day (Pacific/Kiritimati
 * 1994-12-31, Pacific/Apia 2011-12-30, ...) AddDate(0,0,1) from the day before resolves back to the same instant, so
 * upstream's `for !dayMatches(s, t)` loop never terminates -- real Go hangs.  A terminating loop entry needs at most
 * 31 (+ a few DST-fix) day steps, 12 month steps, 25 hour steps, 61 minute or second steps before it exits or wraps;
 * past ORC_STUCK iterations of one loop entry the restatement returns the zero time ("unschedulable cron"). */
#define ORC_STUCK 400

int64_t orc_spec_next(const orc_spec* s, const orc_loc* loc, int64_t t) {
  g_steps = 0;
  int guard;
  /* origLocation / s.Location handling: at this boundary every time is an absolute unix second and `loc` is the
   * effective location (s.Location, or t.Location() when s.Location == time.Local), so In()/In(orig) are no-ops. */
  t = t + 1; /* t.Add(1s - t.Nanosecond()) with whole-second inputs */
  int added = 0;
  civil c = fields_of(loc, t);
  int64_t year_limit = c.year + 5;

WRAP:
  c = fields_of(loc, t);
  if (c.year > year_limit) return ORC_ZERO;

  guard = 0;
  while (((1ULL << (unsigned)c.month) & s->month) == 0) {
    g_steps++;
    if (++guard > ORC_STUCK) return ORC_ZERO;
    if (!added) {
      added = 1;
      t = orc_go_date(c.year, c.month, 1, 0, 0, 0, loc);
    }
    t = add_date(loc, t, 0, 1, 0);
    c = fields_of(loc, t);
    if (c.month == 1) goto WRAP;
  }

  guard = 0;
  while (!day_matches(s, &c)) {
    g_steps++;
    if (++guard > ORC_STUCK) return ORC_ZERO;
    if (!added) {
      added = 1;
      t = orc_go_date(c.year, c.month, c.day, 0, 0, 0, loc);
    }
    t = add_date(loc, t, 0, 0, 1);
    c = fields_of(loc, t);
    /* "Notice if the hour is no longer midnight due to DST." */
    if (c.hour != 0) {
      if (c.hour > 12)
        t += (24 - c.hour) * 3600;
      else
        t -= c.hour * 3600;
      c = fields_of(loc, t);
    }
    if (c.day == 1) goto WRAP;
  }

  guard = 0;
  while (((1ULL << (unsigned)c.hour) & s->hour) == 0) {
    g_steps++;
    if (++guard > ORC_STUCK) return ORC_ZERO;
    if (!added) {
      added = 1;
      t = orc_go_date(c.year, c.month, c.day, c.hour, 0, 0, loc);
    }
    t += 3600;
    c = fields_of(loc, t);
    if (c.hour == 0) goto WRAP;
  }

  guard = 0;
  while (((1ULL << (unsigned)c.minute) & s->minute) == 0) {
    g_steps++;
    if (++guard > ORC_STUCK) return ORC_ZERO;
    if (!added) {
      added = 1;
      t = t - fmod_(t, 60); /* Truncate(time.Minute): absolute, since the zero Time is a multiple of 60 s from 1970 */
    }
    t += 60;
    c = fields_of(loc, t);
    if (c.minute == 0) goto WRAP;
  }

  guard = 0;
  while (((1ULL << (unsigned)c.second) & s->second) == 0) {
    g_steps++;
    if (++guard > ORC_STUCK) return ORC_ZERO;
    if (!added) {
      added = 1; /* Truncate(time.Second) is the identity on whole seconds */
    }
    t += 1;
    c = fields_of(loc, t);
    if (c.second == 0) goto WRAP;
  }
  return t;
}

/* ------------------------------------------------------------------------------------------------
 * robfig/cron v3.0.1 parser.go
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  unsigned min, max;
  int names; /* 0 none, 1 months, 2 dow */
} bounds;
static const bounds B_SECONDS = {0, 59, 0}, B_MINUTES = {0, 59, 0}, B_HOURS = {0, 23, 0}, B_DOM = {1, 31, 0},
                    B_MONTHS = {1, 12, 1}, B_DOW = {0, 6, 2};
static const char* MONTH_NAMES[12] = {"jan", "feb", "mar", "apr", "may", "jun", "jul", "aug", "sep", "oct", "nov", "dec"};
static const char* DOW_NAMES[7] = {"sun", "mon", "tue", "wed", "thu", "fri", "sat"};

typedef struct {
  char* err;
  size_t cap;
  int failed;
} perr;
static void pfail(perr* e, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
#include <stdarg.h>
static void pfail(perr* e, const char* fmt, ...) {
  if (e->failed) return;
  e->failed = 1;
  if (e->err && e->cap) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(e->err, e->cap, fmt, ap);
    va_end(ap);
  }
}

static uint64_t get_bits(unsigned min, unsigned max, unsigned step) {
  if (step == 1) {
    uint64_t hi = (max + 1 >= 64) ? 0 : (UINT64_MAX << (max + 1));
    return ~hi & (UINT64_MAX << min);
  }
  uint64_t bits = 0;
  for (unsigned i = min; i <= max; i += step) bits |= 1ULL << i;
  return bits;
}
static uint64_t all_bits(bounds r) { return get_bits(r.min, r.max, 1) | STAR_BIT; }

/* strconv.Atoi on [s, s+n): optional sign then decimal digits; error on empty/overflow */
static int go_atoi(const char* s, size_t n, long long* out) {
  size_t i = 0;
  int neg = 0;
  if (n == 0) return 0;
  if (s[0] == '+' || s[0] == '-') {
    neg = s[0] == '-';
    i = 1;
    if (n == 1) return 0;
  }
  unsigned long long v = 0;
  for (; i < n; i++) {
    if (s[i] < '0' || s[i] > '9') return 0;
    if (v > (1ULL << 63) / 10) return 0; /* would leave the int64 range */
    v = v * 10 + (unsigned)(s[i] - '0');
    if (v > (1ULL << 63)) return 0;
  }
  if (!neg && v > (unsigned long long)INT64_MAX) return 0;
  *out = neg ? -(long long)v : (long long)v;
  return 1;
}
static int must_parse_int(const char* s, size_t n, unsigned* out, perr* e) {
  long long num;
  if (!go_atoi(s, n, &num)) {
    pfail(e, "failed to parse int from %.*s: invalid syntax", (int)n, s);
    return 0;
  }
  if (num < 0) {
    pfail(e, "negative number (%lld) not allowed: %.*s", num, (int)n, s);
    return 0;
  }
  /* Go's uint is 64-bit; anything this large fails the range checks below just the same */
  *out = num > 0x7fffffff ? 0x7fffffffu : (unsigned)num;
  return 1;
}
static int parse_int_or_name(const char* s, size_t n, int names, unsigned* out, perr* e) {
  if (names && n == 3) {
    char low[4] = {(char)tolower((unsigned char)s[0]), (char)tolower((unsigned char)s[1]),
                   (char)tolower((unsigned char)s[2]), 0};
    if (names == 1) {
      for (unsigned i = 0; i < 12; i++)
        if (strcmp(low, MONTH_NAMES[i]) == 0) { *out = i + 1; return 1; }
    } else {
      for (unsigned i = 0; i < 7; i++)
        if (strcmp(low, DOW_NAMES[i]) == 0) { *out = i; return 1; }
    }
  }
  return must_parse_int(s, n, out, e);
}
/* strings.Split(expr, sep) for a single-byte sep: fills piece offsets, returns the count */
static int split_char(const char* s, size_t n, char sep, size_t* off, size_t* len, int maxp) {
  int k = 0;
  size_t st = 0;
  for (size_t i = 0; i <= n; i++) {
    if (i == n || s[i] == sep) {
      if (k < maxp) { off[k] = st; len[k] = i - st; }
      k++;
      st = i + 1;
    }
  }
  return k;
}
static uint64_t get_range(const char* expr, size_t n, bounds r, perr* e) {
  size_t ro[4], rl[4], lo_[4], ll[4];
  int nrs = split_char(expr, n, '/', ro, rl, 4);
  int nlh = split_char(expr + ro[0], rl[0], '-', lo_, ll, 4);
  const char* p0 = expr + ro[0] + lo_[0];
  int single_digit = nlh == 1;
  unsigned start = 0, end = 0, step = 0;
  uint64_t extra = 0;
  if (ll[0] == 1 && (p0[0] == '*' || p0[0] == '?')) {
    start = r.min;
    end = r.max;
    extra = STAR_BIT;
  } else {
    if (!parse_int_or_name(p0, ll[0], r.names, &start, e)) return 0;
    switch (nlh) {
      case 1: end = start; break;
      case 2:
        if (!parse_int_or_name(expr + ro[0] + lo_[1], ll[1], r.names, &end, e)) return 0;
        break;
      default: pfail(e, "too many hyphens: %.*s", (int)n, expr); return 0;
    }
  }
  switch (nrs) {
    case 1: step = 1; break;
    case 2:
      if (!must_parse_int(expr + ro[1], rl[1], &step, e)) return 0;
      if (single_digit) end = r.max; /* "N/step" means "N-max/step" */
      if (step > 1) extra = 0;
      break;
    default: pfail(e, "too many slashes: %.*s", (int)n, expr); return 0;
  }
  if (start < r.min) { pfail(e, "beginning of range (%u) below minimum (%u): %.*s", start, r.min, (int)n, expr); return 0; }
  if (end > r.max) { pfail(e, "end of range (%u) above maximum (%u): %.*s", end, r.max, (int)n, expr); return 0; }
  if (start > end) { pfail(e, "beginning of range (%u) beyond end of range (%u): %.*s", start, end, (int)n, expr); return 0; }
  if (step == 0) { pfail(e, "step of range should be a positive number: %.*s", (int)n, expr); return 0; }
  return get_bits(start, end, step) | extra;
}
static uint64_t get_field(const char* f, size_t n, bounds r, perr* e) {
  /* strings.FieldsFunc(field, r == ','): empty pieces are dropped */
  uint64_t bits = 0;
  size_t st = 0;
  for (size_t i = 0; i <= n; i++) {
    if (i == n || f[i] == ',') {
      if (i > st) {
        uint64_t b = get_range(f + st, i - st, r, e);
        if (e->failed) return bits;
        bits |= b;
      }
      st = i + 1;
    }
  }
  return bits;
}

static int is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }

/* time.ParseDuration restatement (format.go); result in nanoseconds */
static int leading_int(const char** ps, const char* end, uint64_t* x) {
  const char* s = *ps;
  uint64_t v = 0;
  while (s < end && *s >= '0' && *s <= '9') {
    if (v > (1ULL << 63) / 10) return 0;
    v = v * 10 + (uint64_t)(*s - '0');
    if (v > (1ULL << 63)) return 0;
    s++;
  }
  *x = v;
  *ps = s;
  return 1;
}
static int parse_duration(const char* s, size_t n, int64_t* out) {
  const char* end = s + n;
  uint64_t d = 0;
  int neg = 0;
  if (s < end && (*s == '-' || *s == '+')) {
    neg = *s == '-';
    s++;
  }
  if (end - s == 1 && *s == '0') { *out = 0; return 1; }
  if (s == end) return 0;
  while (s < end) {
    uint64_t v = 0, f = 0;
    double scale = 1;
    if (!(*s == '.' || (*s >= '0' && *s <= '9'))) return 0;
    const char* p = s;
    if (!leading_int(&s, end, &v)) return 0;
    int pre = p != s, post = 0;
    if (s < end && *s == '.') {
      s++;
      const char* q = s;
      int overflow = 0;
      while (s < end && *s >= '0' && *s <= '9') {
        if (!overflow) {
          if (f > ((1ULL << 63) - 1) / 10) {
            overflow = 1;
          } else {
            uint64_t y = f * 10 + (uint64_t)(*s - '0');
            if (y > (1ULL << 63)) {
              overflow = 1;
            } else {
              f = y;
              scale *= 10;
            }
          }
        }
        s++;
      }
      post = q != s;
    }
    if (!pre && !post) return 0;
    const char* u = s;
    while (s < end && !(*s == '.' || (*s >= '0' && *s <= '9'))) s++;
    if (s == u) return 0;
    size_t ul = (size_t)(s - u);
    uint64_t unit;
    if (ul == 2 && memcmp(u, "ns", 2) == 0) unit = 1;
    else if (ul == 2 && memcmp(u, "us", 2) == 0) unit = 1000;
    else if (ul == 3 && memcmp(u, "\xc2\xb5s", 3) == 0) unit = 1000;  /* U+00B5 */
    else if (ul == 3 && memcmp(u, "\xce\xbcs", 3) == 0) unit = 1000;  /* U+03BC */
    else if (ul == 2 && memcmp(u, "ms", 2) == 0) unit = 1000000;
    else if (ul == 1 && u[0] == 's') unit = 1000000000ULL;
    else if (ul == 1 && u[0] == 'm') unit = 60ULL * 1000000000ULL;
    else if (ul == 1 && u[0] == 'h') unit = 3600ULL * 1000000000ULL;
    else return 0;
    if (v > (1ULL << 63) / unit) return 0;
    v *= unit;
    if (f > 0) {
      v += (uint64_t)((double)f * ((double)unit / scale));
      if (v > (1ULL << 63)) return 0;
    }
    d += v;
    if (d > (1ULL << 63)) return 0;
  }
  if (neg) { *out = -(int64_t)d; return 1; }
  if (d > (1ULL << 63) - 1) return 0;
  *out = (int64_t)d;
  return 1;
}

static int parse_impl(const char* spec, size_t len, int with_seconds, orc_spec* out, int* kind, int64_t* every_sec,
                      char* tz, size_t tzcap, char* err, size_t errcap) {
  perr e = {err, errcap, 0};
  if (err && errcap) err[0] = 0;
  if (tz && tzcap) tz[0] = 0;
  *kind = ORC_KIND_SPEC;
  *every_sec = 0;
  memset(out, 0, sizeof *out);
  if (len == 0) { pfail(&e, "empty spec string"); return 1; }

  /* Extract timezone if present */
  if ((len >= 3 && memcmp(spec, "TZ=", 3) == 0) || (len >= 8 && memcmp(spec, "CRON_TZ=", 8) == 0)) {
    const char* sp = (const char*)memchr(spec, ' ', len);
    const char* eq = (const char*)memchr(spec, '=', len);
    if (!sp) { /* upstream v3.0.1 slices spec[eq+1:-1] here and panics; the restatement reports an error */
      pfail(&e, "provided bad location %.*s: missing schedule after time zone", (int)len, spec);
      return 1;
    }
    size_t nl = (size_t)(sp - (eq + 1));
    if (sp < eq + 1) nl = 0;
    if (tz && nl + 1 > tzcap) { pfail(&e, "provided bad location: name too long"); return 1; }
    if (tz) { memcpy(tz, eq + 1, nl); tz[nl] = 0; }
    if (nl == 0 && tz && tzcap >= 4) strcpy(tz, "UTC"); /* LoadLocation("") == UTC */
    size_t i = (size_t)(sp - spec);
    spec += i; len -= i;
    while (len && is_space(spec[0])) { spec++; len--; } /* strings.TrimSpace */
    while (len && is_space(spec[len - 1])) len--;
  }

  /* descriptors */
  if (len >= 1 && spec[0] == '@') {
    uint64_t S0 = 1ULL << B_SECONDS.min, M0 = 1ULL << B_MINUTES.min, H0 = 1ULL << B_HOURS.min;
#define IS(lit) (len == sizeof(lit) - 1 && memcmp(spec, lit, len) == 0)
    if (IS("@yearly") || IS("@annually")) {
      *out = (orc_spec){S0, M0, H0, 1ULL << B_DOM.min, 1ULL << B_MONTHS.min, all_bits(B_DOW)};
      return 0;
    }
    if (IS("@monthly")) { *out = (orc_spec){S0, M0, H0, 1ULL << B_DOM.min, all_bits(B_MONTHS), all_bits(B_DOW)}; return 0; }
    if (IS("@weekly")) { *out = (orc_spec){S0, M0, H0, all_bits(B_DOM), all_bits(B_MONTHS), 1ULL << B_DOW.min}; return 0; }
    if (IS("@daily") || IS("@midnight")) { *out = (orc_spec){S0, M0, H0, all_bits(B_DOM), all_bits(B_MONTHS), all_bits(B_DOW)}; return 0; }
    if (IS("@hourly")) { *out = (orc_spec){S0, M0, all_bits(B_HOURS), all_bits(B_DOM), all_bits(B_MONTHS), all_bits(B_DOW)}; return 0; }
#undef IS
    if (len >= 7 && memcmp(spec, "@every ", 7) == 0) {
      int64_t ns;
      if (!parse_duration(spec + 7, len - 7, &ns)) {
        pfail(&e, "failed to parse duration %.*s", (int)len, spec);
        return 1;
      }
      /* constantdelay.go Every(): clamp to >= 1s, drop the sub-second part */
      if (ns < 1000000000LL) ns = 1000000000LL;
      *kind = ORC_KIND_EVERY;
      *every_sec = ns / 1000000000LL;
      *out = (orc_spec){0, 0, 0, (uint64_t)*every_sec, ORC_EVERY_TAG, 0}; /* batch encoding, see cron_oracle.h */
      return 0;
    }
    pfail(&e, "unrecognized descriptor: %.*s", (int)len, spec);
    return 1;
  }

  /* strings.Fields */
  const char* fp[8];
  size_t fl[8];
  int nf = 0;
  size_t i = 0;
  while (i < len) {
    while (i < len && is_space(spec[i])) i++;
    if (i >= len) break;
    size_t st = i;
    while (i < len && !is_space(spec[i])) i++;
    if (nf < 8) { fp[nf] = spec + st; fl[nf] = i - st; }
    nf++;
  }
  int want = with_seconds ? 6 : 5;
  if (nf != want) {
    pfail(&e, "expected exactly %d fields, found %d: %.*s", want, nf, (int)len, spec);
    return 1;
  }
  /* normalizeFields: defaults {"0","0","0","*","*","*"} for places not in the options */
  const char* ff[6];
  size_t ffl[6];
  if (with_seconds) {
    for (int k = 0; k < 6; k++) { ff[k] = fp[k]; ffl[k] = fl[k]; }
  } else {
    ff[0] = "0"; ffl[0] = 1;
    for (int k = 0; k < 5; k++) { ff[k + 1] = fp[k]; ffl[k + 1] = fl[k]; }
  }
  out->second = get_field(ff[0], ffl[0], B_SECONDS, &e);
  if (!e.failed) out->minute = get_field(ff[1], ffl[1], B_MINUTES, &e);
  if (!e.failed) out->hour = get_field(ff[2], ffl[2], B_HOURS, &e);
  if (!e.failed) out->dom = get_field(ff[3], ffl[3], B_DOM, &e);
  if (!e.failed) out->month = get_field(ff[4], ffl[4], B_MONTHS, &e);
  if (!e.failed) out->dow = get_field(ff[5], ffl[5], B_DOW, &e);
  if (e.failed) { memset(out, 0, sizeof *out); return 1; }
  return 0;
}
int orc_parse_standard(const char* s, size_t len, orc_spec* out, int* kind, int64_t* every_sec, char* tz, size_t tzcap,
                       char* err, size_t errcap) {
  return parse_impl(s, len, 0, out, kind, every_sec, tz, tzcap, err, errcap);
}
int orc_parse_with_seconds(const char* s, size_t len, orc_spec* out, int* kind, int64_t* every_sec, char* tz,
                           size_t tzcap, char* err, size_t errcap) {
  return parse_impl(s, len, 1, out, kind, every_sec, tz, tzcap, err, errcap);
}

/* constantdelay.go: func (schedule ConstantDelaySchedule) Next(t time.Time) time.Time {
 *   return t.Add(schedule.Delay - time.Duration(t.Nanosecond())*time.Nanosecond) }
 * Times are whole seconds at this boundary, so the nanosecond term is zero.  Every() never builds a Delay below one
 * second; a hand-built plane that says so is clamped the same way. */
int64_t orc_sched_next(const orc_spec* s, const orc_loc* loc, int64_t t) {
  if (s->month == ORC_EVERY_TAG) {
    int64_t delay = (int64_t)s->dom;
    if (delay < 1) delay = 1;
    g_steps = 0;
    return t + delay;
  }
  return orc_spec_next(s, loc, t);
}

/* ------------------------------------------------------------------------------------------------
 * internal/controller/cron_controller.go
 * ---------------------------------------------------------------------------------------------- */
/* getNextSchedule, cron_controller.go:389-437 (the parse at :392 is done by the caller) */
int orc_get_next_schedule(const orc_spec* s, const orc_loc* loc, int64_t earliest, int64_t now, int64_t* last_missed,
                          int64_t* next, uint64_t* missed) {
  *last_missed = ORC_ZERO;
  *next = ORC_ZERO;
  *missed = 0;
  if (earliest > now) { /* :404-406 */
    *next = orc_sched_next(s, loc, now);
    return ORC_ST_EARLIEST_AFTER_NOW;
  }
  uint64_t missed_times = 0;
  int64_t last = ORC_ZERO;
  /* :409  for t := sched.Next(earliest); !t.After(now); t = sched.Next(t)   (zero time is "not after now")
   *
   * DELIBERATE DIVERGENCE (second and last): across a DST overlap a chained Next() can return a time at or before
   * its argument (Pacific/Chatham: a fire found at 02:59 on the repeated hour makes the next call reset to 02:00 and
   * return the first 03:00 again) and upstream's loop then cycles forever.  Next() is a pure function of t, so the
   * loop is in a cycle exactly when it revisits a fire time: after each backward step the largest origin is kept and
   * a fire equal to it ends the call with ORC_ST_NONTERMINATING. */
  int64_t anchor = ORC_ZERO;
  int64_t t = orc_sched_next(s, loc, earliest);
  while (t == ORC_ZERO || t <= now) {
    if (t == ORC_ZERO) return ORC_ST_UNSCHEDULABLE; /* :410-412 */
    last = t;
    missed_times++;
    int64_t tn = orc_sched_next(s, loc, t);
    if (tn != ORC_ZERO && tn <= t && t > anchor) anchor = t;
    if (anchor != ORC_ZERO && tn == anchor) return ORC_ST_NONTERMINATING;
    t = tn;
  }
  *last_missed = last;
  *missed = missed_times;
  *next = orc_sched_next(s, loc, now); /* :436 */
  return ORC_ST_OK;
}

/* Reconcile gates :163-180 and decision :184-220 */
int orc_reconcile_decide(const orc_spec* s, const orc_loc* loc, int64_t earliest, int64_t now, int32_t now_nsec,
                         uint8_t policy, uint32_t active, uint8_t flags, int64_t deadline, int64_t* last_missed,
                         int64_t* next, uint64_t* missed, uint8_t* status) {
  *last_missed = ORC_ZERO;
  *next = ORC_ZERO;
  *missed = 0;
  *status = ORC_ST_OK;
  if (flags & ORC_FLAG_DELETED) return ORC_DEC_GATE_DELETED;   /* :163-166 */
  if (flags & ORC_FLAG_SUSPEND) return ORC_DEC_GATE_SUSPENDED; /* :169-173 */
  /* :176  now.After(deadline) with now = now + now_nsec ns and a whole-second deadline */
  if (deadline != ORC_ZERO && (now > deadline || (now == deadline && now_nsec > 0))) return ORC_DEC_GATE_DEADLINE;
  int st = orc_get_next_schedule(s, loc, earliest, now, last_missed, next, missed);
  *status = (uint8_t)st;
  if (st == ORC_ST_UNSCHEDULABLE || st == ORC_ST_NONTERMINATING) return ORC_DEC_ERR_UNSCHEDULABLE; /* :185-190 */
  if (*last_missed == ORC_ZERO) return ORC_DEC_REQUEUE_ONLY;             /* :196-199 */
  if (policy == ORC_POLICY_FORBID && active > 0) return ORC_DEC_SKIP_FORBID; /* :204-207 */
  if (policy == ORC_POLICY_REPLACE) return ORC_DEC_REPLACE_CREATE;       /* :210-220 */
  return ORC_DEC_CREATE;                                                 /* :222-238 */
}

/* sortByCreationTimestamp: slices.SortStableFunc ascending (cron_util.go:117-129); stable insertion sort here.
 * Prune rule: i < n - historyLimit is deleted (cron_controller.go:313-323). */
void orc_history_prune(const int64_t* key, size_t n, int64_t limit, uint32_t* perm, uint8_t* keep) {
  if (n <= 64) {
    for (size_t i = 0; i < n; i++) {
      uint32_t v = (uint32_t)i;
      size_t j = i;
      while (j > 0 && key[perm[j - 1]] > key[v]) {
        perm[j] = perm[j - 1];
        j--;
      }
      perm[j] = v;
    }
  } else { /* bottom-up stable merge sort for long segments */
    uint32_t* tmp = (uint32_t*)malloc(n * sizeof(uint32_t));
    uint32_t *a = perm, *b = tmp;
    for (size_t i = 0; i < n; i++) a[i] = (uint32_t)i;
    for (size_t w = 1; w < n; w *= 2) {
      for (size_t lo = 0; lo < n; lo += 2 * w) {
        size_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
        size_t i = lo, j = mid, o = lo;
        while (i < mid && j < hi) b[o++] = (key[a[j]] < key[a[i]]) ? a[j++] : a[i++];
        while (i < mid) b[o++] = a[i++];
        while (j < hi) b[o++] = a[j++];
      }
      uint32_t* t = a; a = b; b = t;
    }
    if (a != perm) memcpy(perm, a, n * sizeof(uint32_t));
    free(tmp);
  }
  for (size_t i = 0; i < n; i++) {
    if (limit < 0)
      keep[i] = 1; /* ptr.Deref(HistoryLimit, math.MaxInt) */
    else
      keep[i] = !((int64_t)i < (int64_t)n - limit);
  }
}

/* ------------------------------------------------------------------------------------------------
 * batched drivers (pthread, one contiguous shard per thread)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int which;
  size_t lo, hi;
  const uint64_t* const* masks;
  const int32_t* zone_id;
  const orc_loc* const* locs;
  const int64_t* from;
  int k;
  int64_t* out;
  int64_t now;
  int32_t now_nsec;
  const uint8_t* policy;
  const uint32_t* active;
  const uint8_t* flags;
  const int64_t* deadline;
  int64_t *last, *next;
  uint32_t* missed;
  uint8_t *status, *decision;
  const uint64_t* seg_off;
  const int64_t* key;
  const int64_t* limit;
  uint32_t* perm;
  uint8_t* keep;
} job;

static void* worker(void* p) {
  job* j = (job*)p;
  for (size_t i = j->lo; i < j->hi; i++) {
    if (j->which == 2) {
      uint64_t a = j->seg_off[i], b = j->seg_off[i + 1];
      orc_history_prune(j->key + a, (size_t)(b - a), j->limit ? j->limit[i] : -1, j->perm + a, j->keep + a);
      continue;
    }
    orc_spec s = {j->masks[0][i], j->masks[1][i], j->masks[2][i], j->masks[3][i], j->masks[4][i], j->masks[5][i]};
    const orc_loc* loc = (j->zone_id && j->locs) ? j->locs[j->zone_id[i]] : NULL;
    if (j->which == 0) {
      int64_t t = j->from[i];
      for (int q = 0; q < j->k; q++) {
        if (t != ORC_ZERO) t = orc_sched_next(&s, loc, t);
        j->out[i * (size_t)j->k + q] = t;
      }
    } else {
      uint64_t m;
      uint8_t st;
      int d = orc_reconcile_decide(&s, loc, j->from[i], j->now, j->now_nsec, j->policy ? j->policy[i] : 0,
                                   j->active ? j->active[i] : 0, j->flags ? j->flags[i] : 0,
                                   j->deadline ? j->deadline[i] : ORC_ZERO, &j->last[i], &j->next[i], &m, &st);
      j->missed[i] = m > 0xffffffffULL ? 0xffffffffu : (uint32_t)m;
      j->status[i] = st;
      j->decision[i] = (uint8_t)d;
    }
  }
  return NULL;
}
static void run_jobs(job* proto, size_t n, int threads) {
  if (threads < 1) threads = 1;
  if ((size_t)threads > n) threads = n ? (int)n : 1;
  pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * threads);
  job* js = (job*)malloc(sizeof(job) * threads);
  for (int t = 0; t < threads; t++) {
    js[t] = *proto;
    js[t].lo = n * (size_t)t / (size_t)threads;
    js[t].hi = n * (size_t)(t + 1) / (size_t)threads;
    if (threads == 1)
      worker(&js[t]);
    else
      pthread_create(&th[t], NULL, worker, &js[t]);
  }
  if (threads > 1)
    for (int t = 0; t < threads; t++) pthread_join(th[t], NULL);
  free(th);
  free(js);
}
void orc_batch_next_n(const uint64_t* const masks[6], const int32_t* zone_id, const orc_loc* const* locs,
                      const int64_t* from, size_t n, int k, int64_t* out, int threads) {
  job j;
  memset(&j, 0, sizeof j);
  j.which = 0; j.masks = masks; j.zone_id = zone_id; j.locs = locs; j.from = from; j.k = k; j.out = out;
  run_jobs(&j, n, threads);
}
void orc_batch_unmet(const uint64_t* const masks[6], const int32_t* zone_id, const orc_loc* const* locs,
                     const int64_t* earliest, int64_t now, int32_t now_nsec, const uint8_t* policy,
                     const uint32_t* active, const uint8_t* flags, const int64_t* deadline, size_t n, int64_t* last,
                     int64_t* next, uint32_t* missed, uint8_t* status, uint8_t* decision, int threads) {
  job j;
  memset(&j, 0, sizeof j);
  j.which = 1; j.masks = masks; j.zone_id = zone_id; j.locs = locs; j.from = earliest; j.now = now;
  j.now_nsec = now_nsec; j.policy = policy; j.active = active; j.flags = flags; j.deadline = deadline;
  j.last = last; j.next = next; j.missed = missed; j.status = status; j.decision = decision;
  run_jobs(&j, n, threads);
}
void orc_batch_history(const uint64_t* seg_off, const int64_t* key, const int64_t* limit, size_t nseg, uint32_t* perm,
                       uint8_t* keep, int threads) {
  job j;
  memset(&j, 0, sizeof j);
  j.which = 2; j.seg_off = seg_off; j.key = key; j.limit = limit; j.perm = perm; j.keep = keep;
  run_jobs(&j, nseg, threads);
}

/* enumeration of the missed runs themselves (cron_controller.go:409-430 lists nothing; this follows the same chain):
 * list_off = exclusive prefix sum of min(missed, cap) over Crons whose decision is >= REQUEUE_ONLY, list = those
 * Crons' first min(missed, cap) fire times.  Returns the total; writes nothing to `list` beyond list_cap. */
uint64_t orc_batch_unmet_list(const uint64_t* const masks[6], const int32_t* zone_id, const orc_loc* const* locs,
                              const int64_t* earliest, const uint32_t* missed, const uint8_t* decision, size_t n,
                              uint32_t cap, uint64_t* list_off, int64_t* list, size_t list_cap) {
  uint64_t run = 0;
  for (size_t i = 0; i < n; i++) {
    list_off[i] = run;
    uint32_t c = decision[i] >= ORC_DEC_REQUEUE_ONLY ? (missed[i] < cap ? missed[i] : cap) : 0;
    if (c) {
      orc_spec s = {masks[0][i], masks[1][i], masks[2][i], masks[3][i], masks[4][i], masks[5][i]};
      const orc_loc* loc = (zone_id && locs) ? locs[zone_id[i]] : NULL;
      int64_t t = earliest[i];
      for (uint32_t j = 0; j < c; j++) {
        t = orc_sched_next(&s, loc, t);
        if (run + j < list_cap) list[run + j] = t;
      }
    }
    run += c;
  }
  list_off[n] = run;
  return run;
}
