// parse.cpp -- host-side schedule parser of libb200cron.so.
//
// Replaces cronv3.ParseStandard [robfig/cron v3.0.1 parser.go; un-vendored, go.mod:10], which the reference calls on
// every reconcile at internal/controller/cron_controller.go:392.  Grammar and error classes follow upstream:
// optional "TZ="/"CRON_TZ=" prefix, "@" descriptors incl. "@every <duration>", exactly 5 whitespace-separated
// fields (6 with the seconds extension), per field a comma list of `lo[-hi][/step]` terms where `*` and `?` mean
// the full range and set the star bit, a step > 1 clears it, `N/step` means `N-max/step`, and month / day-of-week
// accept three-letter names.
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <string_view>
#include <vector>

#include "../../include/b200cron.h"

namespace {

using sv = std::string_view;
constexpr uint64_t kStar = 1ULL << 63;

struct Range {
  unsigned lo, hi;
  const char* const* names;
  unsigned name_base, name_count;
};
const char* const kMonths[] = {"jan", "feb", "mar", "apr", "may", "jun", "jul", "aug", "sep", "oct", "nov", "dec"};
const char* const kDays[] = {"sun", "mon", "tue", "wed", "thu", "fri", "sat"};
const Range kSecond{0, 59, nullptr, 0, 0}, kMinute{0, 59, nullptr, 0, 0}, kHour{0, 23, nullptr, 0, 0},
    kDom{1, 31, nullptr, 0, 0}, kMonth{1, 12, kMonths, 1, 12}, kDow{0, 6, kDays, 0, 7};

struct Failure {
  std::string text;
  bool set = false;
  void operator()(const char* fmt, ...) __attribute__((format(printf, 2, 3))) {
    if (set) return;
    set = true;
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    text = buf;
  }
};

uint64_t bits_of(unsigned lo, unsigned hi, unsigned step) {
  uint64_t b = 0;
  for (unsigned v = lo; v <= hi; v += step) b |= 1ULL << v;
  return b;
}
uint64_t full(const Range& r) { return bits_of(r.lo, r.hi, 1) | kStar; }

std::vector<sv> split(sv s, char sep) {  // strings.Split: keeps empty pieces
  std::vector<sv> out;
  size_t st = 0;
  for (size_t i = 0; i <= s.size(); ++i)
    if (i == s.size() || s[i] == sep) {
      out.push_back(s.substr(st, i - st));
      st = i + 1;
    }
  return out;
}

// strconv.Atoi: optional sign, decimal digits, int64 range
bool atoi_go(sv s, long long* out) {
  if (s.empty()) return false;
  size_t i = 0;
  bool neg = false;
  if (s[0] == '+' || s[0] == '-') {
    neg = s[0] == '-';
    i = 1;
  }
  if (i == s.size()) return false;
  unsigned long long v = 0;
  for (; i < s.size(); ++i) {
    if (s[i] < '0' || s[i] > '9') return false;
    if (v > (~0ULL - 9) / 10) return false;
    v = v * 10 + (unsigned)(s[i] - '0');
  }
  if (neg ? v > (1ULL << 63) : v > (1ULL << 63) - 1) return false;
  *out = neg ? (long long)(0 - v) : (long long)v;
  return true;
}

bool number(sv s, unsigned* out, Failure& fail) {  // mustParseInt
  long long v;
  if (!atoi_go(s, &v)) {
    fail("failed to parse int from %.*s: invalid syntax", (int)s.size(), s.data());
    return false;
  }
  if (v < 0) {
    fail("negative number (%lld) not allowed: %.*s", v, (int)s.size(), s.data());
    return false;
  }
  *out = v > 1000000 ? 1000000u : (unsigned)v;  // anything this large fails the bounds checks regardless
  return true;
}

bool number_or_name(sv s, const Range& r, unsigned* out, Failure& fail) {  // parseIntOrName
  if (r.names && s.size() == 3) {
    char low[3];
    for (int i = 0; i < 3; ++i) low[i] = (s[i] >= 'A' && s[i] <= 'Z') ? (char)(s[i] + 32) : s[i];
    for (unsigned i = 0; i < r.name_count; ++i)
      if (std::memcmp(low, r.names[i], 3) == 0) {
        *out = r.name_base + i;
        return true;
      }
  }
  return number(s, out, fail);
}

uint64_t term(sv expr, const Range& r, Failure& fail) {  // getRange
  const auto by_slash = split(expr, '/');
  const auto by_dash = split(by_slash[0], '-');
  const bool single = by_dash.size() == 1;
  unsigned lo = 0, hi = 0, step = 1;
  uint64_t extra = 0;
  if (by_dash[0] == "*" || by_dash[0] == "?") {
    lo = r.lo;
    hi = r.hi;
    extra = kStar;
  } else {
    if (!number_or_name(by_dash[0], r, &lo, fail)) return 0;
    if (by_dash.size() == 1) {
      hi = lo;
    } else if (by_dash.size() == 2) {
      if (!number_or_name(by_dash[1], r, &hi, fail)) return 0;
    } else {
      fail("too many hyphens: %.*s", (int)expr.size(), expr.data());
      return 0;
    }
  }
  if (by_slash.size() == 2) {
    if (!number(by_slash[1], &step, fail)) return 0;
    if (single) hi = r.hi;
    if (step > 1) extra = 0;
  } else if (by_slash.size() > 2) {
    fail("too many slashes: %.*s", (int)expr.size(), expr.data());
    return 0;
  }
  const int n = (int)expr.size();
  if (lo < r.lo) { fail("beginning of range (%u) below minimum (%u): %.*s", lo, r.lo, n, expr.data()); return 0; }
  if (hi > r.hi) { fail("end of range (%u) above maximum (%u): %.*s", hi, r.hi, n, expr.data()); return 0; }
  if (lo > hi) { fail("beginning of range (%u) beyond end of range (%u): %.*s", lo, hi, n, expr.data()); return 0; }
  if (step == 0) { fail("step of range should be a positive number: %.*s", n, expr.data()); return 0; }
  return bits_of(lo, hi, step) | extra;
}

uint64_t field(sv f, const Range& r, Failure& fail) {  // getField: FieldsFunc on ',' drops empty pieces
  uint64_t b = 0;
  for (sv piece : split(f, ',')) {
    if (piece.empty()) continue;
    const uint64_t t = term(piece, r, fail);
    if (fail.set) return 0;
    b |= t;
  }
  return b;
}

bool space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }

// time.ParseDuration -> nanoseconds
bool duration(sv s, int64_t* out) {
  static const struct { const char* u; uint64_t ns; } kUnits[] = {
      {"ns", 1}, {"us", 1000}, {"\xc2\xb5s", 1000}, {"\xce\xbcs", 1000}, {"ms", 1000000},
      {"s", 1000000000ULL}, {"m", 60000000000ULL}, {"h", 3600000000000ULL}};
  constexpr uint64_t kTop = 1ULL << 63;
  bool neg = false;
  if (!s.empty() && (s[0] == '-' || s[0] == '+')) {
    neg = s[0] == '-';
    s.remove_prefix(1);
  }
  if (s == "0") { *out = 0; return true; }
  if (s.empty()) return false;
  uint64_t total = 0;
  while (!s.empty()) {
    if (!(s[0] == '.' || (s[0] >= '0' && s[0] <= '9'))) return false;
    uint64_t whole = 0, frac = 0;
    double scale = 1;
    size_t i = 0;
    for (; i < s.size() && s[i] >= '0' && s[i] <= '9'; ++i) {
      if (whole > kTop / 10) return false;
      whole = whole * 10 + (uint64_t)(s[i] - '0');
      if (whole > kTop) return false;
    }
    const bool pre = i > 0;
    s.remove_prefix(i);
    bool post = false;
    if (!s.empty() && s[0] == '.') {
      s.remove_prefix(1);
      bool over = false;
      for (i = 0; i < s.size() && s[i] >= '0' && s[i] <= '9'; ++i) {
        if (over) continue;
        if (frac > (kTop - 1) / 10) { over = true; continue; }
        const uint64_t y = frac * 10 + (uint64_t)(s[i] - '0');
        if (y > kTop) { over = true; continue; }
        frac = y;
        scale *= 10;
      }
      post = i > 0;
      s.remove_prefix(i);
    }
    if (!pre && !post) return false;
    for (i = 0; i < s.size() && !(s[i] == '.' || (s[i] >= '0' && s[i] <= '9')); ++i) {}
    if (i == 0) return false;
    const sv unit = s.substr(0, i);
    s.remove_prefix(i);
    uint64_t ns = 0;
    for (const auto& u : kUnits)
      if (unit == u.u) ns = u.ns;
    if (ns == 0) return false;
    if (whole > kTop / ns) return false;
    whole *= ns;
    if (frac > 0) {
      whole += (uint64_t)((double)frac * ((double)ns / scale));
      if (whole > kTop) return false;
    }
    total += whole;
    if (total > kTop) return false;
  }
  if (neg) { *out = (int64_t)(0 - total); return true; }
  if (total > kTop - 1) return false;
  *out = (int64_t)total;
  return true;
}

void put(char* dst, size_t cap, const std::string& s) {
  if (!dst || cap == 0) return;
  const size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
  std::memcpy(dst, s.data(), n);
  dst[n] = 0;
}

int parse(sv spec, bool seconds, b200cron_spec* out, int* kind, int64_t* every, char* tz, size_t tzcap, char* err,
          size_t errcap) {
  Failure fail;
  b200cron_spec z{};
  *out = z;
  if (kind) *kind = B200CRON_KIND_SPEC;
  if (every) *every = 0;
  put(tz, tzcap, "");
  put(err, errcap, "");
  auto done = [&]() {
    if (!fail.set) return (int)B200CRON_OK;
    *out = z;
    put(err, errcap, fail.text);
    return (int)B200CRON_EPARSE;
  };
  if (spec.empty()) { fail("empty spec string"); return done(); }

  if (spec.substr(0, 3) == "TZ=" || spec.substr(0, 8) == "CRON_TZ=") {
    const size_t sp = spec.find(' ');
    const size_t eq = spec.find('=');
    if (sp == sv::npos) {  // upstream v3.0.1 panics on spec[eq+1:-1]; a library must not, so it is a parse error
      fail("provided bad location %.*s: no schedule after the time zone", (int)spec.size(), spec.data());
      return done();
    }
    std::string name(spec.substr(eq + 1, sp - eq - 1));
    if (name.empty()) name = "UTC";  // time.LoadLocation("") is UTC
    if (name.size() + 1 > tzcap && tz) { fail("provided bad location: name too long"); return done(); }
    put(tz, tzcap, name);
    spec.remove_prefix(sp);
    while (!spec.empty() && space(spec.front())) spec.remove_prefix(1);
    while (!spec.empty() && space(spec.back())) spec.remove_suffix(1);
  }

  if (!spec.empty() && spec[0] == '@') {
    const uint64_t one = 1;
    auto set = [&](uint64_t h, uint64_t d, uint64_t mo, uint64_t w) {
      *out = b200cron_spec{one << kSecond.lo, one << kMinute.lo, h, d, mo, w};
    };
    if (spec == "@yearly" || spec == "@annually") set(one << kHour.lo, one << kDom.lo, one << kMonth.lo, full(kDow));
    else if (spec == "@monthly") set(one << kHour.lo, one << kDom.lo, full(kMonth), full(kDow));
    else if (spec == "@weekly") set(one << kHour.lo, full(kDom), full(kMonth), one << kDow.lo);
    else if (spec == "@daily" || spec == "@midnight") set(one << kHour.lo, full(kDom), full(kMonth), full(kDow));
    else if (spec == "@hourly") set(full(kHour), full(kDom), full(kMonth), full(kDow));
    else if (spec.substr(0, 7) == "@every ") {
      int64_t ns;
      if (!duration(spec.substr(7), &ns)) {
        fail("failed to parse duration %.*s", (int)spec.size(), spec.data());
        return done();
      }
      if (ns < 1000000000LL) ns = 1000000000LL;  // constantdelay.go Every(): at least one second, whole seconds
      if (kind) *kind = B200CRON_KIND_EVERY;
      if (every) *every = ns / 1000000000LL;
      // the batch encoding of a ConstantDelaySchedule (include/b200cron.h): ready to drop into the six planes
      *out = b200cron_spec{0, 0, 0, (uint64_t)(ns / 1000000000LL), B200CRON_EVERY_TAG, 0};
    } else {
      fail("unrecognized descriptor: %.*s", (int)spec.size(), spec.data());
    }
    return done();
  }

  std::vector<sv> f;  // strings.Fields
  for (size_t i = 0; i < spec.size();) {
    while (i < spec.size() && space(spec[i])) ++i;
    const size_t st = i;
    while (i < spec.size() && !space(spec[i])) ++i;
    if (i > st) f.push_back(spec.substr(st, i - st));
  }
  const size_t want = seconds ? 6 : 5;
  if (f.size() != want) {
    fail("expected exactly %zu fields, found %zu: %.*s", want, f.size(), (int)spec.size(), spec.data());
    return done();
  }
  if (!seconds) f.insert(f.begin(), sv("0"));  // normalizeFields: the seconds place defaults to "0"
  const Range* rs[6] = {&kSecond, &kMinute, &kHour, &kDom, &kMonth, &kDow};
  uint64_t m[6];
  for (int i = 0; i < 6 && !fail.set; ++i) m[i] = field(f[i], *rs[i], fail);
  if (!fail.set) *out = b200cron_spec{m[0], m[1], m[2], m[3], m[4], m[5]};
  return done();
}

}  // namespace

extern "C" int b200cron_parse_standard(const char* spec, size_t len, b200cron_spec* out, int* kind, int64_t* every_sec,
                                       char* tz, size_t tzcap, char* err, size_t errcap) {
  if (!spec || !out) return B200CRON_EINVAL;
  return parse(sv(spec, len), false, out, kind, every_sec, tz, tzcap, err, errcap);
}
extern "C" int b200cron_parse_with_seconds(const char* spec, size_t len, b200cron_spec* out, int* kind,
                                           int64_t* every_sec, char* tz, size_t tzcap, char* err, size_t errcap) {
  if (!spec || !out) return B200CRON_EINVAL;
  return parse(sv(spec, len), true, out, kind, every_sec, tz, tzcap, err, errcap);
}
