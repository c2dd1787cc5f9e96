// parse_kernels.cuh -- batched device-side cronv3.ParseStandard.
//
// The reference re-parses every Cron's schedule text on every reconcile (internal/controller/cron_controller.go:392,
// no cache); at batch sizes where evaluation takes microseconds, host parsing becomes the bottleneck.  One thread
// parses one spec with the grammar of robfig/cron v3.0.1 parser.go (same accept/reject decisions as csrc/parse.cpp,
// which is checked against the oracle): optional TZ=/CRON_TZ= prefix, @descriptors, exactly 5 (or 6) whitespace
// separated fields, comma lists of lo[-hi][/step] terms, `*`/`?`, names for month and day-of-week.
// Zone names cannot be resolved on the device: the span of the name inside the spec is returned and the host maps
// it to a zone id.  "@every <duration>" with plain h/m/s terms is parsed here; other durations are flagged (status 2)
// and left to the host.
//
// The grammar lives in parse_one(), which also compiles as plain C++ (PK_HD / PK_LDG below): tests/host_shim builds it
// with g++ so the CPU suite can fuzz the very source the kernel runs against the oracle's parser.
#pragma once
#include <stdint.h>
#ifdef __CUDACC__
#include <cuda_runtime.h>
#define PK_HD __device__
#define PK_LDG(p) __ldg(p)
#else
#define PK_HD
#define PK_LDG(p) (*(p))
#ifndef __forceinline__
#define __forceinline__ inline
#endif
#endif

namespace b200cron {

enum : uint8_t { kParseOk = 0, kParseError = 1, kParseEvery = 2 };

struct Str {
  const char* p;
  int n;
};
PK_HD __forceinline__ char ch(const Str& s, int i) { return PK_LDG(s.p + i); }
PK_HD __forceinline__ bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\v' || c == '\f' || c == '\r'; }
PK_HD __forceinline__ bool equals(const Str& s, const char* lit, int n) {
  if (s.n != n) return false;
  for (int i = 0; i < n; ++i)
    if (ch(s, i) != lit[i]) return false;
  return true;
}
PK_HD __forceinline__ bool starts_with(const Str& s, const char* lit, int n) {
  if (s.n < n) return false;
  for (int i = 0; i < n; ++i)
    if (ch(s, i) != lit[i]) return false;
  return true;
}

// strconv.Atoi + mustParseInt: optional sign, digits, int64 range, negative values rejected.  Values are capped at
// 1e6 (anything that large fails the range checks that follow).
PK_HD inline bool parse_number(const Str& s, uint32_t* out) {
  if (s.n == 0) return false;
  int i = 0;
  bool neg = false;
  const char c0 = ch(s, 0);
  if (c0 == '+' || c0 == '-') { neg = c0 == '-'; i = 1; }
  if (i == s.n) return false;
  uint64_t v = 0;
  for (; i < s.n; ++i) {
    const char c = ch(s, i);
    if (c < '0' || c > '9') return false;
    if (v > (0xFFFFFFFFFFFFFFFFULL - 9) / 10) return false;
    v = v * 10 + (uint64_t)(c - '0');
  }
  if (neg ? v > (1ULL << 63) : v > (1ULL << 63) - 1) return false;
  if (neg && v > 0) return false;  // "negative number not allowed"
  *out = v > 1000000 ? 1000000u : (uint32_t)v;
  return true;
}

PK_HD __forceinline__ char lower(char c) { return (c >= 'A' && c <= 'Z') ? (char)(c + 32) : c; }

// parseIntOrName; names: 0 none, 1 months (jan..dec -> 1..12), 2 weekdays (sun..sat -> 0..6)
PK_HD inline bool parse_number_or_name(const Str& s, int names, uint32_t* out) {
  if (names && s.n == 3) {
    const char a = lower(ch(s, 0)), b = lower(ch(s, 1)), c = lower(ch(s, 2));
    const char* tab = names == 1 ? "janfebmaraprmayjunjulaugsepoctnovdec" : "sunmontuewedthufrisat";
    const int cnt = names == 1 ? 12 : 7;
    for (int k = 0; k < cnt; ++k)
      if (tab[3 * k] == a && tab[3 * k + 1] == b && tab[3 * k + 2] == c) {
        *out = names == 1 ? (uint32_t)(k + 1) : (uint32_t)k;
        return true;
      }
  }
  return parse_number(s, out);
}

// getRange: one `lo[-hi][/step]` term -> bits (with the star bit); false on any upstream error
PK_HD inline bool parse_term(const Str& e, uint32_t lo_b, uint32_t hi_b, int names, uint64_t* bits) {
  int slash = -1, nslash = 0;
  for (int i = 0; i < e.n; ++i)
    if (ch(e, i) == '/') { if (nslash == 0) slash = i; ++nslash; }
  if (nslash > 1) return false;  // "too many slashes"
  const Str r{e.p, nslash ? slash : e.n};
  int dash = -1, ndash = 0;
  for (int i = 0; i < r.n; ++i)
    if (ch(r, i) == '-') { if (ndash == 0) dash = i; ++ndash; }
  const Str first{r.p, ndash ? dash : r.n};
  uint32_t lo = 0, hi = 0, step = 1;
  uint64_t extra = 0;
  if (first.n == 1 && (ch(first, 0) == '*' || ch(first, 0) == '?')) {
    lo = lo_b; hi = hi_b; extra = 1ULL << 63;  // whatever follows the first '-' is ignored, as upstream does
  } else {
    if (!parse_number_or_name(first, names, &lo)) return false;
    if (ndash == 0) hi = lo;
    else if (ndash == 1) {
      const Str second{r.p + dash + 1, r.n - dash - 1};
      if (!parse_number_or_name(second, names, &hi)) return false;
    } else return false;  // "too many hyphens"
  }
  if (nslash == 1) {
    const Str st{e.p + slash + 1, e.n - slash - 1};
    if (!parse_number(st, &step)) return false;
    if (ndash == 0) hi = hi_b;  // "N/step" means "N-max/step"
    if (step > 1) extra = 0;
  }
  if (lo < lo_b || hi > hi_b || lo > hi || step == 0) return false;
  uint64_t b = 0;
  for (uint32_t v = lo; v <= hi; v += step) b |= 1ULL << v;
  *bits = b | extra;
  return true;
}

// getField: comma list, empty pieces dropped
PK_HD inline bool parse_field(const Str& f, uint32_t lo_b, uint32_t hi_b, int names, uint64_t* bits) {
  uint64_t acc = 0;
  int st = 0;
  for (int i = 0; i <= f.n; ++i) {
    if (i == f.n || ch(f, i) == ',') {
      if (i > st) {
        uint64_t b;
        if (!parse_term(Str{f.p + st, i - st}, lo_b, hi_b, names, &b)) return false;
        acc |= b;
      }
      st = i + 1;
    }
  }
  *bits = acc;
  return true;
}

struct ParseOut {
  uint64_t* p[6];
};

// One spec: masks (zeroed on error), status, and the span of the zone name inside the text (relative to its start;
// offset 0xFFFFFFFF = "TZ= ...", the empty name, which LoadLocation reads as UTC).
PK_HD inline uint8_t parse_one(Str s, int with_seconds, uint64_t m[6], uint32_t* zo_out, uint32_t* zl_out) {
#pragma unroll
  for (int p = 0; p < 6; ++p) m[p] = 0;
  uint8_t st = kParseError;
  uint32_t zo = 0, zl = 0;
  const uint64_t kStar = 1ULL << 63;
  const uint64_t all_hour = 0xFFFFFFULL | kStar, all_dom = 0xFFFFFFFEULL | kStar, all_month = 0x1FFEULL | kStar,
                 all_dow = 0x7FULL | kStar;
  do {
    if (s.n == 0) break;  // "empty spec string"
    if (starts_with(s, "TZ=", 3) || starts_with(s, "CRON_TZ=", 8)) {
      int sp = -1, eq = -1;
      for (int k = 0; k < s.n; ++k) {
        const char c = ch(s, k);
        if (c == ' ' && sp < 0) sp = k;
        if (c == '=' && eq < 0) eq = k;
      }
      if (sp < 0) break;  // upstream panics here; a parse error in this library
      zo = (uint32_t)(eq + 1);
      zl = (uint32_t)(sp - eq - 1);
      s.p += sp; s.n -= sp;
      while (s.n && is_space(ch(s, 0))) { ++s.p; --s.n; }
      while (s.n && is_space(ch(s, s.n - 1))) --s.n;
      if (zl == 0) { zo = 0xFFFFFFFFu; }  // "TZ= ..." : LoadLocation("") is UTC; flagged with an impossible offset
    }
    if (s.n && ch(s, 0) == '@') {
      m[0] = 1; m[1] = 1;
      if (equals(s, "@yearly", 7) || equals(s, "@annually", 9)) { m[2] = 1; m[3] = 2; m[4] = 2; m[5] = all_dow; st = kParseOk; }
      else if (equals(s, "@monthly", 8)) { m[2] = 1; m[3] = 2; m[4] = all_month; m[5] = all_dow; st = kParseOk; }
      else if (equals(s, "@weekly", 7)) { m[2] = 1; m[3] = all_dom; m[4] = all_month; m[5] = 1; st = kParseOk; }
      else if (equals(s, "@daily", 6) || equals(s, "@midnight", 9)) { m[2] = 1; m[3] = all_dom; m[4] = all_month; m[5] = all_dow; st = kParseOk; }
      else if (equals(s, "@hourly", 7)) { m[2] = all_hour; m[3] = all_dom; m[4] = all_month; m[5] = all_dow; st = kParseOk; }
      else if (starts_with(s, "@every ", 7)) {
        // time.ParseDuration restricted to what needs no care: one or more <digits><h|m|s> terms, at most 9 digits
        // each (no overflow anywhere near Go's 2^63 ns).  Anything else -- sign, fraction, "ms"/"us"/"ns", a bare "0",
        // an unknown unit -- goes to the host parser (status 2), which is ParseDuration to the letter.
        m[0] = 0; m[1] = 0; st = kParseEvery;
        uint64_t total = 0;
        int k = 7;
        bool simple = k < s.n;
        while (simple && k < s.n) {
          uint64_t v = 0;
          int digits = 0;
          while (k < s.n && ch(s, k) >= '0' && ch(s, k) <= '9' && digits < 10) { v = v * 10 + (uint64_t)(ch(s, k) - '0'); ++k; ++digits; }
          if (digits == 0 || digits > 9 || k >= s.n) { simple = false; break; }
          const char u = ch(s, k++);
          const bool unit_ends = k >= s.n || (ch(s, k) >= '0' && ch(s, k) <= '9');  // the unit is this one letter
          if (!unit_ends || !(u == 'h' || u == 'm' || u == 's')) { simple = false; break; }
          total += v * (u == 'h' ? 3600ULL : u == 'm' ? 60ULL : 1ULL);
          if (total > 9000000000ULL) { simple = false; break; }  // near Go's 2^63 ns: the host parser decides
        }
        if (simple) {  // constantdelay.go Every(): at least one second
          m[3] = total < 1 ? 1 : total;
          m[4] = 1ULL << 62;  // B200CRON_EVERY_TAG
          st = kParseOk;
        }
      }
      else { m[0] = 0; m[1] = 0; }
      break;
    }
    // strings.Fields
    Str f[6];
    int nf = 0, k = 0;
    bool too_many = false;
    while (k < s.n) {
      while (k < s.n && is_space(ch(s, k))) ++k;
      if (k >= s.n) break;
      const int start = k;
      while (k < s.n && !is_space(ch(s, k))) ++k;
      if (nf < 6) f[nf] = Str{s.p + start, k - start};
      else too_many = true;
      ++nf;
    }
    const int want = with_seconds ? 6 : 5;
    if (too_many || nf != want) break;  // "expected exactly N fields"
    int q = 0;
    if (with_seconds) { if (!parse_field(f[q++], 0, 59, 0, &m[0])) break; }
    else m[0] = 1;  // normalizeFields: the seconds place defaults to "0"
    if (!parse_field(f[q++], 0, 59, 0, &m[1])) break;
    if (!parse_field(f[q++], 0, 23, 0, &m[2])) break;
    if (!parse_field(f[q++], 1, 31, 0, &m[3])) break;
    if (!parse_field(f[q++], 1, 12, 1, &m[4])) break;
    if (!parse_field(f[q++], 0, 6, 2, &m[5])) break;
    st = kParseOk;
  } while (false);
  if (st == kParseError) {
#pragma unroll
    for (int p = 0; p < 6; ++p) m[p] = 0;
  }
  *zo_out = zo;
  *zl_out = zl;
  return st;
}

#ifdef __CUDACC__
__global__ void __launch_bounds__(128)
parse_batch_kernel(const char* __restrict__ text, const uint64_t* __restrict__ text_off, size_t n, int with_seconds,
                   ParseOut out, uint8_t* __restrict__ status, uint32_t* __restrict__ tz_off, uint32_t* __restrict__ tz_len) {
  const size_t i = (size_t)blockIdx.x * 128 + threadIdx.x;
  if (i >= n) return;
  const uint64_t a = __ldg(text_off + i), b = __ldg(text_off + i + 1);
  uint64_t m[6];
  uint32_t zo, zl;
  const uint8_t st = parse_one(Str{text + a, (int)(b - a)}, with_seconds, m, &zo, &zl);
#pragma unroll
  for (int p = 0; p < 6; ++p) out.p[p][i] = m[p];
  status[i] = st;
  tz_off[i] = zo;
  tz_len[i] = zl;
}
#endif

}  // namespace b200cron
