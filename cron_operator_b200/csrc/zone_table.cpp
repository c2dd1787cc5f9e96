// zone_table.cpp -- see zone_table.h.  Follows Go's zoneinfo_read.go / zoneinfo.go semantics (go 1.25.5,
// reference go.mod:3) because parity is defined against the reference's Go path; [UPSTREAM, not under /root/reference].
#include "zone_table.h"

#include <dirent.h>
#include <sys/stat.h>

#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstring>
#include <optional>
#include <string_view>

namespace b200cron {
namespace {

constexpr int64_t kAlpha = INT64_MIN;
constexpr int kDay = 86400;

int64_t floordiv(int64_t a, int64_t b) { return a / b - ((a % b) < 0); }
bool leap(int64_t y) { return y % 4 == 0 && (y % 100 != 0 || y % 400 == 0); }
int64_t days_from_civil(int64_t y, int m, int d) {
  y -= m <= 2;
  const int64_t era = floordiv(y, 400);
  const int64_t yoe = y - era * 400;
  const int64_t doy = (153 * (m > 2 ? m - 3 : m + 9) + 2) / 5 + d - 1;
  return era * 146097 + yoe * 365 + yoe / 4 - yoe / 100 + doy - 719468;
}
int64_t year_of_unix(int64_t sec) {
  int64_t z = floordiv(sec, kDay) + 719468;
  const int64_t era = floordiv(z, 146097);
  const int64_t doe = z - era * 146097;
  const int64_t yoe = (doe - doe / 1460 + doe / 36524 - doe / 146096) / 365;
  const int64_t doy = doe - (365 * yoe + yoe / 4 - yoe / 100);
  const int64_t mp = (5 * doy + 2) / 153;
  const int64_t m = mp < 10 ? mp + 3 : mp - 9;
  return yoe + era * 400 + (m <= 2);
}

// ---------------------------------------------------------------------------------------------------------
// POSIX-TZ footer, parsed the way Go's tzset() accepts it
// ---------------------------------------------------------------------------------------------------------
struct Rule {
  enum Kind { Julian, DayOfYear, MonthWeekDay } kind = Julian;
  int day = 0, week = 0, mon = 0, time = 2 * 3600;
};
struct Footer {
  int std_off = 0;  // seconds east of UTC
  int dst_off = 0;
  bool has_dst = false;
  Rule start, end;
};

class Cursor {
 public:
  explicit Cursor(std::string_view s) : s_(s) {}
  bool empty() const { return s_.empty(); }
  char peek() const { return s_.empty() ? '\0' : s_[0]; }
  void skip(size_t n = 1) { s_.remove_prefix(n); }

  bool name() {
    if (s_.empty()) return false;
    if (s_[0] == '<') {
      const size_t p = s_.find('>');
      if (p == std::string_view::npos) return false;
      s_.remove_prefix(p + 1);
      return true;
    }
    for (size_t i = 0; i < s_.size(); ++i) {
      const char c = s_[i];
      if ((c >= '0' && c <= '9') || c == ',' || c == '-' || c == '+') {
        if (i < 3) return false;
        s_.remove_prefix(i);
        return true;
      }
    }
    if (s_.size() < 3) return false;
    s_ = {};
    return true;
  }
  std::optional<int> num(int lo, int hi) {
    if (s_.empty()) return std::nullopt;
    int v = 0;
    size_t i = 0;
    for (; i < s_.size(); ++i) {
      const char c = s_[i];
      if (c < '0' || c > '9') break;
      v = v * 10 + (c - '0');
      if (v > hi) return std::nullopt;
    }
    if (i == 0 || v < lo) return std::nullopt;
    s_.remove_prefix(i);
    return v;
  }
  std::optional<int> offset() {
    if (s_.empty()) return std::nullopt;
    bool neg = false;
    if (peek() == '+') skip();
    else if (peek() == '-') { skip(); neg = true; }
    auto h = num(0, 24 * 7);
    if (!h) return std::nullopt;
    int o = *h * 3600;
    if (peek() == ':') {
      skip();
      auto m = num(0, 59);
      if (!m) return std::nullopt;
      o += *m * 60;
      if (peek() == ':') {
        skip();
        auto sec = num(0, 59);
        if (!sec) return std::nullopt;
        o += *sec;
      }
    }
    return neg ? -o : o;
  }
  std::optional<Rule> rule() {
    Rule r;
    if (s_.empty()) return std::nullopt;
    if (peek() == 'J') {
      skip();
      auto d = num(1, 365);
      if (!d) return std::nullopt;
      r.kind = Rule::Julian;
      r.day = *d;
    } else if (peek() == 'M') {
      skip();
      auto mon = num(1, 12);
      if (!mon || peek() != '.') return std::nullopt;
      skip();
      auto week = num(1, 5);
      if (!week || peek() != '.') return std::nullopt;
      skip();
      auto day = num(0, 6);
      if (!day) return std::nullopt;
      r.kind = Rule::MonthWeekDay;
      r.mon = *mon; r.week = *week; r.day = *day;
    } else {
      auto d = num(0, 365);
      if (!d) return std::nullopt;
      r.kind = Rule::DayOfYear;
      r.day = *d;
    }
    if (peek() == '/') {
      skip();
      auto o = offset();
      if (!o) return std::nullopt;
      r.time = *o;
    }
    return r;
  }

 private:
  std::string_view s_;
};

std::optional<Footer> parse_footer(std::string_view text) {
  Footer f;
  Cursor c(text);
  if (!c.name()) return std::nullopt;
  auto so = c.offset();
  if (!so) return std::nullopt;
  f.std_off = -*so;  // POSIX counts west-positive
  if (c.empty() || c.peek() == ',') return f;
  if (!c.name()) return std::nullopt;
  if (c.empty() || c.peek() == ',') {
    f.dst_off = f.std_off + 3600;
  } else {
    auto d = c.offset();
    if (!d) return std::nullopt;
    f.dst_off = -*d;
  }
  std::string rest_default = ",M3.2.0,M11.1.0";
  Cursor r = c.empty() ? Cursor(rest_default) : c;
  if (r.peek() != ',' && r.peek() != ';') return std::nullopt;
  r.skip();
  auto a = r.rule();
  if (!a || r.peek() != ',') return std::nullopt;
  r.skip();
  auto b = r.rule();
  if (!b || !r.empty()) return std::nullopt;
  f.has_dst = true;
  f.start = *a;
  f.end = *b;
  return f;
}

// seconds from the start of `year` (UTC) at which the rule fires, given the offset in force before it
int rule_second(int64_t year, const Rule& r, int off_before) {
  static const int before[12] = {0, 31, 59, 90, 120, 151, 181, 212, 243, 273, 304, 334};
  static const int mdays[12] = {31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31};
  int s = 0;
  switch (r.kind) {
    case Rule::Julian:
      s = (r.day - 1) * kDay;
      if (leap(year) && r.day >= 60) s += kDay;
      break;
    case Rule::DayOfYear:
      s = r.day * kDay;
      break;
    case Rule::MonthWeekDay: {
      const int m1 = (r.mon + 9) % 12 + 1;
      int64_t yy0 = year - (r.mon <= 2);
      const int64_t yy1 = yy0 / 100, yy2 = yy0 % 100;
      int64_t dow = ((26 * m1 - 2) / 10 + 1 + yy2 + yy2 / 4 + yy1 / 4 - 2 * yy1) % 7;
      if (dow < 0) dow += 7;
      int d = r.day - (int)dow;
      if (d < 0) d += 7;
      const int dim = mdays[r.mon - 1] + (r.mon == 2 && leap(year));
      for (int i = 1; i < r.week; ++i) {
        if (d + 7 >= dim) break;
        d += 7;
      }
      d += before[r.mon - 1];
      if (leap(year) && r.mon > 2) ++d;
      s = d * kDay;
      break;
    }
  }
  return s + r.time - off_before;
}

// offset Go's tzset() reports for unix second `sec`
int footer_offset(const Footer& f, int64_t sec) {
  if (!f.has_dst) return f.std_off;
  const int64_t year = year_of_unix(sec);
  const int64_t ystart = days_from_civil(year, 1, 1) * (int64_t)kDay;
  const int64_t ysec = sec - ystart;
  int64_t a = rule_second(year, f.start, f.std_off);
  int64_t b = rule_second(year, f.end, f.dst_off);
  int outer = f.std_off, inner = f.dst_off;
  if (b < a) { std::swap(a, b); std::swap(outer, inner); }
  return (ysec < a || ysec >= b) ? outer : inner;
}

// ---------------------------------------------------------------------------------------------------------
// TZif
// ---------------------------------------------------------------------------------------------------------
struct RawZone {
  struct Type { int32_t off; bool dst; };
  std::vector<Type> types;
  std::vector<int64_t> when;
  std::vector<uint8_t> idx;
  std::string footer;
};

uint32_t be32(const uint8_t* p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }
int64_t be64(const uint8_t* p) { return (int64_t)((uint64_t)be32(p) << 32 | be32(p + 4)); }

bool parse_tzif(const std::vector<uint8_t>& d, RawZone* z) {
  size_t pos = 0;
  auto need = [&](size_t n) { return pos + n <= d.size(); };
  if (!need(20) || std::memcmp(d.data(), "TZif", 4) != 0) return false;
  const uint8_t ver = d[4];
  if (ver != 0 && ver != '2' && ver != '3') return false;  // Go rejects other versions
  pos = 20;
  size_t cnt[6];  // isutc, isstd, leap, time, type, char
  auto read_counts = [&]() {
    if (!need(24)) return false;
    for (int i = 0; i < 6; ++i) { cnt[i] = be32(d.data() + pos); pos += 4; }
    return true;
  };
  if (!read_counts()) return false;
  size_t tsz = 4;
  if (ver != 0) {
    const size_t skip = cnt[3] * 4 + cnt[3] + cnt[4] * 6 + cnt[5] + cnt[2] * 8 + cnt[1] + cnt[0] + 20;
    if (!need(skip)) return false;
    pos += skip;
    if (!read_counts()) return false;
    tsz = 8;
  }
  const size_t ntime = cnt[3], ntype = cnt[4], nchar = cnt[5];
  if (!need(ntime * tsz + ntime + ntype * 6 + nchar + cnt[2] * (tsz + 4) + cnt[1] + cnt[0])) return false;
  if (ntype == 0) return false;
  const uint8_t* times = d.data() + pos; pos += ntime * tsz;
  const uint8_t* tidx = d.data() + pos; pos += ntime;
  const uint8_t* types = d.data() + pos; pos += ntype * 6;
  pos += nchar + cnt[2] * (tsz + 4) + cnt[1] + cnt[0];
  z->types.resize(ntype);
  for (size_t i = 0; i < ntype; ++i) {
    z->types[i].off = (int32_t)be32(types + 6 * i);
    z->types[i].dst = types[6 * i + 4] != 0;
    if (types[6 * i + 5] >= nchar) return false;
  }
  z->when.resize(ntime);
  z->idx.resize(ntime);
  for (size_t i = 0; i < ntime; ++i) {
    z->when[i] = tsz == 8 ? be64(times + 8 * i) : (int64_t)(int32_t)be32(times + 4 * i);
    if (tidx[i] >= ntype) return false;
    z->idx[i] = tidx[i];
  }
  const size_t rest = d.size() - pos;
  if (rest > 2 && d[pos] == '\n' && d.back() == '\n') z->footer.assign((const char*)d.data() + pos + 1, rest - 2);
  return true;
}

// Go's lookupFirstZone(): the type in force before the first transition
int first_type(const RawZone& z) {
  bool used0 = false;
  for (uint8_t i : z.idx) used0 |= (i == 0);
  if (!used0) return 0;
  if (!z.idx.empty() && z.types[z.idx[0]].dst) {
    for (int i = (int)z.idx[0] - 1; i >= 0; --i)
      if (!z.types[i].dst) return i;
  }
  for (size_t i = 0; i < z.types.size(); ++i)
    if (!z.types[i].dst) return (int)i;
  return 0;
}

void flatten(const RawZone& z, int horizon_year, std::vector<int64_t>* when, std::vector<int32_t>* off) {
  std::vector<std::pair<int64_t, int32_t>> e;
  if (z.when.empty()) {
    e.emplace_back(kAlpha, z.types[0].off);  // Go's fake transition {alpha, index 0}
  } else {
    e.emplace_back(kAlpha, z.types[first_type(z)].off);
    for (size_t i = 0; i < z.when.size(); ++i) e.emplace_back(z.when[i], z.types[z.idx[i]].off);
  }
  // From the last explicit transition on, Go answers from the footer rule (when it parses).
  if (!z.footer.empty()) {
    if (auto f = parse_footer(z.footer)) {
      const int64_t last = e.back().first;
      if (!f->has_dst) {
        e.back().second = f->std_off;
      } else {
        const int64_t horizon = days_from_civil(horizon_year, 1, 1) * (int64_t)kDay;
        const int64_t from = last == kAlpha ? days_from_civil(1800, 1, 1) * (int64_t)kDay : last;
        e.back().second = footer_offset(*f, last == kAlpha ? from : last);
        for (int64_t y = year_of_unix(from); y < horizon_year; ++y) {
          const int64_t ystart = days_from_civil(y, 1, 1) * (int64_t)kDay;
          int64_t bp[3] = {ystart, ystart + rule_second(y, f->start, f->std_off),
                           ystart + rule_second(y, f->end, f->dst_off)};
          std::sort(bp, bp + 3);
          const int64_t yend = days_from_civil(y + 1, 1, 1) * (int64_t)kDay;
          for (int64_t b : bp) {
            if (b <= from || b < ystart || b >= yend || b >= horizon) continue;
            e.emplace_back(b, footer_offset(*f, b));
          }
        }
      }
    }
  }
  // merge periods that do not change the offset (Date()'s result only depends on the offset function)
  for (size_t i = 0; i < e.size(); ++i) {
    if (!when->empty() && i > 0 && off->back() == e[i].second) continue;
    if (i > 0 && e[i].first <= when->back()) {  // out-of-order or duplicate instants: last writer wins
      off->back() = e[i].second;
      continue;
    }
    when->push_back(e[i].first);
    off->push_back(e[i].second);
  }
}

bool read_file(const std::string& path, std::vector<uint8_t>* out) {
  FILE* f = std::fopen(path.c_str(), "rb");
  if (!f) return false;
  uint8_t buf[1 << 16];
  size_t n;
  out->clear();
  while ((n = std::fread(buf, 1, sizeof buf, f)) > 0) out->insert(out->end(), buf, buf + n);
  std::fclose(f);
  return true;
}

void scan(const std::string& root, const std::string& rel, std::vector<std::string>* names) {
  const std::string dir = rel.empty() ? root : root + "/" + rel;
  DIR* dp = opendir(dir.c_str());
  if (!dp) return;
  while (dirent* de = readdir(dp)) {
    const std::string n = de->d_name;
    if (n == "." || n == "..") continue;
    const std::string r = rel.empty() ? n : rel + "/" + n;
    if (rel.empty() && (n == "posix" || n == "right")) continue;
    struct stat st;
    if (stat((root + "/" + r).c_str(), &st) != 0) continue;
    if (S_ISDIR(st.st_mode)) scan(root, r, names);
    else if (S_ISREG(st.st_mode)) names->push_back(r);
  }
  closedir(dp);
}

}  // namespace

int HostZoneTable::find(const std::string& name) const {
  if (name.empty() || name == "UTC") return 0;  // time.LoadLocation("") / ("UTC")
  auto it = std::lower_bound(names.begin() + 1, names.end(), name);
  if (it == names.end() || *it != name) return -1;
  return (int)(it - names.begin());
}

bool build_zone_table(const std::string& zoneinfo_dir, int horizon_year, HostZoneTable* out, std::string* err) {
  std::vector<std::string> files;
  scan(zoneinfo_dir, "", &files);
  std::sort(files.begin(), files.end());
  *out = HostZoneTable();
  out->horizon_year = horizon_year;
  out->horizon = days_from_civil(horizon_year, 1, 1) * (int64_t)kDay;
  out->names.push_back("UTC");
  out->zstart.push_back(0);
  out->when.push_back(kAlpha);
  out->off.push_back(0);
  out->fpre.push_back(0);
  std::vector<uint8_t> data;
  for (const std::string& rel : files) {
    if (rel == "UTC") continue;  // id 0
    RawZone z;
    if (!read_file(zoneinfo_dir + "/" + rel, &data) || !parse_tzif(data, &z)) continue;  // not a TZif file
    out->names.push_back(rel);
    out->zstart.push_back((int32_t)out->when.size());
    std::vector<int64_t> w;
    std::vector<int32_t> o;
    flatten(z, horizon_year, &w, &o);
    out->when.insert(out->when.end(), w.begin(), w.end());
    out->off.insert(out->off.end(), o.begin(), o.end());
    // A transition at T from offset A to B makes the local times [T+min(A,B), T+max(A,B)] missing (gap) or double
    // (overlap); flag it when a local midnight falls in that closed interval.
    int32_t run = 0;
    for (size_t i = 0; i < w.size(); ++i) {
      if (i > 0) {
        const int64_t lo = w[i] + std::min(o[i - 1], o[i]), hi = w[i] + std::max(o[i - 1], o[i]);
        const int64_t first_mid = floordiv(lo + kDay - 1, kDay) * kDay;  // first multiple of a day >= lo
        if (first_mid <= hi) ++run;
      }
      out->fpre.push_back(run);
    }
  }
  out->zstart.push_back((int32_t)out->when.size());
  // year index (see zone_table.h); kYearBase / kYearStep are shared with the walk (cron_walk.cuh)
  const int64_t kYearBase = -2208988800LL, kYearStep = 31536000LL;  // 1900-01-01T00:00:00Z, 365 days
  const int32_t nb = (int32_t)((out->horizon - kYearBase + kYearStep - 1) / kYearStep);
  out->ystride = nb + 1;
  out->yidx.assign(out->names.size() * (size_t)out->ystride, 0);
  for (size_t z = 0; z < out->names.size(); ++z) {
    const int32_t a = out->zstart[z], zn = out->zstart[z + 1] - a;
    int32_t i = 0;
    for (int32_t b = 0; b <= nb; ++b) {
      const int64_t t = kYearBase + (int64_t)b * kYearStep;
      while (i + 1 < zn && out->when[a + i + 1] <= t) ++i;
      out->yidx[z * (size_t)out->ystride + b] = (uint16_t)i;
    }
    if (zn > 65535) {
      if (err) *err = "zone " + out->names[z] + " has too many periods for the year index";
      return false;
    }
  }
  if (out->names.size() <= 1) {
    if (err) *err = "no TZif files found under " + zoneinfo_dir;
    return false;
  }
  return true;
}

}  // namespace b200cron
