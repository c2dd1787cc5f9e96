// zone_table.h -- host-side builder of the flattened IANA zone table the kernels read (pure C++17, no CUDA).
//
// Replaces, for the batched path, what the reference gets from Go's time.LoadLocation when a schedule carries a
// TZ=/CRON_TZ= prefix (robfig/cron v3.0.1 parser.go, reached from internal/controller/cron_controller.go:392) and
// from time.Local otherwise.  Each zone becomes a sorted list of (when, utc offset) periods: the TZif 64-bit body,
// Go's "first zone" rule before the first transition, and the footer POSIX-TZ rule expanded year by year up to a
// horizon, so that the device never parses a rule string.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace b200cron {

struct HostZoneTable {
  std::vector<std::string> names;  // names[0] == "UTC"; the rest sorted
  std::vector<int32_t> zstart;     // names.size() + 1 entries
  std::vector<int64_t> when;       // first entry of every zone is INT64_MIN
  std::vector<int32_t> off;
  // fpre[i] (zone-relative prefix): number of transitions among this zone's entries 1..i whose gap or overlap
  // contains a local midnight -- the only transitions at which Date(y, m, d, 0, 0, 0) is not "natural"
  std::vector<int32_t> fpre;
  // Year index: row z holds, for every 365-day bucket b since 1900-01-01 (kYearBase, kYearStep), the index (relative
  // to the zone) of the period in force at the bucket's start.  A lookup reads two neighbouring entries and is left
  // with the two or three periods of that bucket instead of a binary search over the whole zone (up to ~700 periods
  // with footer rules expanded to 2200) -- on the device each step of that search is a dependent L2 access.
  std::vector<uint16_t> yidx;      // names.size() rows of ystride entries
  int32_t ystride = 0;             // buckets per row + 1
  int64_t horizon = 0;             // unix second up to which footer rules were expanded
  int horizon_year = 0;

  int find(const std::string& name) const;  // -1 when unknown
};

// Scans `zoneinfo_dir` recursively for TZif files (skipping posix/ and right/), builds the table.
// Returns false and fills `err` when the directory is unusable.
bool build_zone_table(const std::string& zoneinfo_dir, int horizon_year, HostZoneTable* out, std::string* err);

}  // namespace b200cron
