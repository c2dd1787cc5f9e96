// b200cron.hpp -- C++ host mirror of the reference's schedule interface over the C ABI of include/b200cron.h.
//
// The reference is Go; its seam for this path is
//   cronv3.ParseStandard(string) (cron.Schedule, error)          internal/controller/cron_controller.go:392
//   cron.Schedule { Next(time.Time) time.Time }                   :405, :409, :436
//   (*CronReconciler).getNextSchedule(ctx, cron, now)             :389-437
//   sortByCreationTimestamp / syncCronHistory                     internal/controller/cron_util.go:117-129, :307-346
// This header keeps those names, argument meanings and error behaviour (errors become exceptions whose text carries
// the reference's "unparsable cron" / "unschedulable cron" substrings) and adds the batched calls the library exists
// for.  Header-only; link with -lb200cron.  Times are Unix seconds; kZeroTime is Go's time.Time{}.
// There is no CPU fallback: every call that computes goes to the CUDA library, and Engine's constructor throws when
// no GPU is usable.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "b200cron.h"

namespace b200cron {

constexpr int64_t kZeroTime = INT64_MIN;  // B200CRON_ZERO_TIME

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& what) : std::runtime_error(what), code(c) {}
};

// robfig/cron v3 SpecSchedule (spec.go) or ConstantDelaySchedule (constantdelay.go, every > 0)
struct Schedule {
  b200cron_spec spec{};
  std::string location;  // "" == time.Local
  int64_t every = 0;     // seconds; > 0 for "@every <duration>"
};

// cronv3.ParseStandard.  Throws Error(B200CRON_EPARSE, upstream's message) -- callers wrap it the way the reference
// does: "unparsable cron %q: %w" (cron_controller.go:394).  Host code, needs no GPU.
inline Schedule ParseStandard(const std::string& text) {
  Schedule s;
  int kind = 0;
  char tz[128] = {0}, err[512] = {0};
  const int rc = b200cron_parse_standard(text.data(), text.size(), &s.spec, &kind, &s.every, tz, sizeof tz, err, sizeof err);
  if (rc != B200CRON_OK) throw Error(rc, err);
  if (kind != B200CRON_KIND_EVERY) s.every = 0;
  s.location = tz;
  return s;
}

// getNextSchedule's result plus Reconcile's decision for one Cron
struct Unmet {
  int64_t last = kZeroTime;  // lastMissed
  int64_t next = kZeroTime;  // Next(now)
  uint32_t missed = 0;
  uint8_t status = 0;    // B200CRON_ST_*
  uint8_t decision = 0;  // B200CRON_DEC_*
};

class Engine {
 public:
  // devices empty = the current device; local_zone is the process's time.Local (zone of specs without TZ=)
  explicit Engine(const std::vector<int>& devices = {}, const char* zoneinfo_dir = nullptr,
                  const std::string& local_zone = "UTC") {
    const int rc = b200cron_ctx_create(devices.empty() ? nullptr : devices.data(), (int)devices.size(), zoneinfo_dir, 0, &ctx_);
    if (rc != B200CRON_OK) throw Error(rc, b200cron_last_error(nullptr));
    local_ = ZoneId(local_zone);
  }
  ~Engine() { b200cron_ctx_destroy(ctx_); }
  Engine(const Engine&) = delete;
  Engine& operator=(const Engine&) = delete;

  // time.LoadLocation(name) as a zone id; throws like ParseStandard's "provided bad location" error
  int32_t ZoneId(const std::string& name) const {
    const int z = b200cron_zone_id(ctx_, name.c_str());
    if (z < 0) throw Error(B200CRON_EPARSE, "provided bad location " + name + ": unknown time zone " + name);
    return z;
  }
  int32_t LocationId(const Schedule& s) const { return (s.location.empty() || s.location == "Local") ? local_ : ZoneId(s.location); }

  // Schedule.Next(t) for one schedule (a batch of one: launch latency dominates -- use NextN for work)
  int64_t Next(const Schedule& s, int64_t t) {
    return NextN({s}, {t}, 1)[0];  // either kind: an "@every" schedule rides in the planes (B200CRON_EVERY_TAG)
  }

  // out[i*k + j] = the (j+1)-th fire time after from[i]
  std::vector<int64_t> NextN(const std::vector<Schedule>& specs, const std::vector<int64_t>& from, int k) {
    Planes p(*this, specs);
    std::vector<int64_t> out(specs.size() * (size_t)k);
    Check(b200cron_next_n(ctx_, p.ptr, p.zone.data(), from.data(), specs.size(), k, out.data()));
    return out;
  }

  // getNextSchedule + gates + Allow/Forbid/Replace for a batch.  policy/active/flags/deadline may be empty
  // (Allow, 0, 0, nil).  now_nsec: the sub-second part of time.Now() (only the deadline gate can see it).
  std::vector<Unmet> UnmetSince(const std::vector<Schedule>& specs, const std::vector<int64_t>& earliest, int64_t now,
                                const std::vector<uint8_t>& policy = {}, const std::vector<uint32_t>& active = {},
                                const std::vector<uint8_t>& flags = {}, const std::vector<int64_t>& deadline = {},
                                int32_t now_nsec = 0) {
    const size_t n = specs.size();
    Planes p(*this, specs);
    std::vector<int64_t> last(n), next(n);
    std::vector<uint32_t> missed(n);
    std::vector<uint8_t> status(n), decision(n);
    Check(b200cron_unmet_since(ctx_, p.ptr, p.zone.data(), earliest.data(), now, now_nsec,
                               policy.empty() ? nullptr : policy.data(), active.empty() ? nullptr : active.data(),
                               flags.empty() ? nullptr : flags.data(), deadline.empty() ? nullptr : deadline.data(), n,
                               last.data(), next.data(), missed.data(), status.data(), decision.data()));
    std::vector<Unmet> out(n);
    for (size_t i = 0; i < n; ++i) out[i] = Unmet{last[i], next[i], missed[i], status[i], decision[i]};
    return out;
  }

  // getNextSchedule for one Cron: {lastMissed, next}; throws the reference's errors (:394, :411)
  std::pair<int64_t, int64_t> getNextSchedule(const std::string& schedule, int64_t earliest, int64_t now) {
    Schedule s;
    try {
      s = ParseStandard(schedule);
    } catch (const Error& e) {
      throw Error(e.code, "unparsable cron \"" + schedule + "\": " + e.what());
    }
    const Unmet u = UnmetSince({s}, {earliest}, now)[0];
    if (u.status == B200CRON_ST_UNSCHEDULABLE || u.status == B200CRON_ST_NONTERMINATING)
      throw Error(B200CRON_ERANGE, "unschedulable cron \"" + schedule + "\"");
    return {u.last, u.next};
  }

  // sortByCreationTimestamp + syncCronHistory's prune over segments: records [seg_off[i], seg_off[i+1]) belong to
  // Cron i; perm = stable ascending order inside each segment (indices relative to the segment), keep[j] = 0 for the
  // sorted positions the reference deletes (limit < 0: HistoryLimit nil)
  void HistoryPrune(const std::vector<uint64_t>& seg_off, const std::vector<int64_t>& created,
                    const std::vector<int64_t>& limit, std::vector<uint32_t>* perm, std::vector<uint8_t>* keep) {
    perm->resize(created.size());
    keep->resize(created.size());
    if (seg_off.size() < 2) return;
    Check(b200cron_history_prune(ctx_, seg_off.data(), created.data(), limit.empty() ? nullptr : limit.data(),
                                 seg_off.size() - 1, perm->data(), keep->data()));
  }

  b200cron_ctx* handle() const { return ctx_; }

  // A resident set of schedules (b200cron_specset_*): a SpecSchedule is immutable once ParseStandard has built it
  // (cron_controller.go:392), so a host that ticks the same Crons again and again parks their masks on the GPU once
  // and then moves only the per-tick arrays.  Destroy it before the Engine.
  class SpecSet {
   public:
    SpecSet(Engine& e, const std::vector<Schedule>& specs) : e_(e), n_(specs.size()) {
      Planes p(e, specs);
      e.Check(b200cron_specset_create(e.ctx_, p.ptr, p.zone.data(), n_, &set_));
    }
    ~SpecSet() { b200cron_specset_destroy(set_); }
    SpecSet(const SpecSet&) = delete;
    SpecSet& operator=(const SpecSet&) = delete;
    size_t size() const { return n_; }
    // the Cron at `index` got a new schedule (its CR was edited)
    void Update(size_t index, const Schedule& s) {
      Planes p(e_, {s});
      e_.Check(b200cron_specset_update(set_, index, 1, p.ptr, p.zone.data()));
    }
    std::vector<int64_t> NextN(const std::vector<int64_t>& from, int k) {
      std::vector<int64_t> out(n_ * (size_t)k);
      e_.Check(b200cron_next_n_set(e_.ctx_, set_, from.data(), k, out.data()));
      return out;
    }
    std::vector<Unmet> UnmetSince(const std::vector<int64_t>& earliest, int64_t now, const std::vector<uint8_t>& policy = {},
                                  const std::vector<uint32_t>& active = {}, const std::vector<uint8_t>& flags = {},
                                  const std::vector<int64_t>& deadline = {}, int32_t now_nsec = 0) {
      std::vector<int64_t> last(n_), next(n_);
      std::vector<uint32_t> missed(n_);
      std::vector<uint8_t> status(n_), decision(n_);
      e_.Check(b200cron_unmet_since_set(e_.ctx_, set_, earliest.data(), now, now_nsec,
                                        policy.empty() ? nullptr : policy.data(), active.empty() ? nullptr : active.data(),
                                        flags.empty() ? nullptr : flags.data(), deadline.empty() ? nullptr : deadline.data(),
                                        last.data(), next.data(), missed.data(), status.data(), decision.data()));
      std::vector<Unmet> out(n_);
      for (size_t i = 0; i < n_; ++i) out[i] = Unmet{last[i], next[i], missed[i], status[i], decision[i]};
      return out;
    }

   private:
    Engine& e_;
    size_t n_;
    b200cron_specset* set_ = nullptr;
  };

 private:
  struct Planes {  // SoA planes + zone ids of a batch of schedules
    std::vector<uint64_t> m[6];
    std::vector<int32_t> zone;
    const uint64_t* ptr[6];
    Planes(const Engine& e, const std::vector<Schedule>& specs) {
      for (auto& v : m) v.reserve(specs.size());
      zone.reserve(specs.size());
      for (const Schedule& s : specs) {
        const uint64_t f[6] = {s.spec.second, s.spec.minute, s.spec.hour, s.spec.dom, s.spec.month, s.spec.dow};
        for (int p = 0; p < 6; ++p) m[p].push_back(f[p]);
        zone.push_back(e.LocationId(s));
      }
      for (int p = 0; p < 6; ++p) ptr[p] = m[p].data();
    }
  };
  void Check(int rc) const {
    if (rc != B200CRON_OK) throw Error(rc, b200cron_last_error(ctx_));
  }
  b200cron_ctx* ctx_ = nullptr;
  int32_t local_ = 0;
};

}  // namespace b200cron
