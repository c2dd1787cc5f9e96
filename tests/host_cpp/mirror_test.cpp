// mirror_test.cpp -- exercises include/b200cron.hpp (the C++ host mirror of the reference's schedule interface).
//   mirror_test cpu   host-only checks: parser, error texts, and that an Engine cannot exist without a GPU
//   mirror_test gpu   the reference's K1-K3 cases, a batch NextN / UnmetSince and a history prune through the library;
//                     prints the results as "key v v v" lines for tests/test_host_cpp.py to compare with the oracle
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <string>

#include "b200cron.hpp"

using namespace b200cron;

static const int64_t kNow = 1767268800;  // 2026-01-01T12:00:00Z

static void print_spec(const char* key, const Schedule& s) {
  std::printf("%s %" PRIu64 " %" PRIu64 " %" PRIu64 " %" PRIu64 " %" PRIu64 " %" PRIu64 " %s %" PRId64 "\n", key, s.spec.second,
              s.spec.minute, s.spec.hour, s.spec.dom, s.spec.month, s.spec.dow, s.location.empty() ? "-" : s.location.c_str(), s.every);
}

static int cpu() {
  print_spec("spec5", ParseStandard("*/5 * * * *"));
  print_spec("spectz", ParseStandard("CRON_TZ=Asia/Tokyo 0 9 * * 1-5"));
  print_spec("every", ParseStandard("@every 1h30m"));
  try {
    ParseStandard("not a schedule");
    std::printf("k1 no-error\n");
  } catch (const Error& e) {
    std::printf("k1 code %d text %s\n", e.code, e.what());
  }
  try {
    Engine e;
    std::printf("engine created\n");
  } catch (const Error& e) {
    std::printf("engine-error %s\n", e.what());
  }
  return 0;
}

static int gpu() {
  Engine e;
  try {
    e.getNextSchedule("not a schedule", kNow - 300, kNow);
    std::printf("k1 no-error\n");
  } catch (const Error& x) {
    std::printf("k1 %s\n", x.what());
  }
  try {
    e.getNextSchedule("0 0 30 2 *", kNow - 300, kNow);
    std::printf("k2 no-error\n");
  } catch (const Error& x) {
    std::printf("k2 %s\n", x.what());
  }
  const auto k3 = e.getNextSchedule("*/1 * * * *", kNow - 300, kNow);
  std::printf("k3 %" PRId64 " %" PRId64 "\n", k3.first, k3.second);
  const auto ev = e.getNextSchedule("@every 90s", kNow - 200, kNow);
  std::printf("every %" PRId64 " %" PRId64 "\n", ev.first, ev.second);
  const std::vector<Schedule> specs = {ParseStandard("*/5 * * * *"), ParseStandard("TZ=America/New_York 30 2 * * *"),
                                       ParseStandard("0 0 29 2 *"), ParseStandard("15 3 * * 0")};
  std::printf("next %" PRId64 "\n", e.Next(specs[0], kNow));
  const auto nn = e.NextN(specs, {kNow, 1772935200 /* 2026-03-08T02:00Z, US spring forward */, kNow, kNow}, 3);
  std::printf("nextn");
  for (int64_t v : nn) std::printf(" %" PRId64, v);
  std::printf("\n");
  const auto um = e.UnmetSince(specs, {kNow - 3600, kNow - 5 * 86400, kNow - 86400, kNow + 10}, kNow, {0, 1, 2, 0}, {0, 2, 1, 0},
                               {0, 0, 0, 0}, {kZeroTime, kZeroTime, kZeroTime, kZeroTime});
  std::printf("unmet");
  for (const Unmet& u : um) std::printf(" %" PRId64 " %" PRId64 " %u %u %u", u.last, u.next, u.missed, (unsigned)u.status, (unsigned)u.decision);
  std::printf("\n");
  {  // the same two calls through a resident set, before and after one Cron's schedule is edited
    Engine::SpecSet set(e, specs);
    const auto sn = set.NextN({kNow, 1772935200, kNow, kNow}, 3);
    const auto su = set.UnmetSince({kNow - 3600, kNow - 5 * 86400, kNow - 86400, kNow + 10}, kNow, {0, 1, 2, 0}, {0, 2, 1, 0});
    bool same = sn == nn && su.size() == um.size();
    for (size_t i = 0; same && i < su.size(); ++i)
      same = su[i].last == um[i].last && su[i].next == um[i].next && su[i].missed == um[i].missed &&
             su[i].status == um[i].status && su[i].decision == um[i].decision;
    set.Update(2, specs[0]);
    const auto sn2 = set.NextN({kNow, 1772935200, kNow, kNow}, 3);
    same = same && sn2[6] == nn[0] && sn2[7] == nn[1] && sn2[8] == nn[2] && sn2[0] == nn[0] && sn2[9] == nn[9];
    std::printf("specset %s\n", same ? "same" : "DIFFERENT");
  }
  std::vector<uint32_t> perm;
  std::vector<uint8_t> keep;
  e.HistoryPrune({0, 5, 5, 8}, {5, 3, 9, 3, 7, 2, 2, 1}, {3, 1, -1}, &perm, &keep);
  std::printf("history");
  for (size_t i = 0; i < perm.size(); ++i) std::printf(" %u:%u", perm[i], (unsigned)keep[i]);
  std::printf("\n");
  return 0;
}

int main(int argc, char** argv) {
  try {
    return (argc > 1 && !std::strcmp(argv[1], "gpu")) ? gpu() : cpu();
  } catch (const std::exception& e) {
    std::printf("FAILED %s\n", e.what());
    return 1;
  }
}
