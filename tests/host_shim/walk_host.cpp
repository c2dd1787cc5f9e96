// walk_host.cpp -- TEST-ONLY host build of the kernels' per-spec walk (cron_walk.cuh) and zone table builder.
//
// This box has no GPU, so the CPU test suite compiles the *same* __host__ __device__ source the kernels use with g++
// and checks it against the oracle.  It is test infrastructure: nothing in the product library links or calls it,
// and the product has no CPU execution path.
#include <cstdint>
#include <cstring>
#include <string>

#include "cron_walk.cuh"
#include "parse_kernels.cuh"
#include "unmet_walk.cuh"
#include "zone_table.h"

using namespace cronwalk;

extern "C" {

void* shim_table_build(const char* dir, int horizon_year) {
  auto* t = new b200cron::HostZoneTable();
  std::string err;
  if (!b200cron::build_zone_table(dir, horizon_year, t, &err)) {
    delete t;
    return nullptr;
  }
  return t;
}
void shim_table_free(void* t) { delete (b200cron::HostZoneTable*)t; }
int shim_zone_count(void* t) { return (int)((b200cron::HostZoneTable*)t)->names.size(); }
const char* shim_zone_name(void* t, int i) { return ((b200cron::HostZoneTable*)t)->names[i].c_str(); }
int shim_zone_find(void* t, const char* name) { return ((b200cron::HostZoneTable*)t)->find(name); }
int shim_zone_entries(void* t, int z, const int64_t** when, const int32_t** off) {
  auto* h = (b200cron::HostZoneTable*)t;
  *when = h->when.data() + h->zstart[z];
  *off = h->off.data() + h->zstart[z];
  return h->zstart[z + 1] - h->zstart[z];
}
int64_t shim_total_entries(void* t) { return (int64_t)((b200cron::HostZoneTable*)t)->when.size(); }

static ZoneTable view(void* t) {
  auto* h = (b200cron::HostZoneTable*)t;
  ZoneTable z;
  z.when = h->when.data();
  z.off = h->off.data();
  z.fpre = h->fpre.data();
  z.zstart = h->zstart.data();
  z.yidx = h->yidx.data();
  z.ystride = h->ystride;
  z.nzones = (int32_t)h->names.size();
  z.horizon = h->horizon;
  return z;
}

void shim_next_n(void* table, const uint64_t* const masks[6], const int32_t* zone_id, const int64_t* from, size_t n,
                 int k, int64_t* out) {
  ZoneTable zt = view(table);
  for (size_t i = 0; i < n; ++i) {
    Spec s = make_spec(masks[0][i], masks[1][i], masks[2][i], masks[3][i], masks[4][i], masks[5][i]);
    Walk w;
    const int32_t z = zone_id ? zone_id[i] : 0;
    if (z == 0) w.bind_utc(); else w.bind_zone(zt, z);
    if (k != 1 && !s.never) w.tod_last = spec_tod_last(s);  // as next_n_kernel does
    int64_t t = from[i];
    bool resident = false;
    // as next_n_tile_kernel does: a zoned spec deep inside one zone period runs on that period's clock, and falls
    // back to the general walk when its chain leaves the period
    int32_t shift = 0;
    int64_t nhi = kOmega;
    if (k > 1 && z != 0 && !s.every &&
        period_clock_next(zt, z, t, masks[4][i], -62135596800LL, 253402300799LL, zt.horizon - 6LL * 366 * 86400, &shift, &nhi) &&
        next_n_on_period_clock(s, t, k, shift, nhi, out + i * (size_t)k))
      continue;
    for (int q = 0; q < k; ++q) {
      if (t != kZero) {
        // zone_id == NULL: the all-UTC specialisation, as next1_kernel<true> is chosen for launches without zone ids
        t = zone_id ? next_from(s, w, t, resident) : next_from_t<false, true>(s, w, t, resident);
        resident = true;
      }
      out[i * (size_t)k + q] = t;
    }
  }
}

static uint8_t* g_klass = nullptr;  // optional: which class answered each Cron of the next shim_unmet call
void shim_set_klass_out(uint8_t* p) { g_klass = p; }

void shim_unmet(void* table, const uint64_t* const masks[6], const int32_t* zone_id, const int64_t* earliest,
                int64_t now, int32_t now_nsec, const uint8_t* policy, const uint32_t* active, const uint8_t* flags,
                const int64_t* deadline, size_t n, int64_t* last, int64_t* next, uint32_t* missed, uint8_t* status,
                uint8_t* decision) {
  ZoneTable zt = view(table);
  for (size_t i = 0; i < n; ++i) {
    Spec s = make_spec(masks[0][i], masks[1][i], masks[2][i], masks[3][i], masks[4][i], masks[5][i]);
    UnmetIn in;
    in.earliest = earliest[i];
    in.now = now;
    in.now_nsec = now_nsec;
    in.policy = policy ? policy[i] : 0;
    in.active = active ? active[i] : 0;
    in.flags = flags ? flags[i] : 0;
    in.deadline = deadline ? deadline[i] : kZero;
    // as the kernels do: class A (UTC, or one zone period holding the whole window) takes the UTC machine on that
    // period's clock, everything else -- and what that machine rejects -- the general one
    const int32_t z = zone_id ? zone_id[i] : 0;
    int klass = 0;
    UnmetOut o = unmet_decide_classed(s, zt, z, in, -62135596800LL, 253402300799LL, zt.horizon - 6LL * 366 * 86400, &klass);
    if (g_klass) g_klass[i] = (uint8_t)klass;
    last[i] = o.last; next[i] = o.next; missed[i] = o.missed; status[i] = o.status; decision[i] = o.decision;
  }
}

// The two-phase scheme of unmet_kernel: `easy_rounds` rounds of unmet_step_easy, and the general state machine from
// scratch for the Crons that did not finish.  rounds[i] = easy rounds used, deferred[i] = 1 when the general machine
// had to answer.  (easy_rounds <= 0: the general machine alone; rounds[i] then counts its steps.)
void shim_unmet_two_phase(void* table, const uint64_t* const masks[6], const int32_t* zone_id, const int64_t* earliest,
                          int64_t now, int32_t now_nsec, const uint8_t* policy, const uint32_t* active,
                          const uint8_t* flags, const int64_t* deadline, size_t n, int easy_rounds, int64_t* last,
                          int64_t* next, uint32_t* missed, uint8_t* status, uint8_t* decision, uint8_t* rounds,
                          uint8_t* deferred) {
  ZoneTable zt = view(table);
  for (size_t i = 0; i < n; ++i) {
    Spec s = make_spec(masks[0][i], masks[1][i], masks[2][i], masks[3][i], masks[4][i], masks[5][i]);
    UnmetIn in;
    in.earliest = earliest[i];
    in.now = now;
    in.now_nsec = now_nsec;
    in.policy = policy ? policy[i] : 0;
    in.active = active ? active[i] : 0;
    in.flags = flags ? flags[i] : 0;
    in.deadline = deadline ? deadline[i] : kZero;
    const int32_t z = zone_id ? zone_id[i] : 0;
    UnmetState st;
    unmet_begin(zt, z, in, st);
    bool defer = false;
    int r = 0;
    if (s.every) {
      st.o = unmet_decide_every((int64_t)s.second, in);
      st.done = true;
    }
    if (easy_rounds > 0) {
      while (!st.done) {
        if (r >= easy_rounds) { defer = true; break; }
        unmet_step_easy(s, in, st, &defer);
        ++r;
      }
    } else {
      defer = !st.done;
    }
    if (defer) {
      unmet_begin(zt, z, in, st);
      int g = 0;
      while (!st.done) { unmet_step(s, in, st); ++g; }
      if (easy_rounds <= 0) r = g > 255 ? 255 : g;
    }
    rounds[i] = (uint8_t)r; deferred[i] = defer ? 1 : 0;
    last[i] = st.o.last; next[i] = st.o.next; missed[i] = st.o.missed; status[i] = st.o.status; decision[i] = st.o.decision;
  }
}

// parse_batch_kernel's per-spec body (parse_kernels.cuh parse_one) over a batch laid out as the ABI takes it:
// concatenated text + n+1 offsets; tz_off relative to each spec's start, as the kernel writes it
void shim_parse_batch(const char* text, const uint64_t* text_off, size_t n, int with_seconds, uint64_t* const masks[6],
                      uint8_t* status, uint32_t* tz_off, uint32_t* tz_len) {
  for (size_t i = 0; i < n; ++i) {
    uint64_t m[6];
    status[i] = b200cron::parse_one(b200cron::Str{text + text_off[i], (int)(text_off[i + 1] - text_off[i])}, with_seconds, m,
                                    &tz_off[i], &tz_len[i]);
    for (int p = 0; p < 6; ++p) masks[p][i] = m[p];
  }
}

}  // extern "C"
