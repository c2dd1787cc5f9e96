// kernels.cuh -- sm_100a kernels of libb200cron.so.  All are HBM-bound integer kernels (no contraction, so no
// tensor cores): SoA mask planes are read with coalesced 128-bit loads, one lane walks one spec (cron_walk.cuh),
// and outputs leave through shared-memory staging so that every warp writes full 128-byte lines.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "cron_walk.cuh"
#include "unmet_walk.cuh"

namespace b200cron {

using cronwalk::kZero;
using cronwalk::Spec;
using cronwalk::Walk;
using cronwalk::ZoneTable;

// supported absolute-time range: years 1..9999 (k8s metav1.Time's RFC 3339 range)
constexpr int64_t kMinTime = -62135596800LL;
constexpr int64_t kMaxTime = 253402300799LL;
constexpr int64_t kSixYears = 6LL * 366 * 86400;

enum : int { kErrRange = 1, kErrZone = 2 };

struct MaskPlanes {
  const uint64_t* p[6];
};

__device__ __forceinline__ bool bind_checked(Walk& w, const ZoneTable& zt, int32_t zone, int64_t t, int* err) {
  w.bind_utc();
  if (t < kMinTime || t > kMaxTime) { atomicOr(err, kErrRange); return false; }
  if (zone == 0) { w.bind_utc(); return true; }
  if (zone < 0 || zone >= zt.nzones) { atomicOr(err, kErrZone); return false; }
  w.bind_zone(zt, zone);
  if (w.zn > 1 && t >= zt.horizon - kSixYears) { atomicOr(err, kErrRange); return false; }
  return true;
}

// -------------------------------------------------------------------------------------------------------------
// next_n: out[i*k + j] = (j+1)-th chained Next() after from[i]              (spec.go Next; cron_controller.go:405,409,436)
//
// Layout: each warp owns 64 consecutive specs.  Lane l loads specs (2l, 2l+1) of that tile from each of the 6
// planes and `from` with one 128-bit load per array (the warp reads 512 contiguous bytes per plane) and walks them
// one after the other.  k == 1: the two results leave as one 128-bit store.  k > 1: results are staged in the
// warp's slice of shared memory and written back row-major so that every warp store covers 256 contiguous bytes.
//   KT = 2,4,8,16,32 (compile-time k): element (row, j) lives at row*KT + ((j + lane_of_row) & (KT-1)); the rotation
//     spreads the lanes' simultaneous writes of column j over all banks, and the write-back un-rotates for free.
//   KT = 0 (any other k): transposed staging [j][row] and a division per element on the way out.
// Algorithmic bytes: 48 (masks) + 8 (from) [+ 4 zone] in, 8k out per spec.
// -------------------------------------------------------------------------------------------------------------
#ifndef B200CRON_NEXTN_STAGE
#define B200CRON_NEXTN_STAGE 1  // 0: results go straight to global memory, no shared-memory staging (tuning)
#endif
#ifndef B200CRON_NEXTN_MINBLOCKS
#define B200CRON_NEXTN_MINBLOCKS 12
#endif
constexpr bool kNextStage = B200CRON_NEXTN_STAGE != 0;
constexpr int kNextThreads = 64;
constexpr int kNextWarpTile = 64;
constexpr int kNextTile = kNextWarpTile * (kNextThreads / 32);

template <int KT, bool kVec>
__global__ void __launch_bounds__(kNextThreads, B200CRON_NEXTN_MINBLOCKS)
next_n_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ from, size_t n, int k_rt,
              int64_t* __restrict__ out, ZoneTable zt, int* __restrict__ err) {
  extern __shared__ __align__(16) int64_t stage_all[];  // [warps][64 * k], unused when KT == 1
  const int k = KT > 0 ? KT : k_rt;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const size_t wbase = (size_t)blockIdx.x * kNextTile + (size_t)warp * kNextWarpTile;
  if (wbase >= n) return;
  const size_t i0 = wbase + 2 * (size_t)lane;
  int64_t* stage = stage_all + (size_t)warp * kNextWarpTile * k;

  uint64_t mk[6][2];
  int64_t fr[2] = {0, 0};
  int32_t zz[2] = {0, 0};
  const bool full = i0 + 1 < n;
  if (kVec && full) {
#pragma unroll
    for (int p = 0; p < 6; ++p) {
      if (p == 0 && m.p[0] == nullptr) { mk[0][0] = 1; mk[0][1] = 1; continue; }  // 5-field schedules: second 0
      const ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2*>(m.p[p] + i0));
      mk[p][0] = v.x; mk[p][1] = v.y;
    }
    const longlong2 f = __ldg(reinterpret_cast<const longlong2*>(from + i0));
    fr[0] = f.x; fr[1] = f.y;
    if (zone_id) {
      const int2 z = __ldg(reinterpret_cast<const int2*>(zone_id + i0));
      zz[0] = z.x; zz[1] = z.y;
    }
  } else {
#pragma unroll
    for (int e = 0; e < 2; ++e) {
#pragma unroll
      for (int p = 0; p < 6; ++p)
        mk[p][e] = (i0 + e < n) ? ((p == 0 && m.p[0] == nullptr) ? 1ULL : __ldg(m.p[p] + i0 + e)) : 0;
      if (i0 + e < n) {
        fr[e] = __ldg(from + i0 + e);
        if (zone_id) zz[e] = __ldg(zone_id + i0 + e);
      }
    }
  }

  // The lane's two specs are walked one after the other through ONE copy of the walk in the instruction stream
  // (operands picked with selects): unrolling both into lock step doubled the code and spilled, and measured slower.
  Spec sp[2];
#pragma unroll
  for (int e = 0; e < 2; ++e) sp[e] = cronwalk::make_spec(mk[0][e], mk[1][e], mk[2][e], mk[3][e], mk[4][e], mk[5][e]);
  const int kk = KT == 1 ? 1 : k;
  int64_t ra = kZero, rb = kZero;
#pragma unroll 1
  for (int e = 0; e < 2; ++e) {
    const int row = 2 * lane + e;
    const bool live = i0 + e < n;  // out-of-range lanes stay in the loops (doing nothing) so the warp can re-sync
    const Spec ss = e ? sp[1] : sp[0];
    Walk w;
    int64_t t = e ? fr[1] : fr[0];
    bool ok = live && bind_checked(w, zt, e ? zz[1] : zz[0], t, err);
    if (!live) w.bind_utc();
    if (KT != 1 && !ss.never) w.tod_last = cronwalk::spec_tod_last(ss);  // lets chained calls skip exhausted days
#pragma unroll 1
    for (int j = 0; j < kk; ++j) {
      // Every lane runs exactly kk iterations: re-converge here, otherwise lanes that took the short path of one
      // iteration run ahead and the warp executes each path of each iteration separately.
      __syncwarp();
      if (ok && t != kZero) {
        t = cronwalk::next_from(ss, w, t, j > 0);
        // chained calls keep walking: re-check the zoned horizon (UTC never trips it)
        if (KT != 1 && t != kZero && w.zn > 1 && t >= zt.horizon - kSixYears) { atomicOr(err, kErrRange); ok = false; t = kZero; }
      } else {
        t = kZero;
      }
      if (!kNextStage) { if (KT != 1 && live) out[(i0 + e) * (size_t)k + j] = t; }
      else if (KT > 1) stage[row * KT + ((j + lane) & (KT - 1))] = t;
      else if (KT == 0) stage[j * kNextWarpTile + row] = t;
    }
    if (e) rb = t; else ra = t;
  }

  if (KT == 1) {
    if (kVec && full) {
      *reinterpret_cast<longlong2*>(out + i0) = make_longlong2(ra, rb);
    } else {
      if (i0 < n) out[i0] = ra;
      if (i0 + 1 < n) out[i0 + 1] = rb;
    }
    return;
  }
  if (!kNextStage) return;
  __syncwarp();
  const int rows = (int)((n - wbase) < (size_t)kNextWarpTile ? (n - wbase) : (size_t)kNextWarpTile);
  int64_t* dst = out + wbase * (size_t)k;
  if (KT > 1) {
    constexpr int kRowsPerIter = 32 / (KT > 1 ? KT : 1);
    const int rr = lane / (KT > 1 ? KT : 1), j = lane % (KT > 1 ? KT : 1);
    for (int r = rr; r < rows; r += kRowsPerIter)
      dst[(size_t)r * KT + j] = stage[r * KT + ((j + (r >> 1)) & (KT - 1))];
  } else {
    const int words = rows * k;
    for (int x = lane; x < words; x += 32) {
      const int r = x / k, j = x - r * k;
      dst[x] = stage[j * kNextWarpTile + r];
    }
  }
}


// -------------------------------------------------------------------------------------------------------------
// next1: the k == 1 case with ONE spec per lane (64-bit loads, still 256 contiguous bytes per plane per warp).
// A 1M-spec launch is only ~30 us long: with two specs per lane a resident warp gets ~4.4 tiles and the last,
// partly filled wave leaves SMs idle for ~20 % of the kernel; halving the tile halves that quantisation loss,
// which measured as more than the 128-bit loads save.  Used when k == 1; next_n_kernel handles k > 1.
// -------------------------------------------------------------------------------------------------------------
#ifndef B200CRON_NEXT1_THREADS
#define B200CRON_NEXT1_THREADS 128
#endif
#ifndef B200CRON_NEXT1_MINBLOCKS
#define B200CRON_NEXT1_MINBLOCKS 16
#endif
constexpr int kNext1Threads = B200CRON_NEXT1_THREADS;

template <bool kUtcOnly>
__global__ void __launch_bounds__(kNext1Threads, B200CRON_NEXT1_MINBLOCKS)
next1_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ from, size_t n,
             int64_t* __restrict__ out, ZoneTable zt, int* __restrict__ err) {
  // Programmatic dependent launch: when the host marks consecutive launches as independent
  // (b200cron_ctx_set_overlap), the next launch on the stream may start filling SMs as soon as every CTA of this one
  // has got this far, instead of idling through this launch's last partial wave.  A no-op for ordinary launches.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  const size_t i = (size_t)blockIdx.x * kNext1Threads + threadIdx.x;
  if (i >= n) return;
  const Spec s = cronwalk::make_spec(m.p[0] ? __ldg(m.p[0] + i) : 1ULL, __ldg(m.p[1] + i), __ldg(m.p[2] + i), __ldg(m.p[3] + i),
                                     __ldg(m.p[4] + i), __ldg(m.p[5] + i));
  const int64_t t = __ldg(from + i);
  const int32_t zone = (!kUtcOnly && zone_id) ? __ldg(zone_id + i) : 0;
  Walk w;
  const bool ok = bind_checked(w, zt, zone, t, err);
  out[i] = ok ? cronwalk::next_from_t<false, kUtcOnly>(s, w, t, false) : kZero;
}

// -------------------------------------------------------------------------------------------------------------
// unmet_since: getNextSchedule + gates + decision per Cron                    (cron_controller.go:163-220, 389-437)
// One lane per Cron; every array is read and written with lane-consecutive (coalesced) accesses.
// Algorithmic bytes: 48 + 8 + 4 + 1 + 4 + 1 + 8 = 74 in, 8 + 8 + 4 + 1 + 1 = 22 out per Cron.
// -------------------------------------------------------------------------------------------------------------
constexpr int kUnmetThreads = 128;

struct UnmetArgs {
  MaskPlanes m;
  const int32_t* zone_id;
  const int64_t* earliest;
  int64_t now;
  int32_t now_nsec;
  const uint8_t* policy;
  const uint32_t* active;
  const uint8_t* flags;
  const int64_t* deadline;
  size_t n;
  int64_t* last;
  int64_t* next;
  uint32_t* missed;
  uint8_t* status;
  uint8_t* decision;
};

// 12 CTAs/SM (40 registers, some spills): the kernel is bound by instruction-fetch redirects in branchy, divergent
// code (ncu: 21 % issue-active at 24 warps/SM); more resident warps hide them -- measured 8.3 -> 6.9 ms per 10M Crons.
//
// Modes.  kUnmetSingle: one launch does everything.  kUnmetEasy + kUnmetHard (the default pair): the easy kernel lets
// only the closed form answer, gives every Cron at most kUnmetEasyRounds rounds (first fire + fast-forward, a zone
// transition crossed + fast-forward, ..., end of chain, Next(now)) and appends the Crons that need more -- a step only
// the emulation can take, a backward step -- to a compact index list; the hard kernel then runs the general state
// machine over that list only, in dense warps.  On the C4 workload (10M Crons, B200) 3 rounds defer 8 % of the Crons
// and take 5.23 ms, 4 rounds 4.6 % / 4.79 ms, 5 rounds 2.4 % / 4.34 ms, 6 rounds the same 2.4 % / 4.35 ms: what is
// left after five rounds needs the step emulation, which in the single kernel (7.30 ms) made every second warp run it
// for one or two lanes.  The easy kernel has no step emulation in its instruction stream.
enum { kUnmetSingle = 0, kUnmetEasy = 1, kUnmetHard = 2 };
#ifndef B200CRON_UNMET_EASY_ROUNDS
#define B200CRON_UNMET_EASY_ROUNDS 5
#endif
constexpr int kUnmetEasyRounds = B200CRON_UNMET_EASY_ROUNDS;
#ifndef B200CRON_UNMET_EASY_MINBLOCKS
#define B200CRON_UNMET_EASY_MINBLOCKS 8
#endif
#ifndef B200CRON_UNMET_HARD_MINBLOCKS
#define B200CRON_UNMET_HARD_MINBLOCKS 12
#endif

// one Cron per lane; all 32 lanes of the warp must call this together (live = false for padding lanes)
template <int kMode>
__device__ __forceinline__ void unmet_one(const UnmetArgs& a, const ZoneTable& zt, int* __restrict__ err, size_t i,
                                          bool live, uint32_t* __restrict__ defer_idx,
                                          uint32_t* __restrict__ defer_count) {
  const size_t ii = live ? i : 0;  // out-of-range lanes stay in the warp-uniform loop below, doing nothing
  const Spec s = cronwalk::make_spec(__ldg(a.m.p[0] + ii), __ldg(a.m.p[1] + ii), __ldg(a.m.p[2] + ii),
                                     __ldg(a.m.p[3] + ii), __ldg(a.m.p[4] + ii), __ldg(a.m.p[5] + ii));
  cronwalk::UnmetIn in;
  in.earliest = __ldg(a.earliest + ii);
  in.now = a.now;
  in.now_nsec = a.now_nsec;
  in.policy = a.policy ? __ldg(a.policy + ii) : 0;
  in.active = a.active ? __ldg(a.active + ii) : 0;
  in.flags = a.flags ? __ldg(a.flags + ii) : 0;
  in.deadline = a.deadline ? __ldg(a.deadline + ii) : kZero;
  const int32_t zone = a.zone_id ? __ldg(a.zone_id + ii) : 0;

  bool ok = live;
  if (ok) {
    const int64_t lo = in.earliest < in.now ? in.earliest : in.now;
    if (lo < kMinTime || in.now > kMaxTime || in.earliest > kMaxTime) { atomicOr(err, kErrRange); ok = false; }
    if (ok && zone != 0) {
      if (zone < 0 || zone >= zt.nzones) { atomicOr(err, kErrZone); ok = false; }
      else if (__ldg(zt.zstart + zone + 1) - __ldg(zt.zstart + zone) > 1 && in.now >= zt.horizon - kSixYears) {
        atomicOr(err, kErrRange); ok = false;
      }
    }
  }
  cronwalk::UnmetState st;
  st.o.last = kZero; st.o.next = kZero; st.o.missed = 0; st.o.status = cronwalk::kStOk; st.o.decision = 0;
  st.done = true;
  st.w.bind_utc();
  if (ok) cronwalk::unmet_begin(zt, zone, in, st);
  // all 32 lanes go round together: a full-mask vote per iteration is the reconvergence point
  bool defer = false;
  if (kMode == kUnmetEasy) {
#pragma unroll 1
    for (int round = 0; round < kUnmetEasyRounds; ++round) {
      if (!__any_sync(0xffffffffu, !st.done)) break;
      if (!st.done) cronwalk::unmet_step_easy(s, in, st, &defer);
    }
    defer = defer || !st.done;  // still going after kUnmetEasyRounds rounds: a longer chain, the general kernel's business
    // warp-aggregated append of the deferred Crons to the compact list (ballot + prefix popcount, one atomic per warp)
    const unsigned lane = threadIdx.x & 31u;
    const unsigned mask = __ballot_sync(0xffffffffu, defer);
    if (mask) {
      const int leader = __ffs((int)mask) - 1;
      uint32_t base = 0;
      if ((int)lane == leader) base = atomicAdd(defer_count, (uint32_t)__popc(mask));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (defer) defer_idx[base + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = (uint32_t)i;
    }
  } else {
    while (__any_sync(0xffffffffu, !st.done)) {
      if (!st.done) cronwalk::unmet_step(s, in, st);
    }
  }
  if (live && !defer) {
    a.last[i] = st.o.last;
    a.next[i] = st.o.next;
    a.missed[i] = st.o.missed;
    a.status[i] = st.o.status;
    a.decision[i] = st.o.decision;
  }
}

// kOrdered: the launch walks the Crons in the order of an index list (zone_order_* kernels below) instead of input order
template <int kMode, bool kOrdered = false>
__global__ void __launch_bounds__(kUnmetThreads, kMode == kUnmetEasy ? B200CRON_UNMET_EASY_MINBLOCKS : 12)
unmet_kernel(UnmetArgs a, ZoneTable zt, int* __restrict__ err, uint32_t* __restrict__ defer_idx,
             uint32_t* __restrict__ defer_count, const uint32_t* __restrict__ order = nullptr) {
  const size_t q = (size_t)blockIdx.x * kUnmetThreads + threadIdx.x;
  const bool live = q < a.n;
  const size_t i = kOrdered ? (live ? (size_t)__ldg(order + q) : 0) : q;
  unmet_one<kMode>(a, zt, err, i, live, defer_idx, defer_count);
}

// the general state machine over the compact list the easy kernel left behind (grid-stride: the host does not know
// the list's length when it launches)
__global__ void __launch_bounds__(kUnmetThreads, B200CRON_UNMET_HARD_MINBLOCKS)
unmet_hard_kernel(UnmetArgs a, ZoneTable zt, int* __restrict__ err, const uint32_t* __restrict__ defer_idx,
                  const uint32_t* __restrict__ defer_count) {
  const uint32_t cnt = *defer_count;
  for (size_t base = (size_t)blockIdx.x * kUnmetThreads; base < cnt; base += (size_t)gridDim.x * kUnmetThreads) {
    const size_t q = base + threadIdx.x;
    const bool live = q < cnt;
    unmet_one<kUnmetHard>(a, zt, err, live ? defer_idx[q] : 0, live, nullptr, nullptr);
  }
}

// -------------------------------------------------------------------------------------------------------------
// Zone order (opt-in, B200CRON_UNMET_ZONE_ORDER=1; written at the end of round 1 and NOT yet run on a GPU -- see
// DESIGN.md section 8).  The easy kernel spends its time in zone-dependent code that 4-5 of a warp's 32 lanes execute
// together: half of the lanes are UTC and the zoned ones sit in different zones, hence in different periods.  A
// counting sort of the Cron indices by zone id (histogram, scan, scatter; one pass over zone_id each) lets the easy
// kernel walk zone-pure warps through an index list; inputs are then gathered and outputs scattered 8 bytes at a
// time, which is cheap next to the divergence it removes.  Ids outside [0, nzones) are binned with UTC; unmet_one
// reports them as before.
// -------------------------------------------------------------------------------------------------------------
constexpr int kZoneOrderThreads = 256;
constexpr int kZoneOrderItems = 16;  // Crons per thread: a CTA bins 4096 consecutive Crons
constexpr int kZoneOrderTile = kZoneOrderThreads * kZoneOrderItems;

__device__ __forceinline__ uint32_t zone_bin(const int32_t* __restrict__ zone_id, size_t i, int32_t nzones) {
  const int32_t z = __ldg(zone_id + i);
  return (z < 0 || z >= nzones) ? 0u : (uint32_t)z;
}

// the CTA's histogram of its tile in shared memory (bins [0, nzones)); all threads call it
__device__ __forceinline__ void zone_tile_hist(const int32_t* __restrict__ zone_id, size_t n, int32_t nzones,
                                               uint32_t* __restrict__ sh) {
  for (int b = threadIdx.x; b < nzones; b += kZoneOrderThreads) sh[b] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * kZoneOrderTile;
#pragma unroll 4
  for (int it = 0; it < kZoneOrderItems; ++it) {
    const size_t i = base + (size_t)it * kZoneOrderThreads + threadIdx.x;
    if (i < n) atomicAdd(&sh[zone_bin(zone_id, i, nzones)], 1u);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kZoneOrderThreads)
zone_order_hist_kernel(const int32_t* __restrict__ zone_id, size_t n, int32_t nzones, uint32_t* __restrict__ hist) {
  extern __shared__ uint32_t zo_sh[];
  zone_tile_hist(zone_id, n, nzones, zo_sh);
  for (int b = threadIdx.x; b < nzones; b += kZoneOrderThreads)
    if (zo_sh[b]) atomicAdd(hist + b, zo_sh[b]);
}

// exclusive prefix sum of the histogram into `cursor` (one CTA; a thread owns a contiguous run of bins)
__global__ void __launch_bounds__(kZoneOrderThreads)
zone_order_scan_kernel(const uint32_t* __restrict__ hist, int32_t nzones, uint32_t* __restrict__ cursor) {
  __shared__ uint32_t part[kZoneOrderThreads];
  const int per = (nzones + kZoneOrderThreads - 1) / kZoneOrderThreads;
  const int lo = threadIdx.x * per, hi = min(lo + per, (int)nzones);
  uint32_t sum = 0;
  for (int b = lo; b < hi; ++b) sum += hist[b];
  part[threadIdx.x] = sum;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = 0;
    for (int t = 0; t < kZoneOrderThreads; ++t) { const uint32_t v = part[t]; part[t] = run; run += v; }
  }
  __syncthreads();
  uint32_t run = part[threadIdx.x];
  for (int b = lo; b < hi; ++b) { cursor[b] = run; run += hist[b]; }
}

// every CTA reserves, per zone, a range of the output for its tile's Crons and fills it
__global__ void __launch_bounds__(kZoneOrderThreads)
zone_order_scatter_kernel(const int32_t* __restrict__ zone_id, size_t n, int32_t nzones, uint32_t* __restrict__ cursor,
                          uint32_t* __restrict__ order) {
  extern __shared__ uint32_t zo_sh[];  // [nzones] counts, then local cursors; [nzones] bases
  uint32_t* sh_base = zo_sh + nzones;
  zone_tile_hist(zone_id, n, nzones, zo_sh);
  for (int b = threadIdx.x; b < nzones; b += kZoneOrderThreads) {
    const uint32_t c = zo_sh[b];
    sh_base[b] = c ? atomicAdd(cursor + b, c) : 0u;
    zo_sh[b] = 0;
  }
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * kZoneOrderTile;
#pragma unroll 4
  for (int it = 0; it < kZoneOrderItems; ++it) {
    const size_t i = base + (size_t)it * kZoneOrderThreads + threadIdx.x;
    if (i < n) {
      const uint32_t b = zone_bin(zone_id, i, nzones);
      order[sh_base[b] + atomicAdd(&zo_sh[b], 1u)] = (uint32_t)i;
    }
  }
}

// -------------------------------------------------------------------------------------------------------------
// history_prune: stable ascending order of each Cron's finished workloads by creation time + keep/delete flags
//                                                           (cron_util.go:117-129, cron_controller.go:313-323)
//
// A CTA owns kHistSegs consecutive segments, i.e. one contiguous run of records.  The run's keys are brought into
// shared memory with coalesced loads, every record finds its segment by binary search over the tile's offsets and
// its stable rank by counting the records of its segment that sort before it (segments are short: the reference
// keeps historyLimit of them), and perm/keep go straight to global memory (sector-coalesced, see below).
// Runs longer than kHistCap records are ranked straight from global memory (L2-resident) -- correct for any
// length, but quadratic in the segment length.
// Algorithmic bytes: 8 (key) in, 4 (perm) + 1 (keep) out per record; 8 (offset) + 8 (limit) in per segment.
// -------------------------------------------------------------------------------------------------------------
constexpr int kHistThreads = 256;
constexpr int kHistSegs = 256;
constexpr int kHistCap = 3072;  // records per CTA tile held in shared memory (31 KB static with seg ids/offsets/limits)

__global__ void __launch_bounds__(kHistThreads)
history_prune_kernel(const uint64_t* __restrict__ seg_off, const int64_t* __restrict__ key,
                     const int64_t* __restrict__ limit, size_t nseg, uint32_t* __restrict__ perm,
                     uint8_t* __restrict__ keep) {
  __shared__ uint64_t s_off[kHistSegs + 1];
  __shared__ int64_t s_key[kHistCap];
  const size_t seg0 = (size_t)blockIdx.x * kHistSegs;
  const int nsegs = (int)((nseg - seg0) < (size_t)kHistSegs ? (nseg - seg0) : (size_t)kHistSegs);
  for (int x = threadIdx.x; x <= nsegs; x += kHistThreads) s_off[x] = __ldg(seg_off + seg0 + x);
  __syncthreads();
  const uint64_t r0 = s_off[0];
  const uint64_t total = s_off[nsegs] - r0;
  const bool in_smem = total <= (uint64_t)kHistCap;
  if (in_smem) {
    for (uint32_t x = threadIdx.x; x < total; x += kHistThreads) s_key[x] = __ldg(key + r0 + x);
    __syncthreads();
  }
  if (in_smem) {
    // Shared-memory path, all 32-bit.  Each segment's thread stamps its id over its own records (segments are
    // short), so a record finds its segment with one load instead of a binary search over the offsets.
    __shared__ uint8_t s_seg[kHistCap];
    __shared__ uint32_t s_rel[kHistSegs + 1];
    __shared__ int32_t s_lim[kHistSegs];
    const uint32_t tot = (uint32_t)total;
    for (int sg = threadIdx.x; sg <= nsegs; sg += kHistThreads) s_rel[sg] = (uint32_t)(s_off[sg] - r0);
    for (int sg = threadIdx.x; sg < nsegs; sg += kHistThreads) {
      const int64_t lim = limit ? __ldg(limit + seg0 + sg) : -1;
      s_lim[sg] = (lim < 0 || lim > (int64_t)kHistCap) ? (int32_t)kHistCap : (int32_t)lim;  // >= len keeps everything
      const uint32_t a0 = (uint32_t)(s_off[sg] - r0), b0 = (uint32_t)(s_off[sg + 1] - r0);
      for (uint32_t j = a0; j < b0; ++j) s_seg[j] = (uint8_t)sg;
    }
    __syncthreads();
    for (uint32_t x = threadIdx.x; x < tot; x += kHistThreads) {
      const int sg = s_seg[x];
      const uint32_t a = s_rel[sg], b = s_rel[sg + 1];
      const int64_t kx = s_key[x];
      uint32_t rank = 0;
      for (uint32_t j = a; j < b; ++j) {
        const int64_t kj = s_key[j];
        rank += (kj < kx) || (kj == kx && j < x);
      }
      // sorted position `rank` is deleted when rank < len - limit.  Written straight to global memory: the lanes of a
      // warp hold consecutive records, i.e. a handful of whole segments, and a permutation inside a segment touches
      // the same 32-byte sectors as the identity would -- staging the outputs in shared memory bought nothing and
      // cost 15 KB per CTA (4 resident CTAs per SM instead of 7).
      perm[r0 + a + rank] = x - a;
      keep[r0 + a + rank] = !((int32_t)rank < (int32_t)(b - a) - s_lim[sg]);
    }
    return;
  }
  // Long runs (more than kHistCap records in the tile): rank straight from global memory.
  for (uint64_t x = threadIdx.x; x < total; x += kHistThreads) {
    int lo = 0, hi = nsegs;  // largest sg with s_off[sg] - r0 <= x
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_off[mid] - r0 <= x) lo = mid; else hi = mid;
    }
    const uint64_t a = s_off[lo] - r0, b = s_off[lo + 1] - r0;
    const uint64_t len = b - a;
    const int64_t kx = __ldg(key + r0 + x);
    uint32_t rank = 0;
    for (uint64_t j = a; j < b; ++j) {
      const int64_t kj = __ldg(key + r0 + j);
      rank += (kj < kx) || (kj == kx && j < x);
    }
    const int64_t lim = limit ? __ldg(limit + seg0 + lo) : -1;
    // sorted position `rank` is deleted when rank < len - limit (limit < 0: HistoryLimit nil keeps everything)
    perm[r0 + a + rank] = (uint32_t)(x - a);
    keep[r0 + a + rank] = (lim < 0) ? 1 : !((int64_t)rank < (int64_t)len - lim);
  }
}

// -------------------------------------------------------------------------------------------------------------
// unmet list: the first min(missed, cap) missed fire times of every Cron, packed back to back
//                                   (the enumeration of cron_controller.go:409-430, which upstream only counts)
//
// list_off = exclusive prefix sum of the per-Cron counts (three small kernels: per-tile sums into a scratch array,
// a single-CTA scan of the tile sums, per-tile offsets = warp-shuffle scan + tile base); unmet_list_emit_kernel then
// re-walks each chain and writes its fire times at list + list_off[i].  Gated / unschedulable Crons contribute 0.
// Scratch: one uint64 per tile of kScanTile Crons.
// -------------------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 4;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ uint32_t list_count(uint32_t missed, uint8_t decision, uint32_t cap) {
  const bool listed = decision >= cronwalk::kDecRequeueOnly;  // REQUEUE_ONLY, SKIP_FORBID, REPLACE_CREATE, CREATE
  return listed ? (missed < cap ? missed : cap) : 0u;
}

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) >= d) v += o;
  }
  return v;
}

// block-wide exclusive scan of one value per thread; returns the thread's exclusive prefix, *total = block sum
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* total) {
  __shared__ uint32_t warp_sums[kScanThreads / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t inc = warp_inclusive_scan(v);
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const uint32_t ws = lane < kScanThreads / 32 ? warp_sums[lane] : 0u;
    const uint32_t wi = warp_inclusive_scan(ws);
    if (lane < kScanThreads / 32) warp_sums[lane] = wi - ws;  // exclusive warp bases
    if (lane == kScanThreads / 32 - 1) *total = wi;
  }
  __syncthreads();
  const uint32_t base = warp_sums[warp];
  __syncthreads();
  return base + inc - v;
}

__global__ void __launch_bounds__(kScanThreads)
list_tile_sums_kernel(const uint32_t* __restrict__ missed, const uint8_t* __restrict__ decision, size_t n,
                      uint32_t cap, uint64_t* __restrict__ tile_sums) {
  __shared__ uint32_t total;
  const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  uint32_t v = 0;
#pragma unroll
  for (int q = 0; q < kScanItems; ++q)
    if (base + q < n) v += list_count(__ldg(missed + base + q), __ldg(decision + base + q), cap);
  block_exclusive_scan(v, &total);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = total;
}

// single CTA: in-place exclusive scan of the tile sums; tile_sums[ntiles] receives the grand total
__global__ void __launch_bounds__(kScanThreads)
list_scan_sums_kernel(uint64_t* __restrict__ tile_sums, size_t ntiles) {
  __shared__ uint64_t carry;
  __shared__ uint64_t part[kScanThreads];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (size_t c = 0; c < ntiles; c += kScanThreads) {
    const size_t i = c + threadIdx.x;
    const uint64_t v = i < ntiles ? tile_sums[i] : 0;
    part[threadIdx.x] = v;
    __syncthreads();
    for (int d = 1; d < kScanThreads; d <<= 1) {  // Hillis-Steele over one chunk
      const uint64_t o = threadIdx.x >= d ? part[threadIdx.x - d] : 0;
      __syncthreads();
      part[threadIdx.x] += o;
      __syncthreads();
    }
    if (i < ntiles) tile_sums[i] = carry + part[threadIdx.x] - v;
    __syncthreads();
    if (threadIdx.x == kScanThreads - 1) carry += part[kScanThreads - 1];
    __syncthreads();
  }
  if (threadIdx.x == 0) tile_sums[ntiles] = carry;
}

__global__ void __launch_bounds__(kScanThreads)
list_offsets_kernel(const uint32_t* __restrict__ missed, const uint8_t* __restrict__ decision, size_t n, uint32_t cap,
                    const uint64_t* __restrict__ tile_sums, uint64_t* __restrict__ list_off) {
  __shared__ uint32_t total;
  const size_t base = (size_t)blockIdx.x * kScanTile + (size_t)threadIdx.x * kScanItems;
  uint32_t c[kScanItems];
  uint32_t v = 0;
#pragma unroll
  for (int q = 0; q < kScanItems; ++q) {
    c[q] = (base + q < n) ? list_count(__ldg(missed + base + q), __ldg(decision + base + q), cap) : 0u;
    v += c[q];
  }
  uint64_t run = tile_sums[blockIdx.x] + block_exclusive_scan(v, &total);
#pragma unroll
  for (int q = 0; q < kScanItems; ++q) {
    if (base + q < n) list_off[base + q] = run;
    run += c[q];
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) list_off[n] = tile_sums[gridDim.x];
}

__global__ void __launch_bounds__(kUnmetThreads)
unmet_list_emit_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ earliest,
                       size_t n, const uint64_t* __restrict__ list_off, int64_t* __restrict__ list, ZoneTable zt) {
  const size_t i = (size_t)blockIdx.x * kUnmetThreads + threadIdx.x;
  if (i >= n) return;
  const uint64_t a = __ldg(list_off + i), b = __ldg(list_off + i + 1);
  if (b == a) return;
  const Spec s = cronwalk::make_spec(__ldg(m.p[0] + i), __ldg(m.p[1] + i), __ldg(m.p[2] + i), __ldg(m.p[3] + i),
                                     __ldg(m.p[4] + i), __ldg(m.p[5] + i));
  const int32_t zone = zone_id ? __ldg(zone_id + i) : 0;
  Walk w;
  if (zone == 0) w.bind_utc(); else w.bind_zone(zt, zone);  // ids and ranges were validated by unmet_kernel
  int64_t t = __ldg(earliest + i);
  bool resident = false;
  for (uint64_t j = a; j < b; ++j) {
    t = cronwalk::next_from_t<true>(s, w, t, resident);
    resident = true;
    list[j] = t;
  }
}

}  // namespace b200cron
