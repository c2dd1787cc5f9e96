// kernels.cuh -- sm_100a kernels of libb200cron.so.  Integer kernels bound by the scope table to the HBM line (no
// contraction, so no tensor cores): SoA mask planes are read with coalesced loads, one lane walks one spec
// (cron_walk.cuh), and warps are made class-pure (UTC / one zone period's clock / general zoned) by ballot compaction
// before the walk, so that the short instantiations of the walk run in dense warps.
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

enum : int { kErrRange = 1, kErrZone = 2, kErrSegLen = 4 };

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

#ifndef B200CRON_NEXTN_SPLIT
#define B200CRON_NEXTN_SPLIT 1  // 0 (tuning builds): k > 1 launches use next_n_kernel over the specs in input order
#endif

#if !B200CRON_NEXTN_SPLIT
// -------------------------------------------------------------------------------------------------------------
// (tuning builds only: round 1's kernel, kept as the A/B baseline of next_n_tile_kernel)
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
#define B200CRON_NEXTN_STAGE 0  // 1: stage results in shared memory and write full rows (measured 6 % slower on C3)
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
        if (KT != 1 && t != kZero && !ss.every && w.zn > 1 && t >= zt.horizon - kSixYears) { atomicOr(err, kErrRange); ok = false; t = kZero; }
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


#endif  // !B200CRON_NEXTN_SPLIT

// -------------------------------------------------------------------------------------------------------------
// next_n: out[i*k + j] = (j+1)-th chained Next() after from[i]              (spec.go Next; cron_controller.go:405,409,436)
// over class-pure warps (k > 1).  In input order half of a C3 warp's lanes are UTC and ride
// through the zoned instruction stream (zone prerequisites, period lookups, the step emulation), and the kernel sits at
// 80 registers for everybody.  Here a CTA takes a tile of kNextTileSpecs consecutive specs, deals the indices of ONE
// class -- evaluated in UTC (kUtc) or in a zone -- into shared memory with a ballot compaction and walks only those, in
// dense warps; the host launches the kernel once per class.  The UTC instantiation has no zone logic and no step
// emulation in its instruction stream (in UTC the closed form always answers) and half the registers, so twice the
// resident warps.  Loads are gathers inside the tile (every sector is used by one class or the other; the second
// launch finds most of them in L2), results go straight to the spec's row (the pattern B200CRON_NEXTN_STAGE=0 measured
// best).  A launch without zone ids is all UTC: no classification.
// -------------------------------------------------------------------------------------------------------------
constexpr int kNextTileThreads = 128;
constexpr int kNextTileSpecs = 512;         // class-A launch: specs per CTA
constexpr int kNextTileSpecsZoned = 2048;   // zoned launch: it takes ~7 % of the specs, so a CTA looks at four times as
                                            // many to keep all of its warps busy (with 512 the launch was latency-bound)
#ifndef B200CRON_NEXTN_UTC_MINBLOCKS
#define B200CRON_NEXTN_UTC_MINBLOCKS 8  // 64 registers (no spills): 1.90 ms per 8M x 16 against 2.11 at 10 and 2.42 at 12
#endif
#ifndef B200CRON_NEXTN_ZONED_OUTLINE
#define B200CRON_NEXTN_ZONED_OUTLINE 0  // 1 (tuning builds): the zoned launch calls the step emulation out of line
#endif
#ifndef B200CRON_NEXTN_PAIR
#define B200CRON_NEXTN_PAIR 1
#endif
#ifndef B200CRON_NEXTN_ZONED_MINBLOCKS
#define B200CRON_NEXTN_ZONED_MINBLOCKS 8  // 64 registers: the launch is latency-bound, more resident warps win (2.11 vs 2.19 ms)
#endif

// A row whose first entry is kNextRejected was handed over by the class-A launch -- a zoned spec near a transition, or a
// chain that left the zone period it started in -- and the zoned launch (next on the stream) walks it with the general
// code.
constexpr int64_t kNextRejected = INT64_MIN + 1;  // never a result: results are >= kMinTime, or kZero

template <int KT, bool kUtc>
__global__ void __launch_bounds__(kNextTileThreads, kUtc ? B200CRON_NEXTN_UTC_MINBLOCKS : B200CRON_NEXTN_ZONED_MINBLOCKS)
next_n_tile_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ from, size_t n, int k_rt,
                   int64_t* out, ZoneTable zt, int* __restrict__ err) {
  // kUtc launch = class A: specs evaluated in UTC, and zoned specs whose first Date() reset target lies naturally in
  // the zone period of their start -- their chain runs on that period's clock (cronwalk::period_clock_next) and is
  // rejected if a fire leaves the period.
  // The other launch = everything else: zoned specs near a transition, and the rejected rows.
  constexpr int kTile = kUtc ? kNextTileSpecs : kNextTileSpecsZoned;
  __shared__ uint16_t s_idx[kTile];
  __shared__ int32_t s_shift[kUtc ? kTile : 1];
  __shared__ int64_t s_nhi[kUtc ? kTile : 1];
  __shared__ int s_n;
  const int k = KT > 0 ? KT : k_rt;
  const size_t base = (size_t)blockIdx.x * kTile;
  const int cnt = (int)((n - base) < (size_t)kTile ? (n - base) : (size_t)kTile);
  const unsigned lane = threadIdx.x & 31u;
  int mine = cnt;
  if (zone_id != nullptr) {
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
#pragma unroll 1
    for (int it = 0; it < kTile / kNextTileThreads; ++it) {
      const int q = it * kNextTileThreads + threadIdx.x;
      bool take = false;
      int32_t shift = 0;
      int64_t nhi = cronwalk::kOmega;
      if (q < cnt) {
        const int32_t z = __ldg(zone_id + base + q);
        if (kUtc) {
          // class A?  Everything else is handed to the zoned launch by the same mark a rejected chain gets.
          take = z == 0 || cronwalk::period_clock_next(zt, z, __ldg(from + base + q), __ldg(m.p[4] + base + q), kMinTime,
                                                       kMaxTime, zt.horizon - kSixYears, &shift, &nhi);
          if (!take) out[(base + q) * (size_t)k] = kNextRejected;
        } else {
          take = z != 0 && out[(base + q) * (size_t)k] == kNextRejected;
        }
      }
      const unsigned mk = __ballot_sync(0xffffffffu, take);
      int b = 0;
      if (lane == 0 && mk) b = atomicAdd(&s_n, __popc(mk));
      b = __shfl_sync(0xffffffffu, b, 0);
      if (take) {
        const int slot = b + __popc(mk & ((1u << lane) - 1u));
        s_idx[slot] = (uint16_t)q;
        if (kUtc) { s_shift[slot] = shift; s_nhi[slot] = nhi; }
      }
    }
    __syncthreads();
    mine = s_n;
  } else if (!kUtc) {
    return;  // no zone ids: nothing is zoned
  }
  for (int q = threadIdx.x; q < ((mine + 31) & ~31); q += kNextTileThreads) {
    const bool live = q < mine;
    const size_t i = base + (live ? (zone_id ? (int)s_idx[q] : q) : 0);
    const Spec ss = cronwalk::make_spec(m.p[0] ? __ldg(m.p[0] + i) : 1ULL, __ldg(m.p[1] + i), __ldg(m.p[2] + i),
                                        __ldg(m.p[3] + i), __ldg(m.p[4] + i), __ldg(m.p[5] + i));
    int64_t t = __ldg(from + i);
    const int32_t zone = kUtc ? 0 : __ldg(zone_id + i);
    const int32_t shift = (kUtc && live && zone_id && !ss.every) ? s_shift[q] : 0;
    const int64_t nhi = (kUtc && live && zone_id && !ss.every) ? s_nhi[q] : cronwalk::kOmega;
    const bool on_period = kUtc && (shift != 0 || nhi != cronwalk::kOmega);
    Walk w;
    bool ok = live && bind_checked(w, zt, zone, t, err);
    if (!live) w.bind_utc();
    if (!ss.never) w.tod_last = cronwalk::spec_tod_last(ss);  // lets chained calls skip exhausted days
    t += shift;  // onto the period's clock (0 for UTC)
    bool left = false;
    int64_t* row = out + i * (size_t)k;
    // compile-time even k and a 16-byte aligned output: results leave two at a time (half the store instructions)
    const bool pair = B200CRON_NEXTN_PAIR && KT > 0 && (KT & 1) == 0 && (reinterpret_cast<uintptr_t>(out) & 15u) == 0;
    int64_t held = kZero;
#pragma unroll 1
    for (int j = 0; j < k; ++j) {
      __syncwarp();  // every lane runs exactly k iterations: re-converge, or lanes on a short path run ahead for good
      int64_t r = kZero;
      if (ok && t != kZero) {
        t = cronwalk::next_from_t<(!kUtc && B200CRON_NEXTN_ZONED_OUTLINE), kUtc>(ss, w, t, j > 0);
        if (!kUtc && t != kZero && !ss.every && w.zn > 1 && t >= zt.horizon - kSixYears) { atomicOr(err, kErrRange); ok = false; t = kZero; }
        if (on_period) left = left || t == kZero || t >= nhi;  // the chain left its period: the general walk answers
        r = t == kZero ? kZero : t - shift;
      } else {
        t = kZero;
      }
      if (pair) {
        if (j & 1) { if (live) *reinterpret_cast<longlong2*>(row + j - 1) = make_longlong2(held, r); }
        else held = r;
      } else if (live) {
        row[j] = r;
      }
    }
    if (kUtc && live && left) row[0] = kNextRejected;
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
#ifndef B200CRON_NEXT1_VOTE5
#define B200CRON_NEXT1_VOTE5 1  // 0 (tuning builds): six-plane UTC launches always run the general walk
#endif

// kFive: the launch has no seconds plane (5-field schedules, the only kind ParseStandard builds: second 0).  The
// seconds mask is then the compile-time constant 1 and everything derived from it folds away (first matching second,
// the seconds search of next_tod_ge, the sentinel scans of make_spec).
template <bool kUtcOnly, bool kFive>
__global__ void __launch_bounds__(kNext1Threads, B200CRON_NEXT1_MINBLOCKS)
next1_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ from, size_t n,
             int64_t* __restrict__ out, ZoneTable zt, int* __restrict__ err) {
  // Programmatic dependent launch: when the host marks consecutive launches as independent
  // (b200cron_ctx_set_overlap), the next launch on the stream may start filling SMs as soon as every CTA of this one
  // has got this far, instead of idling through this launch's last partial wave.  A no-op for ordinary launches.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // 32-bit element index (the host keeps a launch below 2^32 specs): loads and the store address their arrays as
  // uniform base + 32-bit offset instead of a 64-bit add per array
  const uint32_t i = blockIdx.x * (uint32_t)kNext1Threads + threadIdx.x;
  if (i >= (uint32_t)n) { asm volatile("griddepcontrol.wait;" ::: "memory"); return; }
  const uint64_t sec = kFive ? 1ULL : (m.p[0] ? __ldg(m.p[0] + i) : 1ULL);
  const uint64_t mi = __ldg(m.p[1] + i), ho = __ldg(m.p[2] + i), dm = __ldg(m.p[3] + i), mo = __ldg(m.p[4] + i), dw = __ldg(m.p[5] + i);
  const int64_t t = __ldg(from + i);
  const int32_t zone = (!kUtcOnly && zone_id) ? __ldg(zone_id + i) : 0;
  Walk w;
  const bool ok = bind_checked(w, zt, zone, t, err);
  int64_t r = kZero;
  // A launch that does carry a seconds plane still runs the 5-field code when all of the warp's specs fire at second
  // 0 only (what ParseStandard always yields; C2): the warp-uniform vote picks the copy of the walk in which the
  // seconds mask is the constant 1.  UTC launches only: the zoned walk is too large to hold twice.
  if (kFive || (kUtcOnly && B200CRON_NEXT1_VOTE5 && __all_sync(__activemask(), sec == 1ULL))) {
    const Spec s = cronwalk::make_spec(1ULL, mi, ho, dm, mo, dw);
    if (ok) r = cronwalk::next_from_t<false, kUtcOnly>(s, w, t, false);
  } else {
    const Spec s = cronwalk::make_spec(sec, mi, ho, dm, mo, dw);
    if (ok) r = cronwalk::next_from_t<false, kUtcOnly>(s, w, t, false);
  }
  // The other half of the programmatic launch: wait for the launch this one may have overlapped to complete (and
  // flush) before publishing anything.  Completion is then transitive along the stream -- when this grid has
  // finished, so has every earlier one -- which is what later stream work (copies, event records, other kernels)
  // relies on.  Returns at once for a launch that overlapped nothing.
  asm volatile("griddepcontrol.wait;" ::: "memory");
  out[i] = r;
}

// -------------------------------------------------------------------------------------------------------------
// next1 with persistent warps and an asynchronous input ring.  TUNING BUILD ONLY (-DB200CRON_NEXT1_RING=1): measured
// slower than next1_kernel (18.8 vs 13.7 us per 1M specs), kept as the record of that experiment.
//
// next1_kernel gives every warp one tile: issue seven loads, wait a DRAM round trip, walk, store, exit -- the wait is
// the kernel's top stall, and a 1M-spec launch ends in a ragged last wave.  Here the grid is sized to the machine
// (resident CTAs only) and every warp walks a strided sequence of 32-spec tiles.  A warp owns kRingStages stages of
// shared memory (7 planes x 256 B + 128 B of zone ids each) and one mbarrier per stage; lane 0 fills a stage with 1-D
// bulk async copies (cp.async.bulk, one per plane, completion counted in bytes on the stage's mbarrier), so the loads of
// tile i+2 are in flight while tile i is walked and a warp never waits for DRAM after its first tile.  Warps do not
// synchronise with each other at all: stage reuse is ordered by __syncwarp (the warp is the stage's only reader).
// Bulk copies read THIS GPU's global memory only (the host passes device-resident buffers; pinned host buffers take
// next1_kernel).  The n % 32 leftover specs are walked by warp 0 with plain loads.
// -------------------------------------------------------------------------------------------------------------
#if defined(B200CRON_NEXT1_RING) && B200CRON_NEXT1_RING
#ifndef B200CRON_RING_STAGES
#define B200CRON_RING_STAGES 2
#endif
#ifndef B200CRON_RING_MINBLOCKS
#define B200CRON_RING_MINBLOCKS 12
#endif
constexpr int kRingThreads = 128;
constexpr int kRingWarps = kRingThreads / 32;
constexpr int kRingStages = B200CRON_RING_STAGES;
constexpr int kRingStageBytes = 7 * 256 + 128;  // six mask planes + from (32 x 8 B each), zone ids (32 x 4 B)
constexpr int kRingSmemBytes = kRingWarps * kRingStages * kRingStageBytes + kRingWarps * kRingStages * 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity) : "memory");
}
// 1-D bulk copy global -> this CTA's shared memory, completion signalled on `bar` in bytes
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

template <bool kUtcOnly>
__global__ void __launch_bounds__(kRingThreads, B200CRON_RING_MINBLOCKS)
next1_ring_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ from, size_t n,
                  int64_t* __restrict__ out, ZoneTable zt, int* __restrict__ err) {
  extern __shared__ __align__(128) unsigned char ring_smem[];
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // see next1_kernel
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned char* my = ring_smem + (size_t)warp * kRingStages * kRingStageBytes;
  const uint32_t my_u32 = smem_u32(my);
  const uint32_t bar0 = smem_u32(ring_smem + kRingWarps * kRingStages * kRingStageBytes) + (uint32_t)warp * kRingStages * 8u;
  const bool have_sec = m.p[0] != nullptr;
  const bool have_zone = !kUtcOnly && zone_id != nullptr;
  const uint32_t stage_tx = (have_sec ? 7u : 6u) * 256u + (have_zone ? 128u : 0u);
  const size_t tiles = n / 32;                                   // full tiles; the rest is the tail below
  const size_t W = (size_t)gridDim.x * kRingWarps;               // warps in the grid
  const size_t g = (size_t)blockIdx.x + (size_t)gridDim.x * warp;  // consecutive tiles go to different CTAs
  if (lane == 0) {
#pragma unroll
    for (int st = 0; st < kRingStages; ++st) mbar_init(bar0 + 8u * st, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  auto fill = [&](size_t tile, int st) {  // lane 0 only
    const uint32_t dst = my_u32 + (uint32_t)st * kRingStageBytes, bar = bar0 + 8u * st;
    const size_t e = tile * 32;
    mbar_expect_tx(bar, stage_tx);
#pragma unroll
    for (int p = 1; p < 6; ++p) bulk_g2s(dst + p * 256, m.p[p] + e, 256, bar);
    if (have_sec) bulk_g2s(dst, m.p[0] + e, 256, bar);
    bulk_g2s(dst + 6 * 256, from + e, 256, bar);
    if (have_zone) bulk_g2s(dst + 7 * 256, zone_id + e, 128, bar);
  };
  if (lane == 0) {
#pragma unroll
    for (int st = 0; st < kRingStages; ++st)
      if (g + (size_t)st * W < tiles) fill(g + (size_t)st * W, st);
  }
  bool waited = false;
  int st = 0;
  uint32_t parity = 0;
  // full tiles come through the ring; the n % 32 leftover specs form one more tile that is read with plain loads
  // (same loop body, so the walk exists once in the instruction stream)
  const size_t all_tiles = (n + 31) / 32;
  for (size_t tile = g; tile < all_tiles; tile += W) {
    const size_t i = tile * 32 + lane;
    uint64_t mk[6];
    int64_t t;
    int32_t zone = 0;
    if (tile < tiles) {  // warp-uniform
      mbar_wait(bar0 + 8u * st, parity);
      const unsigned char* sp = my + st * kRingStageBytes;
      const uint64_t* pl = reinterpret_cast<const uint64_t*>(sp) + lane;
      mk[0] = have_sec ? pl[0] : 1ULL;
#pragma unroll
      for (int p = 1; p < 6; ++p) mk[p] = pl[32 * p];
      t = (int64_t)pl[192];
      if (have_zone) zone = reinterpret_cast<const int32_t*>(sp + 7 * 256)[lane];
      __syncwarp();  // every lane has taken its operands out of the stage: lane 0 may refill it
      if (lane == 0 && tile + (size_t)kRingStages * W < tiles) fill(tile + (size_t)kRingStages * W, st);
      if (++st == kRingStages) { st = 0; parity ^= 1u; }
    } else {
      const size_t ii = i < n ? i : n - 1;
      mk[0] = have_sec ? __ldg(m.p[0] + ii) : 1ULL;
#pragma unroll
      for (int p = 1; p < 6; ++p) mk[p] = __ldg(m.p[p] + ii);
      t = __ldg(from + ii);
      if (have_zone) zone = __ldg(zone_id + ii);
    }
    const Spec s = cronwalk::make_spec(mk[0], mk[1], mk[2], mk[3], mk[4], mk[5]);
    Walk w;
    const bool ok = bind_checked(w, zt, zone, t, err);
    const int64_t r = ok ? cronwalk::next_from_t<false, kUtcOnly>(s, w, t, false) : kZero;
    if (!waited) { asm volatile("griddepcontrol.wait;" ::: "memory"); waited = true; }  // see next1_kernel
    if (i < n) out[i] = r;
  }
  if (!waited) asm volatile("griddepcontrol.wait;" ::: "memory");
}
#endif  // B200CRON_NEXT1_RING

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

// The unmet path runs as up to three kinds of launch (plans in launch_unmet, b200cron.cu):
//   unmet_tile_kernel   walks the Crons in input order, a tile of kUnmetTile per CTA.  Inside the tile the Crons are
//                       first dealt into two classes -- evaluated in UTC / evaluated in a zone -- with a ballot
//                       compaction in shared memory, so that warps are class-pure: UTC warps run unmet_step_utc (no zone
//                       logic, no step emulation, nothing to defer: ~5x fewer instructions), zoned warps run the easy
//                       rounds below.  On C4 half of the Crons are UTC and, mixed into every warp in input order, they
//                       used to ride through the whole zoned instruction stream.
//   unmet_ordered_kernel the easy rounds over an index list grouped by zone id (zone_order_* kernels): a warp's 32
//                       Crons then share their zone's periods and transition, and the zone-dependent code stops
//                       diverging.  Used for the zoned class when the plan says so (the tile kernel then only does UTC).
//   unmet_hard_kernel   the general state machine (with the step emulation) over the compact list of Crons the easy
//                       rounds gave up on, in dense warps.
// Easy rounds: only the closed form may answer, at most kUnmetEasyRounds rounds per Cron (first fire + fast-forward, a
// zone transition crossed + fast-forward, ..., end of chain, Next(now)); a Cron that needs more -- a step only the
// emulation can take, a backward step -- is appended (ballot + one atomic per warp) to the defer list.  On C4 five
// rounds defer 2.4 % of the Crons (all in zones whose transitions sit on local midnight).
enum { kUnmetSingle = 0, kUnmetEasy = 1, kUnmetHard = 2, kUnmetUtc = 3 };
#ifndef B200CRON_UNMET_EASY_ROUNDS
#define B200CRON_UNMET_EASY_ROUNDS 5
#endif
constexpr int kUnmetEasyRounds = B200CRON_UNMET_EASY_ROUNDS;
#ifndef B200CRON_UNMET_EASY_MINBLOCKS
#define B200CRON_UNMET_EASY_MINBLOCKS 8
#endif
#ifndef B200CRON_UNMET_TILE_MINBLOCKS
#define B200CRON_UNMET_TILE_MINBLOCKS 8  // the class-A tile kernel (UTC machine only)
#endif
#ifndef B200CRON_UNMET_ORDERED_SINGLE
#define B200CRON_UNMET_ORDERED_SINGLE 0
#endif
#ifndef B200CRON_UNMET_HARD_MINBLOCKS
#define B200CRON_UNMET_HARD_MINBLOCKS 12
#endif

// one Cron per lane; all 32 lanes of the warp must call this together (live = false for padding lanes)
// kUnmetUtc takes a clock shift: `shift` = 0 for a Cron evaluated in UTC, or the UTC offset of the ONE zone period that
// holds the Cron's whole window (class A of unmet_tile_kernel) -- inside a period local time is UTC moved by the offset,
// so the UTC machine runs on the local clock and the answers are moved back.  `nhi` bounds the period's natural window
// on that clock: a Next() beyond it (or one that found nothing in five years) left the period, the lane is then
// *rejected -- nothing is written -- and the caller hands the Cron to the general path.
// kFive: the launch has no seconds plane (5-field schedules): the seconds mask is the constant 1 and folds away.
template <int kMode, bool kFive = false>
__device__ __forceinline__ void unmet_one(const UnmetArgs& a, const ZoneTable& zt, int* __restrict__ err, size_t i,
                                          bool live, uint32_t* __restrict__ defer_idx,
                                          uint32_t* __restrict__ defer_count, int32_t shift = 0,
                                          int64_t nhi = cronwalk::kOmega, bool* rejected = nullptr) {
  const size_t ii = live ? i : 0;  // out-of-range lanes stay in the warp-uniform loop below, doing nothing
  // masks[0] == NULL: 5-field schedules, second 0 (what ParseStandard always yields)
  const Spec s = cronwalk::make_spec(kFive ? 1ULL : (a.m.p[0] ? __ldg(a.m.p[0] + ii) : 1ULL), __ldg(a.m.p[1] + ii),
                                     __ldg(a.m.p[2] + ii), __ldg(a.m.p[3] + ii), __ldg(a.m.p[4] + ii), __ldg(a.m.p[5] + ii));
  cronwalk::UnmetIn in;
  in.earliest = __ldg(a.earliest + ii);
  in.now = a.now;
  in.now_nsec = a.now_nsec;
  in.policy = a.policy ? __ldg(a.policy + ii) : 0;
  in.active = a.active ? __ldg(a.active + ii) : 0;
  in.flags = a.flags ? __ldg(a.flags + ii) : 0;
  in.deadline = a.deadline ? __ldg(a.deadline + ii) : kZero;
  const int32_t zone = (kMode != kUnmetUtc && a.zone_id) ? __ldg(a.zone_id + ii) : 0;

  bool ok = live;
  if (ok) {
    const int64_t lo = in.earliest < in.now ? in.earliest : in.now;
    if (lo < kMinTime || in.now > kMaxTime || in.earliest > kMaxTime) { atomicOr(err, kErrRange); ok = false; }
    if (kMode != kUnmetUtc && ok && zone != 0) {
      if (zone < 0 || zone >= zt.nzones) { atomicOr(err, kErrZone); ok = false; }
      else if (!s.every && __ldg(zt.zstart + zone + 1) - __ldg(zt.zstart + zone) > 1 && in.now >= zt.horizon - kSixYears) {
        atomicOr(err, kErrRange); ok = false;
      }
    }
  }
  cronwalk::UnmetState st;
  st.o.last = kZero; st.o.next = kZero; st.o.missed = 0; st.o.status = cronwalk::kStOk; st.o.decision = 0;
  st.done = true;
  st.w.bind_utc();
  if (ok) {
    if (s.every) {
      st.o = cronwalk::unmet_decide_every((int64_t)s.second, in);  // "@every <dur>": closed form, no walk
    } else if (kMode == kUnmetUtc) {
      cronwalk::unmet_begin_shifted(zt, in, shift, st);
    } else {
      cronwalk::unmet_begin(zt, zone, in, st);
    }
  }
  // all 32 lanes go round together: a full-mask vote per iteration is the reconvergence point
  bool defer = false;
  if (kMode == kUnmetUtc) {
    while (__any_sync(0xffffffffu, !st.done)) {
      if (!st.done) cronwalk::unmet_step_utc(s, in, st);
    }
    if (ok && !s.every) defer = cronwalk::unmet_finish_shifted(shift, nhi, st);  // did every call stay in the period?
    if (rejected) *rejected = defer;
  } else if (kMode == kUnmetEasy) {
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

// the whole job in one launch with the general state machine: only for batches beyond 2^32 Crons (the index lists of
// the other kernels are 32-bit)
__global__ void __launch_bounds__(kUnmetThreads, 12)
unmet_single_kernel(UnmetArgs a, ZoneTable zt, int* __restrict__ err) {
  const size_t q = (size_t)blockIdx.x * kUnmetThreads + threadIdx.x;
  unmet_one<kUnmetSingle>(a, zt, err, q, q < a.n, nullptr, nullptr);
}

// Input-order tiles with class-pure warps (see above).  Class A = Crons the UTC machine can answer: evaluated in UTC,
// or in a zone whose single period [ps, pe) holds the Cron's whole window -- from 32 days before min(earliest, now)
// (which covers every Date() reset target: the midnight of the day or of the 1st of the month) to now; the machine then
// runs on that period's local clock (unmet_one's `shift`) and a lane whose Next() leaves the period is rejected into
// class B.  Class B = the zoned Crons with a transition in their window: kZoned handles them here (easy rounds);
// otherwise they are appended to `z1_list` for the zone-ordered kernel.  A launch without zone ids is all class A: no
// classification, no lists.
constexpr int kUnmetTile = 512;  // Crons per CTA: ~12 + 4 warps' worth on a half-UTC workload
template <bool kZoned, bool kFive>
__global__ void __launch_bounds__(kUnmetThreads, kZoned ? B200CRON_UNMET_EASY_MINBLOCKS : B200CRON_UNMET_TILE_MINBLOCKS)
unmet_tile_kernel(UnmetArgs a, ZoneTable zt, int* __restrict__ err, uint32_t* __restrict__ defer_idx,
                  uint32_t* __restrict__ defer_count, uint32_t* __restrict__ z1_list, uint32_t* __restrict__ z1_count) {
  const size_t base = (size_t)blockIdx.x * kUnmetTile;
  const int cnt = (int)((a.n - base) < (size_t)kUnmetTile ? (a.n - base) : (size_t)kUnmetTile);
  // class A indices packed from the front of s_idx, class B from the back; a class-A entry keeps its clock shift and
  // the end of its period's natural window on the local clock (zone-less launch: identity order, no shift)
  __shared__ uint16_t s_idx[kUnmetTile];
  __shared__ int32_t s_shift[kUnmetTile];
  __shared__ int64_t s_nhi[kUnmetTile];
  __shared__ uint16_t s_rej[kUnmetTile];  // class-A entries whose Next() left their period: they join class B
  __shared__ int s_n[3];
  if (threadIdx.x < 3) s_n[threadIdx.x] = threadIdx.x == 0 && a.zone_id == nullptr ? cnt : 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31u;
  const unsigned below = (1u << lane) - 1u;
#pragma unroll 1
  for (int it = 0; it < kUnmetTile / kUnmetThreads; ++it) {
    if (a.zone_id == nullptr) break;
    const int q = it * kUnmetThreads + threadIdx.x;
    const bool in = q < cnt;
    int32_t shift = 0;
    int64_t nhi = cronwalk::kOmega;
    bool cls_b = false;
    if (in) {
      const int32_t z = __ldg(a.zone_id + base + q);
      // invalid ids are class B too: the general path reports them
      if (z != 0) cls_b = !cronwalk::unmet_class_a(zt, z, __ldg(a.earliest + base + q), a.now, kMinTime, kMaxTime,
                                                   zt.horizon - kSixYears, &shift, &nhi);
    }
    const unsigned mb = __ballot_sync(0xffffffffu, cls_b), ma = __ballot_sync(0xffffffffu, in && !cls_b);
    int ba = 0, bb = 0;
    if (lane == 0) {
      if (ma) ba = atomicAdd(&s_n[0], __popc(ma));
      if (mb) bb = atomicAdd(&s_n[1], __popc(mb));
    }
    ba = __shfl_sync(0xffffffffu, ba, 0);
    bb = __shfl_sync(0xffffffffu, bb, 0);
    if (in && !cls_b) {
      const int slot = ba + __popc(ma & below);
      s_idx[slot] = (uint16_t)q; s_shift[slot] = shift; s_nhi[slot] = nhi;
    }
    if (cls_b) s_idx[kUnmetTile - 1 - (bb + __popc(mb & below))] = (uint16_t)q;
  }
  __syncthreads();
  const int na = s_n[0];
  for (int q = threadIdx.x; q < ((na + 31) & ~31); q += kUnmetThreads) {
    const bool live = q < na;
    const bool plain = a.zone_id == nullptr;
    const int li = live ? (plain ? q : (int)s_idx[q]) : 0;
    bool rejected = false;
    unmet_one<kUnmetUtc, kFive>(a, zt, err, base + li, live, nullptr, nullptr, (live && !plain) ? s_shift[q] : 0,
                         (live && !plain) ? s_nhi[q] : cronwalk::kOmega, &rejected);
    const unsigned mr = __ballot_sync(0xffffffffu, rejected);
    if (mr) {
      int br = 0;
      if (lane == 0) br = atomicAdd(&s_n[2], __popc(mr));
      br = __shfl_sync(0xffffffffu, br, 0);
      if (rejected) s_rej[br + __popc(mr & below)] = (uint16_t)li;
    }
  }
  if (a.zone_id == nullptr) return;
  __syncthreads();
  const int nb = s_n[1], nr = s_n[2];
  auto class_b = [&](int q) { return q < nb ? s_idx[kUnmetTile - 1 - q] : s_rej[q - nb]; };
  if (kZoned) {
    for (int q = threadIdx.x; q < ((nb + nr + 31) & ~31); q += kUnmetThreads) {
      const bool live = q < nb + nr;
      unmet_one<kUnmetEasy>(a, zt, err, base + (live ? class_b(q) : 0), live, defer_idx, defer_count);
    }
  } else if (nb + nr > 0) {
    __shared__ uint32_t s_base;
    if (threadIdx.x == 0) s_base = atomicAdd(z1_count, (uint32_t)(nb + nr));
    __syncthreads();
    for (int q = threadIdx.x; q < nb + nr; q += kUnmetThreads) z1_list[s_base + q] = (uint32_t)(base + class_b(q));
  }
}

// the easy rounds over the zone-ordered index list (count = its length, known only on the device)
__global__ void __launch_bounds__(kUnmetThreads, B200CRON_UNMET_EASY_MINBLOCKS)
unmet_ordered_kernel(UnmetArgs a, ZoneTable zt, int* __restrict__ err, uint32_t* __restrict__ defer_idx,
                     uint32_t* __restrict__ defer_count, const uint32_t* __restrict__ order,
                     const uint32_t* __restrict__ order_count) {
  const uint32_t cnt = __ldg(order_count);
  const size_t q0 = (size_t)blockIdx.x * kUnmetThreads + (threadIdx.x & ~31u);
  if (q0 >= cnt) return;  // warp-uniform
  const size_t q = q0 + (threadIdx.x & 31u);
  const bool live = q < cnt;
#if B200CRON_UNMET_ORDERED_SINGLE
  // tuning build: the general state machine right here (step emulation out of line), no defer list, no hard kernel --
  // zone-pure warps either need the emulation in every lane or in none
  unmet_one<kUnmetSingle>(a, zt, err, live ? (size_t)__ldg(order + q) : 0, live, nullptr, nullptr);
#else
  unmet_one<kUnmetEasy>(a, zt, err, live ? (size_t)__ldg(order + q) : 0, live, defer_idx, defer_count);
#endif
}

// the general state machine over the compact list the easy rounds left behind (grid-stride: the host does not know
// the list's length when it launches)
__global__ void __launch_bounds__(kUnmetThreads, B200CRON_UNMET_HARD_MINBLOCKS)
unmet_hard_kernel(UnmetArgs a, ZoneTable zt, int* __restrict__ err, const uint32_t* __restrict__ defer_idx,
                  const uint32_t* __restrict__ defer_count) {
  // The list is dealt out in equal contiguous shares over all warps of the grid (the host does not know its length
  // when it launches), so a list a little longer than the grid costs every warp two more lanes instead of a few warps a
  // second full round.  Measured neutral on C4 (2.73 vs 2.74 ms): the kernel is bound by instruction fetch in the
  // step emulation (ncu: 6.8 of 32 lanes active, 15 % issue-active), not by how its lanes are spread.
  const uint32_t cnt = *defer_count;
  const uint32_t warps = gridDim.x * (kUnmetThreads / 32);
  const uint32_t share = (cnt + warps - 1) / warps;
  const uint32_t gw = blockIdx.x * (kUnmetThreads / 32) + (threadIdx.x >> 5);
  const uint64_t lo = (uint64_t)gw * share;
  const uint64_t hi = lo + share < cnt ? lo + share : cnt;
  for (uint32_t r = 0; r < share; r += 32) {  // the same trip count in every lane of the warp
    const uint64_t q = lo + r + (threadIdx.x & 31u);
    const bool live = q < hi;
    if (!__any_sync(0xffffffffu, live)) break;
    unmet_one<kUnmetHard>(a, zt, err, live ? defer_idx[q] : 0, live, nullptr, nullptr);
  }
}

// -------------------------------------------------------------------------------------------------------------
// Zone order: a counting sort by zone id of the Crons in `list` (the class-B Crons unmet_tile_kernel collected; its
// length *count is known only on the device, so the grids cover n and CTAs past the end leave at once): histogram,
// scan, scatter -- one pass over the list each.  Ids outside [0, nzones) are binned with zone 1 and reported by
// unmet_one as before.  cursor[nzones] receives the list's length.
// -------------------------------------------------------------------------------------------------------------
constexpr int kZoneOrderThreads = 256;
constexpr int kZoneOrderItems = 16;  // Crons per thread: a CTA bins 4096 consecutive list entries
constexpr int kZoneOrderTile = kZoneOrderThreads * kZoneOrderItems;

__device__ __forceinline__ uint32_t zone_bin(const int32_t* __restrict__ zone_id, uint32_t i, int32_t nzones) {
  const int32_t z = __ldg(zone_id + i);
  return (z <= 0 || z >= nzones) ? 1u : (uint32_t)z;
}

// the CTA's histogram of its tile of the list in shared memory (bins [0, nzones)); all threads call it
__device__ __forceinline__ void zone_tile_hist(const int32_t* __restrict__ zone_id, const uint32_t* __restrict__ list,
                                               uint32_t count, int32_t nzones, uint32_t* __restrict__ sh) {
  for (int b = threadIdx.x; b < nzones; b += kZoneOrderThreads) sh[b] = 0;
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * kZoneOrderTile;
#pragma unroll 4
  for (int it = 0; it < kZoneOrderItems; ++it) {
    const size_t q = base + (size_t)it * kZoneOrderThreads + threadIdx.x;
    if (q < count) atomicAdd(&sh[zone_bin(zone_id, __ldg(list + q), nzones)], 1u);
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kZoneOrderThreads)
zone_order_hist_kernel(const int32_t* __restrict__ zone_id, const uint32_t* __restrict__ list,
                       const uint32_t* __restrict__ count, int32_t nzones, uint32_t* __restrict__ hist) {
  extern __shared__ uint32_t zo_sh[];
  const uint32_t cnt = __ldg(count);
  if ((size_t)blockIdx.x * kZoneOrderTile >= cnt) return;
  zone_tile_hist(zone_id, list, cnt, nzones, zo_sh);
  for (int b = threadIdx.x; b < nzones; b += kZoneOrderThreads)
    if (zo_sh[b]) atomicAdd(hist + b, zo_sh[b]);
}

// exclusive prefix sum of the histogram into `cursor` (one CTA; a thread owns a contiguous run of bins);
// cursor[nzones] = total
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
    cursor[nzones] = run;
  }
  __syncthreads();
  uint32_t run = part[threadIdx.x];
  for (int b = lo; b < hi; ++b) { cursor[b] = run; run += hist[b]; }
}

// every CTA reserves, per zone, a range of the output for its tile's Crons and fills it
__global__ void __launch_bounds__(kZoneOrderThreads)
zone_order_scatter_kernel(const int32_t* __restrict__ zone_id, const uint32_t* __restrict__ list,
                          const uint32_t* __restrict__ count, int32_t nzones, uint32_t* __restrict__ cursor,
                          uint32_t* __restrict__ order) {
  extern __shared__ uint32_t zo_sh[];  // [nzones] counts, then local cursors; [nzones] bases
  uint32_t* sh_base = zo_sh + nzones;
  const uint32_t cnt = __ldg(count);
  if ((size_t)blockIdx.x * kZoneOrderTile >= cnt) return;
  zone_tile_hist(zone_id, list, cnt, nzones, zo_sh);
  for (int b = threadIdx.x; b < nzones; b += kZoneOrderThreads) {
    const uint32_t c = zo_sh[b];
    sh_base[b] = c ? atomicAdd(cursor + b, c) : 0u;
    zo_sh[b] = 0;
  }
  __syncthreads();
  const size_t base = (size_t)blockIdx.x * kZoneOrderTile;
#pragma unroll 4
  for (int it = 0; it < kZoneOrderItems; ++it) {
    const size_t q = base + (size_t)it * kZoneOrderThreads + threadIdx.x;
    if (q < cnt) {
      const uint32_t i = __ldg(list + q);
      const uint32_t b = zone_bin(zone_id, i, nzones);
      order[sh_base[b] + atomicAdd(&zo_sh[b], 1u)] = i;
    }
  }
}

// -------------------------------------------------------------------------------------------------------------
// history_prune: stable ascending order of each Cron's finished workloads by creation time + keep/delete flags
//                                                           (cron_util.go:117-129, cron_controller.go:313-323)
//
// A CTA owns kHistSegs consecutive segments, i.e. one contiguous run of records, and works through it in chunks of
// whole segments that fit kHistCap records (sized by records, not by segments: a tile of long histories simply takes
// more chunks).  A chunk's keys are brought into shared memory with coalesced loads as 32-bit differences to the
// chunk's first key (creation timestamps of one Cron's workloads lie within days of each other; a chunk whose keys
// span more than 2^31 keeps them as 64-bit values), every record finds its segment through a one-byte stamp and its
// stable rank by counting the records of its segment that sort before it (segments are short: the reference keeps
// historyLimit of them), and perm/keep go straight to global memory (a warp's lanes hold consecutive records, so a
// permutation inside a segment touches the same 32-byte sectors as the identity).
// A single segment longer than kHistCap is ranked by the whole CTA straight from global memory -- quadratic, so
// bounded: above kHistMaxSeg records the launch flags kErrSegLen (B200CRON_EINVAL) and leaves that segment in input
// order with every record kept.
// Algorithmic bytes: 8 (key) in, 4 (perm) + 1 (keep) out per record; 8 (offset) + 8 (limit) in per segment.
// -------------------------------------------------------------------------------------------------------------
#ifndef B200CRON_HIST_PACK
#define B200CRON_HIST_PACK 1  // 0 (tuning builds): never use the packed key+position form of the rank loop
#endif
constexpr int kHistThreads = 256;
constexpr int kHistSegs = 256;
constexpr int kHistCap = 2048;          // records per chunk held in shared memory
constexpr uint64_t kHistMaxSeg = 65536;  // longest segment the quadratic path will rank

__global__ void __launch_bounds__(kHistThreads)
history_prune_kernel(const uint64_t* __restrict__ seg_off, const int64_t* __restrict__ key,
                     const int64_t* __restrict__ limit, size_t nseg, uint32_t* __restrict__ perm,
                     uint8_t* __restrict__ keep, int* __restrict__ err) {
  __shared__ uint64_t s_off[kHistSegs + 1];
  __shared__ __align__(8) unsigned char s_keys[kHistCap * 8];  // int32 differences, or int64 keys for a wide chunk
  __shared__ uint8_t s_seg[kHistCap];
  __shared__ int32_t s_lim[kHistSegs];  // HistoryLimit clamped to 31 bits; nil (negative) = keep everything
  const size_t seg0 = (size_t)blockIdx.x * kHistSegs;
  const int nsegs = (int)((nseg - seg0) < (size_t)kHistSegs ? (nseg - seg0) : (size_t)kHistSegs);
  for (int x = threadIdx.x; x <= nsegs; x += kHistThreads) s_off[x] = __ldg(seg_off + seg0 + x);
  for (int x = threadIdx.x; x < nsegs; x += kHistThreads) {
    const int64_t lim = limit ? __ldg(limit + seg0 + x) : -1;
    s_lim[x] = (lim < 0 || lim > 0x7fffffff) ? 0x7fffffff : (int32_t)lim;
  }
  __syncthreads();
  int32_t* k32 = reinterpret_cast<int32_t*>(s_keys);
  int64_t* k64 = reinterpret_cast<int64_t*>(s_keys);
  int sa = 0;
  while (sa < nsegs) {  // every thread computes the same chunk boundaries from shared memory
    const uint64_t r0 = s_off[sa];
    int lo = sa, hi = nsegs + 1;  // largest sb in (sa, nsegs] with s_off[sb] - r0 <= kHistCap, at least sa + 1
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_off[mid] - r0 <= (uint64_t)kHistCap) lo = mid; else hi = mid;
    }
    if (lo == sa) {
      // ---- one segment longer than a chunk: ranked by the whole CTA from global memory ----
      const uint64_t len = s_off[sa + 1] - r0;
      const int64_t lim = limit ? __ldg(limit + seg0 + sa) : -1;
      if (len > kHistMaxSeg) {
        if (threadIdx.x == 0) atomicOr(err, kErrSegLen);
        for (uint64_t x = threadIdx.x; x < len; x += kHistThreads) { perm[r0 + x] = (uint32_t)x; keep[r0 + x] = 1; }
      } else {
        for (uint64_t x = threadIdx.x; x < len; x += kHistThreads) {
          const int64_t kx = __ldg(key + r0 + x);
          uint32_t rank = 0;
          for (uint64_t j = 0; j < len; ++j) {
            const int64_t kj = __ldg(key + r0 + j);
            rank += (kj < kx) || (kj == kx && j < x);
          }
          perm[r0 + rank] = (uint32_t)x;
          keep[r0 + rank] = (lim < 0) ? 1 : !((int64_t)rank < (int64_t)len - lim);
        }
      }
      sa += 1;
      continue;
    }
    const int sb = lo;
    const uint32_t tot = (uint32_t)(s_off[sb] - r0);
    // ---- load the chunk: keys as 32-bit differences when they all fit, segment stamps ----
    __syncthreads();  // the previous chunk's readers are done with s_keys / s_seg
    const int64_t base = tot ? __ldg(key + r0) : 0;
    bool wide = false, roomy = false;  // roomy: this thread saw something the packed form below cannot hold
    for (uint32_t x = threadIdx.x; x < tot; x += kHistThreads) {
      const int64_t d = __ldg(key + r0 + x) - base;  // wraps for absurd spans; caught by the check on the stored value
      k32[x] = (int32_t)d;
      wide = wide || d != (int64_t)(int32_t)d || (__ldg(key + r0 + x) < base) != (d < 0);
      roomy = roomy || d >= (1 << 26) || d < -(1 << 26);
    }
    for (int sg = sa + threadIdx.x; sg < sb; sg += kHistThreads) {
      const uint32_t a0 = (uint32_t)(s_off[sg] - r0), b0 = (uint32_t)(s_off[sg + 1] - r0);
      for (uint32_t j = a0; j < b0; ++j) s_seg[j] = (uint8_t)(sg - sa);
      roomy = roomy || b0 - a0 > 32;
    }
    const bool any_wide = __syncthreads_or(wide);
    // Packed form (B200CRON_HIST_PACK): when every key of the chunk lies within 2^26 of the first one (two years of
    // seconds) and no segment has more than 32 records -- the reference's shape: creation times of one Cron's last few
    // workloads -- key difference and position inside the segment share one 32-bit word, which orders records exactly as
    // the stable sort does: the rank loop is then one load, one compare and one add per record pair.
    const bool packed = B200CRON_HIST_PACK && !__syncthreads_or(roomy) && !any_wide;
    if (any_wide) {  // keys too far apart for 32 bits: reload them whole
      for (uint32_t x = threadIdx.x; x < tot; x += kHistThreads) k64[x] = __ldg(key + r0 + x);
      __syncthreads();
    } else if (packed) {
      for (uint32_t x = threadIdx.x; x < tot; x += kHistThreads) {
        const int sg = sa + s_seg[x];
        k32[x] = k32[x] * 32 + (int32_t)(x - (uint32_t)(s_off[sg] - r0));
      }
      __syncthreads();
    }
    for (uint32_t x = threadIdx.x; x < tot; x += kHistThreads) {
      const int sg = sa + s_seg[x];
      const uint32_t a = (uint32_t)(s_off[sg] - r0), b = (uint32_t)(s_off[sg + 1] - r0);
      uint32_t rank = 0;
      if (packed) {
        const int32_t cx = k32[x];
#pragma unroll 4
        for (uint32_t j = a; j < b; ++j) rank += k32[j] < cx;
      } else if (!any_wide) {
        const int32_t kx = k32[x];
#pragma unroll 4
        for (uint32_t j = a; j < b; ++j) {
          const int32_t kj = k32[j];
          rank += (kj < kx) | ((kj == kx) & (j < x));
        }
      } else {
        const int64_t kx = k64[x];
        for (uint32_t j = a; j < b; ++j) {
          const int64_t kj = k64[j];
          rank += (kj < kx) || (kj == kx && j < x);
        }
      }
      // sorted position `rank` is deleted when rank < len - limit
      perm[r0 + a + rank] = x - a;
      keep[r0 + a + rank] = !((int64_t)rank < (int64_t)(b - a) - (int64_t)s_lim[sg]);
    }
    sa = sb;
  }
}

// -------------------------------------------------------------------------------------------------------------
// unmet list: the first min(missed, cap) missed fire times of every Cron, packed back to back
//                                   (the enumeration of cron_controller.go:409-430, which upstream only counts)
//
// list_off = exclusive prefix sum of the per-Cron counts (three small kernels: per-tile sums into a scratch array,
// a single-CTA scan of the tile sums, per-tile offsets = warp-shuffle scan + tile base); unmet_list_emit_kernel then
// re-walks each chain and writes its fire times at list + list_off[i] (through a per-warp tile, see below).  Gated / unschedulable Crons contribute 0.
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

// Warp-cooperative emission: a warp owns 32 consecutive Crons, whose lists are one contiguous run of the output.  The
// run is produced window by window: every lane walks its own chain and drops the fire times that fall into the current
// window into the warp's shared-memory tile (at their final position), then the warp writes the window out with
// lane-consecutive stores -- full 256-byte rows instead of 32 separate 8-byte streams 800 bytes apart.
constexpr int kEmitWindow = 512;  // list entries per window (4 KB per warp)
__global__ void __launch_bounds__(kUnmetThreads)
unmet_list_emit_kernel(MaskPlanes m, const int32_t* __restrict__ zone_id, const int64_t* __restrict__ earliest,
                       size_t n, const uint64_t* __restrict__ list_off, int64_t* __restrict__ list, ZoneTable zt) {
  __shared__ int64_t s_tile[kUnmetThreads / 32][kEmitWindow];
  const unsigned lane = threadIdx.x & 31u;
  int64_t* tile = s_tile[threadIdx.x >> 5];
  const size_t i0 = (size_t)blockIdx.x * kUnmetThreads + (threadIdx.x & ~31u);
  if (i0 >= n) return;  // warp-uniform
  const size_t i = i0 + lane;
  const bool live = i < n;
  const size_t ii = live ? i : n - 1;
  const uint64_t a = live ? __ldg(list_off + ii) : 0, b = live ? __ldg(list_off + ii + 1) : 0;
  const size_t ilast = (i0 + 31 < n) ? i0 + 31 : n - 1;
  const uint64_t run_lo = __ldg(list_off + i0), run_hi = __ldg(list_off + ilast + 1);
  if (run_hi == run_lo) return;  // warp-uniform: nothing listed for these 32 Crons
  const Spec s = cronwalk::make_spec(m.p[0] ? __ldg(m.p[0] + ii) : 1ULL, __ldg(m.p[1] + ii), __ldg(m.p[2] + ii),
                                     __ldg(m.p[3] + ii), __ldg(m.p[4] + ii), __ldg(m.p[5] + ii));
  const int32_t zone = zone_id ? __ldg(zone_id + ii) : 0;
  Walk w;
  if (zone == 0) w.bind_utc(); else w.bind_zone(zt, zone);  // ids and ranges were validated by the unmet kernels
  int64_t t = __ldg(earliest + ii);
  bool resident = false;
  uint64_t j = a;
  for (uint64_t win = run_lo; win < run_hi; win += kEmitWindow) {
    const uint64_t wend = (run_hi - win) < (uint64_t)kEmitWindow ? run_hi : win + kEmitWindow;
    for (;;) {  // one chained call per round; the vote is the reconvergence point
      const bool go = j < b && j < wend;
      if (!__any_sync(0xffffffffu, go)) break;
      if (go) {
        t = cronwalk::next_from_t<true>(s, w, t, resident);
        resident = true;
        tile[j - win] = t;
        ++j;
      }
    }
    __syncwarp();
    for (uint64_t x = lane; x < wend - win; x += 32) list[win + x] = tile[x];
    __syncwarp();
  }
}

}  // namespace b200cron
