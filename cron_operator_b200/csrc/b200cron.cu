// b200cron.cu -- context, host<->device plumbing and the extern "C" entry points of libb200cron.so (include/b200cron.h).
//
// Host-pointer calls shard `n` contiguously over the context's devices (no collective: every spec is independent,
// SURVEY.md section 8e) and, per device, pipeline chunks over kSlots streams so that H2D copies, the kernel and D2H
// copies of neighbouring chunks overlap.  `_dev` calls only enqueue a kernel on the caller's stream.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <type_traits>
#include <vector>

#include "../../include/b200cron.h"
#include "kernels.cuh"
#include "parse_kernels.cuh"
#include "synth_kernels.cuh"
#include "zone_table.h"

namespace {

using namespace b200cron;

constexpr int kSlots = 3;
constexpr size_t kChunk = 1u << 18;       // specs per pipelined chunk of the host-pointer next_n calls
constexpr size_t kUnmetChunk = 1u << 20;  // Crons per chunk of the host-pointer unmet calls (six launches per chunk)

// Error text of the calling thread's last failed call: entry points validate their arguments before they take the
// context's mutex, and callers on different threads (cgo runs Reconcile on up to MaxConcurrentReconciles OS threads)
// must not see each other's messages.
thread_local std::string t_error;

struct Slot {
  cudaStream_t stream = nullptr;
  void* buf = nullptr;  // device staging
  size_t cap = 0;
};

// Compact index list through which the easy unmet kernel hands Crons to the general one.  `done` is recorded after
// the pair that used the list; the next pair to use it waits on that event first, so calls on different streams
// never share a list that is still in flight.
struct DeferList {
  uint32_t* idx = nullptr;
  uint32_t* count = nullptr;
  size_t cap = 0;  // entries
  cudaEvent_t done = nullptr;
};
constexpr int kDeferLists = 3;

struct Device {
  int id = 0;
  int sm_count = 148;
  cudaStream_t stream = nullptr;  // default stream of the _dev entry points
  int64_t* d_when = nullptr;
  int32_t* d_off = nullptr;
  int32_t* d_fpre = nullptr;
  int32_t* d_zstart = nullptr;
  uint16_t* d_yidx = nullptr;
  int* d_err = nullptr;
  int* h_err = nullptr;  // pinned mirror
  Slot slots[kSlots];
  uint64_t* d_scratch = nullptr;  // per-tile sums of the unmet-list prefix sum
  size_t scratch_cap = 0;         // entries
  DeferList defer[kDeferLists];       // scratch of the unmet kernel pair, handed out round-robin (see launch_unmet)
  int defer_next = 0;
  int defer_last = -1;  // the list the most recent unmet launch used (b200cron_debug_last_deferred)
  uint64_t* d_static_sums = nullptr;  // device address of g_static_tile_sums on this device (resolved on first use)
  ZoneTable zt{};
};

}  // namespace

struct b200cron_ctx {
  std::vector<Device> devs;
  HostZoneTable zones;
  std::mutex mu;
  std::atomic<uint64_t> launches{0};
  std::atomic<int> overlap{0};  // b200cron_ctx_set_overlap
};

namespace {

int fail(b200cron_ctx* ctx, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  (void)ctx;
  t_error = buf;
  return code;
}

#define CU(ctx, call)                                                                                  \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess)                                                                             \
      return fail(ctx, B200CRON_ECUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

int ensure_slot(b200cron_ctx* ctx, Slot& s, size_t bytes) {
  if (bytes <= s.cap) return 0;
  if (s.buf) CU(ctx, cudaFree(s.buf));
  s.buf = nullptr;
  s.cap = 0;
  CU(ctx, cudaMalloc(&s.buf, bytes));
  s.cap = bytes;
  return 0;
}

// Pinned (page-locked, mapped) host memory?  *dev receives the address kernels on the CURRENT device must use for
// it: equal to the host address under unified addressing, but not for cudaHostRegister'd memory on systems where
// canUseHostPointerForRegisteredMem is 0.
bool pinned_alias(const void* p, const void** dev) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  if (a.type != cudaMemoryTypeHost || a.devicePointer == nullptr) return false;
  *dev = a.devicePointer;
  return true;
}
bool is_pinned_host(const void* p) {
  const void* d;
  return pinned_alias(p, &d);
}
template <typename T>
const T* dev_alias(const T* p) {  // p is known to be pinned (is_pinned_host) or NULL
  const void* d = p;
  if (p) pinned_alias(p, &d);
  return static_cast<const T*>(d);
}
template <typename T>
T* dev_alias(T* p) { return const_cast<T*>(dev_alias(static_cast<const T*>(p))); }

[[maybe_unused]] bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// B200CRON_NEXTN_SPLIT (kernels.cuh): 1 = k > 1 launches walk class-pure warps (next_n_tile_kernel, one launch per
// class); 0 (tuning builds) = one launch of next_n_kernel over the specs in input order.

// ---- launches -------------------------------------------------------------------------------------------------
#if !B200CRON_NEXTN_SPLIT
template <int KT>
cudaError_t launch_next_kt(bool vec, unsigned grid, size_t smem, cudaStream_t st, const MaskPlanes& m,
                           const int32_t* zone, const int64_t* from, size_t n, int k, int64_t* out,
                           const ZoneTable& zt, int* err) {
  if (vec) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(next_n_kernel<KT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    next_n_kernel<KT, true><<<grid, kNextThreads, smem, st>>>(m, zone, from, n, k, out, zt, err);
  } else {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(next_n_kernel<KT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    next_n_kernel<KT, false><<<grid, kNextThreads, smem, st>>>(m, zone, from, n, k, out, zt, err);
  }
  return cudaGetLastError();
}
#endif

#ifndef B200CRON_NEXT1_FIVE
#define B200CRON_NEXT1_FIVE 1  // 0 (tuning builds): launches without a seconds plane use the general instantiation
#endif
// B200CRON_NEXT1_RING (tuning builds): 1 = k == 1 launches on device-resident inputs use the persistent kernel with the
// asynchronous input ring (next1_ring_kernel); 0 = always next1_kernel.  Measured on B200 (profiles/README.md, r2c):
// the ring is bit-exact through the whole GPU suite but SLOWER -- 18.8 vs 13.7 us per 1M specs with overlapping
// launches, 24.6 vs 18.5 us without -- because next1 is bound by instruction issue, not by load latency, and the
// producer lane's bookkeeping (expect_tx + seven bulk copies per 32-spec tile, mbarrier waits) costs issue slots that
// the plain kernel's seven LDGs do not.  So it is off, and compiled only into tuning builds.
#ifndef B200CRON_NEXT1_RING
#define B200CRON_NEXT1_RING 0
#endif

// device_resident: every buffer lives in this device's memory (the bulk-copy ring reads this GPU's memory only;
// pinned host buffers used in place take the plain-load kernel)
cudaError_t launch_next(b200cron_ctx* ctx, const Device& d, cudaStream_t st, const MaskPlanes& m, const int32_t* zone,
                        const int64_t* from, size_t n, int k, int64_t* out, bool device_resident) {
  if (n == 0) return cudaSuccess;
#if B200CRON_NEXT1_RING || !B200CRON_NEXTN_SPLIT  // only the tuning-build kernels use 128-bit accesses
  bool vec = aligned16(from) && aligned16(out) && (!zone || (reinterpret_cast<uintptr_t>(zone) & 7u) == 0);
  for (int p = 0; p < 6; ++p) vec = vec && (m.p[p] == nullptr || aligned16(m.p[p]));
#endif
#if !B200CRON_NEXTN_SPLIT
  const unsigned grid = (unsigned)((n + kNextTile - 1) / kNextTile);
  const size_t smem = (k == 1 || !kNextStage) ? 0 : (size_t)kNextTile * k * sizeof(int64_t);
#endif
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  if (k == 1) {  // one spec per lane (finer waves than next_n_kernel's two)
    if (n > 0xffffff00ULL) return cudaErrorInvalidValue;  // next1_kernel indexes with 32 bits (4G specs = 240 GB of inputs)
    const unsigned g1 = (unsigned)((n + kNext1Threads - 1) / kNext1Threads);
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(g1);
    cfg.blockDim = dim3(kNext1Threads);
    cfg.stream = st;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr.val.programmaticStreamSerializationAllowed = 1;
    if (ctx->overlap.load(std::memory_order_relaxed)) { cfg.attrs = &attr; cfg.numAttrs = 1; }
    int* err = d.d_err;
#if B200CRON_NEXT1_RING
    if (device_resident && vec && (!zone || aligned16(zone)) && n >= 32) {
      const size_t cta_tiles = (n / 32 + kRingWarps - 1) / kRingWarps;
      const size_t resident = (size_t)d.sm_count * B200CRON_RING_MINBLOCKS;
      cfg.gridDim = dim3((unsigned)(cta_tiles < resident ? cta_tiles : resident));
      cfg.blockDim = dim3(kRingThreads);
      cfg.dynamicSmemBytes = kRingSmemBytes;
      if (zone) return cudaLaunchKernelEx(&cfg, next1_ring_kernel<false>, m, zone, from, n, out, d.zt, err);
      return cudaLaunchKernelEx(&cfg, next1_ring_kernel<true>, m, zone, from, n, out, d.zt, err);
    }
#else
    (void)device_resident;
#endif
    const bool five = B200CRON_NEXT1_FIVE && m.p[0] == nullptr;  // no seconds plane: the 5-field instantiation
    if (zone) {
      if (five) return cudaLaunchKernelEx(&cfg, next1_kernel<false, true>, m, zone, from, n, out, d.zt, err);
      return cudaLaunchKernelEx(&cfg, next1_kernel<false, false>, m, zone, from, n, out, d.zt, err);
    }
    if (five) return cudaLaunchKernelEx(&cfg, next1_kernel<true, true>, m, zone, from, n, out, d.zt, err);  // all-UTC launch
    return cudaLaunchKernelEx(&cfg, next1_kernel<true, false>, m, zone, from, n, out, d.zt, err);
  }
#if B200CRON_NEXTN_SPLIT
  {  // k > 1: class-pure warps, one launch per class (kernels.cuh); a launch without zone ids is all UTC
    const unsigned tgrid = (unsigned)((n + kNextTileSpecs - 1) / kNextTileSpecs);
    auto go = [&](auto kt) {
      constexpr int KT = decltype(kt)::value;
      next_n_tile_kernel<KT, true><<<tgrid, kNextTileThreads, 0, st>>>(m, zone, from, n, k, out, d.zt, d.d_err);
      if (zone) {
        ctx->launches.fetch_add(1, std::memory_order_relaxed);
        const unsigned zgrid = (unsigned)((n + kNextTileSpecsZoned - 1) / kNextTileSpecsZoned);
        next_n_tile_kernel<KT, false><<<zgrid, kNextTileThreads, 0, st>>>(m, zone, from, n, k, out, d.zt, d.d_err);
      }
      return cudaGetLastError();
    };
    if (k == 16) return go(std::integral_constant<int, 16>{});
    return go(std::integral_constant<int, 0>{});
  }
#else
  switch (k) {
    case 2: return launch_next_kt<2>(vec, grid, smem, st, m, zone, from, n, k, out, d.zt, d.d_err);
    case 4: return launch_next_kt<4>(vec, grid, smem, st, m, zone, from, n, k, out, d.zt, d.d_err);
    case 8: return launch_next_kt<8>(vec, grid, smem, st, m, zone, from, n, k, out, d.zt, d.d_err);
    case 16: return launch_next_kt<16>(vec, grid, smem, st, m, zone, from, n, k, out, d.zt, d.d_err);
    case 32: return launch_next_kt<32>(vec, grid, smem, st, m, zone, from, n, k, out, d.zt, d.d_err);
    default: return launch_next_kt<0>(vec, grid, smem, st, m, zone, from, n, k, out, d.zt, d.d_err);
  }
#endif
}

// B200CRON_UNMET_PLAN (tuning builds, profiles/build_variants.sh): 0 = the tile kernel handles both classes in input
// order; 1 = the tile kernel handles the UTC class, the zoned class goes through a zone-ordered index list.
#ifndef B200CRON_UNMET_PLAN
#define B200CRON_UNMET_PLAN 1
#endif

cudaError_t launch_unmet(b200cron_ctx* ctx, Device& d, cudaStream_t st, const UnmetArgs& a) {
  if (a.n == 0) return cudaSuccess;
  if (a.n > 0xffffffffULL) {  // beyond the 32-bit index lists: one launch of the general machine, no scratch
    ctx->launches.fetch_add(1, std::memory_order_relaxed);
    unmet_single_kernel<<<(unsigned)((a.n + kUnmetThreads - 1) / kUnmetThreads), kUnmetThreads, 0, st>>>(a, d.zt, d.d_err);
    return cudaGetLastError();
  }
  const unsigned tiles = (unsigned)((a.n + kUnmetTile - 1) / kUnmetTile);
  if (!a.zone_id) {  // every Cron in UTC: the closed form always answers, nothing can be deferred, no scratch
    ctx->launches.fetch_add(1, std::memory_order_relaxed);
    if (a.m.p[0]) unmet_tile_kernel<false, false><<<tiles, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, nullptr, nullptr, nullptr, nullptr);
    else unmet_tile_kernel<false, true><<<tiles, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, nullptr, nullptr, nullptr, nullptr);
    return cudaGetLastError();
  }
  DeferList& b = d.defer[d.defer_next];
  d.defer_last = d.defer_next;
  d.defer_next = (d.defer_next + 1) % kDeferLists;
  cudaError_t e;
  const int32_t nz = d.zt.nzones;
  if (!b.done) {
    if ((e = cudaEventCreateWithFlags(&b.done, cudaEventDisableTiming)) != cudaSuccess) return e;
    // counters: [0] defer count | [1] class-B count | [2 .. nz+1] zone histogram | [nz+2 .. 2nz+2] cursors (last = total)
    if ((e = cudaMalloc(&b.count, (2 * (size_t)nz + 3) * sizeof(uint32_t))) != cudaSuccess) return e;
  } else if ((e = cudaStreamWaitEvent(st, b.done, 0)) != cudaSuccess) {
    return e;
  }
  const bool ordered = B200CRON_UNMET_PLAN == 1;
  // defer list (every Cron may be deferred) [+ class-B list + zone-ordered list]
  const size_t need = ordered ? 3 * a.n : a.n;
  if (need > b.cap) {
    if (b.idx) {
      if ((e = cudaEventSynchronize(b.done)) != cudaSuccess) return e;
      if ((e = cudaFree(b.idx)) != cudaSuccess) return e;
      b.idx = nullptr;
      b.cap = 0;
    }
    if ((e = cudaMalloc(&b.idx, need * sizeof(uint32_t))) != cudaSuccess) return e;
    b.cap = need;
  }
  const unsigned grid = (unsigned)((a.n + kUnmetThreads - 1) / kUnmetThreads);
  if (ordered) {
    uint32_t* z1_count = b.count + 1;
    uint32_t* hist = b.count + 2;
    uint32_t* cursor = hist + nz;
    uint32_t* z1_list = b.idx + a.n;
    uint32_t* order = b.idx + 2 * a.n;
    if ((e = cudaMemsetAsync(b.count, 0, (size_t)(nz + 2) * sizeof(uint32_t), st)) != cudaSuccess) return e;
    const unsigned zgrid = (unsigned)((a.n + kZoneOrderTile - 1) / kZoneOrderTile);
    ctx->launches.fetch_add(6, std::memory_order_relaxed);
    if (a.m.p[0]) unmet_tile_kernel<false, false><<<tiles, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, nullptr, nullptr, z1_list, z1_count);
    else unmet_tile_kernel<false, true><<<tiles, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, nullptr, nullptr, z1_list, z1_count);
    zone_order_hist_kernel<<<zgrid, kZoneOrderThreads, (size_t)nz * sizeof(uint32_t), st>>>(a.zone_id, z1_list, z1_count, nz, hist);
    zone_order_scan_kernel<<<1, kZoneOrderThreads, 0, st>>>(hist, nz, cursor);
    zone_order_scatter_kernel<<<zgrid, kZoneOrderThreads, 2 * (size_t)nz * sizeof(uint32_t), st>>>(a.zone_id, z1_list, z1_count,
                                                                                                  nz, cursor, order);
    // the scatter advanced cursor[z] to the end of zone z's range; cursor[nz] (the total) is untouched
    unmet_ordered_kernel<<<grid, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, b.idx, b.count, order, cursor + nz);
  } else {
#if B200CRON_UNMET_PLAN != 1
    if ((e = cudaMemsetAsync(b.count, 0, sizeof(uint32_t), st)) != cudaSuccess) return e;
    ctx->launches.fetch_add(2, std::memory_order_relaxed);
    unmet_tile_kernel<true, false><<<tiles, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, b.idx, b.count, nullptr, nullptr);
#endif
  }
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  if (B200CRON_UNMET_ORDERED_SINGLE && ordered) return cudaEventRecord(b.done, st);  // tuning build: nothing was deferred
  const unsigned hard_max = (unsigned)d.sm_count * B200CRON_UNMET_HARD_MINBLOCKS;
  unmet_hard_kernel<<<grid < hard_max ? grid : hard_max, kUnmetThreads, 0, st>>>(a, d.zt, d.d_err, b.idx, b.count);
  if ((e = cudaGetLastError()) != cudaSuccess) return e;
  return cudaEventRecord(b.done, st);
}

cudaError_t launch_history(b200cron_ctx* ctx, const Device& d, cudaStream_t st, const uint64_t* seg_off, const int64_t* key,
                           const int64_t* limit, size_t nseg, uint32_t* perm, uint8_t* keep) {
  if (nseg == 0) return cudaSuccess;
  const unsigned grid = (unsigned)((nseg + kHistSegs - 1) / kHistSegs);
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  history_prune_kernel<<<grid, kHistThreads, 0, st>>>(seg_off, key, limit, nseg, perm, keep, d.d_err);
  return cudaGetLastError();
}

int check_err_flags(b200cron_ctx* ctx) {
  int bits = 0;
  for (Device& d : ctx->devs) {
    bits |= *d.h_err;
    *d.h_err = 0;
  }
  if (bits & kErrSegLen)
    return fail(ctx, B200CRON_EINVAL, "history_prune: a segment has more than %llu records", (unsigned long long)kHistMaxSeg);
  if (bits & kErrZone) return fail(ctx, B200CRON_ERANGE, "zone_id outside [0, %d)", (int)ctx->zones.names.size());
  if (bits & kErrRange)
    return fail(ctx, B200CRON_ERANGE,
                "time outside the supported range (years 1..9999; zoned evaluations must stay below year %d)",
                ctx->zones.horizon_year - 6);
  return B200CRON_OK;
}

// fetch the device error word into the pinned mirror and clear it on the device (enqueued on `st`)
cudaError_t collect_err(Device& d, cudaStream_t st) {
  cudaError_t e = cudaMemcpyAsync(d.h_err, d.d_err, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e != cudaSuccess) return e;
  return cudaMemsetAsync(d.d_err, 0, sizeof(int), st);
}

struct Shard {
  size_t lo, hi;
};
Shard shard_of(size_t n, int g, int G) {
  // contiguous, even-sized boundaries so that 128-bit loads stay aligned on every device
  size_t lo = n * (size_t)g / (size_t)G, hi = n * (size_t)(g + 1) / (size_t)G;
  lo &= ~(size_t)1;
  if (g + 1 < G) hi &= ~(size_t)1;
  return {lo, hi};
}

Device* device_at(b200cron_ctx* ctx, int dev_index) {
  if (!ctx || dev_index < 0 || dev_index >= (int)ctx->devs.size()) return nullptr;
  return &ctx->devs[dev_index];
}

}  // namespace

// =================================================================================================================
// context
// =================================================================================================================
extern "C" int b200cron_ctx_create(const int* devices, int ndev, const char* zoneinfo_dir, int horizon_year,
                                   b200cron_ctx** out) {
  if (!out) return B200CRON_EINVAL;
  *out = nullptr;
  int count = 0;
  cudaError_t ce = cudaGetDeviceCount(&count);
  if (ce != cudaSuccess || count == 0)
    return fail(nullptr, B200CRON_ECUDA, "no usable CUDA device (%s); this library has no CPU path",
                ce == cudaSuccess ? "device count is 0" : cudaGetErrorString(ce));
  std::vector<int> ids;
  if (!devices || ndev <= 0) {
    int cur = 0;
    CU(nullptr, cudaGetDevice(&cur));
    ids.push_back(cur);
  } else {
    for (int i = 0; i < ndev; ++i) {
      if (devices[i] < 0 || devices[i] >= count) return fail(nullptr, B200CRON_EINVAL, "device %d does not exist", devices[i]);
      ids.push_back(devices[i]);
    }
  }
  std::string dir = zoneinfo_dir ? zoneinfo_dir : (getenv("ZONEINFO") ? getenv("ZONEINFO") : "/usr/share/zoneinfo");
  if (horizon_year <= 0) horizon_year = 2200;
  if (horizon_year < 2040 || horizon_year > 9000) return fail(nullptr, B200CRON_EINVAL, "horizon_year out of range");

  auto* ctx = new b200cron_ctx();
  std::string err;
  if (!build_zone_table(dir, horizon_year, &ctx->zones, &err)) {
    delete ctx;
    return fail(nullptr, B200CRON_EINVAL, "zone table: %s", err.c_str());
  }
  const HostZoneTable& z = ctx->zones;
  ctx->devs.resize(ids.size());
  auto cleanup = [&](int code) { b200cron_ctx_destroy(ctx); return code; };
  for (size_t i = 0; i < ids.size(); ++i) {
    Device& d = ctx->devs[i];
    d.id = ids[i];
#define CUC(call)                                                                                       \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) {                                                                            \
      fail(nullptr, B200CRON_ECUDA, "%s failed: %s", #call, cudaGetErrorString(e_));                    \
      return cleanup(B200CRON_ECUDA);                                                                   \
    }                                                                                                   \
  } while (0)
    CUC(cudaSetDevice(d.id));
    CUC(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, d.id));
    CUC(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
    for (Slot& s : d.slots) CUC(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    CUC(cudaMalloc(&d.d_when, z.when.size() * sizeof(int64_t)));
    CUC(cudaMalloc(&d.d_off, z.off.size() * sizeof(int32_t)));
    CUC(cudaMalloc(&d.d_fpre, z.fpre.size() * sizeof(int32_t)));
    CUC(cudaMalloc(&d.d_zstart, z.zstart.size() * sizeof(int32_t)));
    CUC(cudaMalloc(&d.d_yidx, z.yidx.size() * sizeof(uint16_t)));
    CUC(cudaMemcpy(d.d_yidx, z.yidx.data(), z.yidx.size() * sizeof(uint16_t), cudaMemcpyHostToDevice));
    CUC(cudaMalloc(&d.d_err, sizeof(int)));
    CUC(cudaHostAlloc(&d.h_err, sizeof(int), cudaHostAllocDefault));
    *d.h_err = 0;
    CUC(cudaMemcpy(d.d_when, z.when.data(), z.when.size() * sizeof(int64_t), cudaMemcpyHostToDevice));
    CUC(cudaMemcpy(d.d_off, z.off.data(), z.off.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CUC(cudaMemcpy(d.d_fpre, z.fpre.data(), z.fpre.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CUC(cudaMemcpy(d.d_zstart, z.zstart.data(), z.zstart.size() * sizeof(int32_t), cudaMemcpyHostToDevice));
    CUC(cudaMemset(d.d_err, 0, sizeof(int)));
#undef CUC
    d.zt.when = d.d_when;
    d.zt.off = d.d_off;
    d.zt.fpre = d.d_fpre;
    d.zt.zstart = d.d_zstart;
    d.zt.yidx = d.d_yidx;
    d.zt.ystride = z.ystride;
    d.zt.nzones = (int32_t)z.names.size();
    d.zt.horizon = z.horizon;
  }
  *out = ctx;
  return B200CRON_OK;
}

extern "C" void b200cron_ctx_destroy(b200cron_ctx* ctx) {
  if (!ctx) return;
  for (Device& d : ctx->devs) {
    cudaSetDevice(d.id);
    for (Slot& s : d.slots) {
      if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
      if (s.buf) cudaFree(s.buf);
    }
    if (d.stream) { cudaStreamSynchronize(d.stream); cudaStreamDestroy(d.stream); }
    if (d.d_when) cudaFree(d.d_when);
    if (d.d_off) cudaFree(d.d_off);
    if (d.d_fpre) cudaFree(d.d_fpre);
    if (d.d_zstart) cudaFree(d.d_zstart);
    if (d.d_yidx) cudaFree(d.d_yidx);
    if (d.d_err) cudaFree(d.d_err);
    if (d.d_scratch) cudaFree(d.d_scratch);
    for (DeferList& b : d.defer) {
      if (b.idx) cudaFree(b.idx);
      if (b.count) cudaFree(b.count);
      if (b.done) cudaEventDestroy(b.done);
    }
    if (d.h_err) cudaFreeHost(d.h_err);
  }
  delete ctx;
}

extern "C" const char* b200cron_last_error(const b200cron_ctx* ctx) {
  (void)ctx;
  return t_error.c_str();
}
extern "C" int b200cron_device_count(const b200cron_ctx* ctx) { return ctx ? (int)ctx->devs.size() : 0; }
extern "C" int b200cron_zone_id(const b200cron_ctx* ctx, const char* name) {
  if (!ctx || !name) return -1;
  return ctx->zones.find(name);
}
extern "C" int b200cron_zone_count(const b200cron_ctx* ctx) { return ctx ? (int)ctx->zones.names.size() : 0; }
extern "C" const char* b200cron_zone_name(const b200cron_ctx* ctx, int id) {
  if (!ctx || id < 0 || id >= (int)ctx->zones.names.size()) return nullptr;
  return ctx->zones.names[id].c_str();
}
extern "C" int b200cron_zone_periods(const b200cron_ctx* ctx, int id, const int64_t** when, const int32_t** off) {
  if (!ctx || id < 0 || id >= (int)ctx->zones.names.size() || !when || !off) return B200CRON_EINVAL;
  const int a = ctx->zones.zstart[id], b = ctx->zones.zstart[id + 1];
  *when = ctx->zones.when.data() + a;
  *off = ctx->zones.off.data() + a;
  return b - a;
}
extern "C" void* b200cron_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) return nullptr;
  return p;
}
extern "C" void b200cron_host_free(void* p) {
  if (p) cudaFreeHost(p);
}
extern "C" int b200cron_sync_check(b200cron_ctx* ctx, int dev_index, void* stream) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
  CU(ctx, collect_err(*d, st));
  CU(ctx, cudaStreamSynchronize(st));
  const int bits = *d->h_err;
  *d->h_err = 0;
  if (bits & kErrSegLen)
    return fail(ctx, B200CRON_EINVAL, "history_prune: a segment has more than %llu records", (unsigned long long)kHistMaxSeg);
  if (bits & kErrZone) return fail(ctx, B200CRON_ERANGE, "zone_id outside [0, %d)", (int)ctx->zones.names.size());
  if (bits & kErrRange)
    return fail(ctx, B200CRON_ERANGE,
                "time outside the supported range (years 1..9999; zoned evaluations must stay below year %d)",
                ctx->zones.horizon_year - 6);
  return B200CRON_OK;
}
extern "C" uint64_t b200cron_launch_count(const b200cron_ctx* ctx) { return ctx ? ctx->launches.load() : 0; }

// =================================================================================================================
// Next / NextN
// =================================================================================================================
extern "C" int b200cron_ctx_set_overlap(b200cron_ctx* ctx, int on) {
  if (!ctx) return B200CRON_EINVAL;
  ctx->overlap.store(on ? 1 : 0, std::memory_order_relaxed);
  return B200CRON_OK;
}

extern "C" int b200cron_next_n_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint64_t* const masks[6],
                                   const int32_t* zone_id, const int64_t* from, size_t n, int k, int64_t* out) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!masks || !from || !out || k < 1 || k > 64) return fail(ctx, B200CRON_EINVAL, "next_n: bad argument (1 <= k <= 64)");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "next_n: only the seconds plane (masks[0]) may be NULL");
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  MaskPlanes m;
  for (int p = 0; p < 6; ++p) m.p[p] = masks[p];
  cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
  CU(ctx, launch_next(ctx, *d, st, m, zone_id, from, n, k, out, true));
  return B200CRON_OK;
}

extern "C" int b200cron_next_n(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id,
                               const int64_t* from, size_t n, int k, int64_t* out) {
  if (!ctx) return B200CRON_EINVAL;
  if (!masks || !from || !out || k < 1 || k > 64) return fail(ctx, B200CRON_EINVAL, "next_n: bad argument (1 <= k <= 64)");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "next_n: only the seconds plane (masks[0]) may be NULL");
  std::lock_guard<std::mutex> lk(ctx->mu);
  const int G = (int)ctx->devs.size();
  // Zero-copy path: when every buffer is pinned host memory (b200cron_host_alloc / cudaHostAlloc / cudaHostRegister)
  // the kernels read and write it in place over PCIe -- no staging copies, the transfer overlaps the walk by itself.
  {
    bool pinned = is_pinned_host(from) && is_pinned_host(out) && (!zone_id || is_pinned_host(zone_id));
    for (int p = 0; p < 6 && pinned; ++p) pinned = (p == 0 && !masks[0]) || is_pinned_host(masks[p]);
    if (pinned && n > 0) {
      for (int g = 0; g < G; ++g) {
        Device& d = ctx->devs[g];
        const Shard sh = shard_of(n, g, G);
        if (sh.hi <= sh.lo) continue;
        CU(ctx, cudaSetDevice(d.id));
        MaskPlanes m;
        for (int p = 0; p < 6; ++p) m.p[p] = masks[p] ? dev_alias(masks[p]) + sh.lo : nullptr;
        CU(ctx, launch_next(ctx, d, d.stream, m, zone_id ? dev_alias(zone_id) + sh.lo : nullptr, dev_alias(from) + sh.lo,
                            sh.hi - sh.lo, k, dev_alias(out) + sh.lo * (size_t)k, false));
        CU(ctx, collect_err(d, d.stream));
      }
      for (Device& d : ctx->devs) {
        CU(ctx, cudaSetDevice(d.id));
        CU(ctx, cudaStreamSynchronize(d.stream));
      }
      return check_err_flags(ctx);
    }
  }
  // per-chunk staging layout: 6 mask planes | from | out (k per spec) | zone ids
  const size_t per_spec = 6 * 8 + 8 + (size_t)k * 8 + 4;
  for (int g = 0; g < G; ++g) {
    Device& d = ctx->devs[g];
    const Shard sh = shard_of(n, g, G);
    if (sh.hi <= sh.lo) continue;
    CU(ctx, cudaSetDevice(d.id));
    const size_t len = sh.hi - sh.lo;
    const size_t chunk = std::min(kChunk, align_up(len, 2));
    for (Slot& s : d.slots) {
      int rc = ensure_slot(ctx, s, align_up(chunk, 64) * per_spec + 1024);
      if (rc) return rc;
    }
    int slot = 0;
    for (size_t c = sh.lo; c < sh.hi; c += chunk, slot = (slot + 1) % kSlots) {
      Slot& s = d.slots[slot];
      const size_t cn = std::min(chunk, sh.hi - c);
      const size_t stride = align_up(chunk, 64);
      uint64_t* d_masks = reinterpret_cast<uint64_t*>(s.buf);
      int64_t* d_from = reinterpret_cast<int64_t*>(d_masks + 6 * stride);
      int64_t* d_out = d_from + stride;
      int32_t* d_zone = reinterpret_cast<int32_t*>(d_out + stride * k);
      MaskPlanes m;
      for (int p = 0; p < 6; ++p) {
        if (!masks[p]) { m.p[p] = nullptr; continue; }  // only plane 0 may be NULL (checked above)
        m.p[p] = d_masks + p * stride;
        CU(ctx, cudaMemcpyAsync(d_masks + p * stride, masks[p] + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      }
      CU(ctx, cudaMemcpyAsync(d_from, from + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      if (zone_id) CU(ctx, cudaMemcpyAsync(d_zone, zone_id + c, cn * 4, cudaMemcpyHostToDevice, s.stream));
      CU(ctx, launch_next(ctx, d, s.stream, m, zone_id ? d_zone : nullptr, d_from, cn, k, d_out, true));
      CU(ctx, cudaMemcpyAsync(out + c * (size_t)k, d_out, cn * (size_t)k * 8, cudaMemcpyDeviceToHost, s.stream));
    }
  }
  for (Device& d : ctx->devs) {
    CU(ctx, cudaSetDevice(d.id));
    for (Slot& s : d.slots) CU(ctx, cudaStreamSynchronize(s.stream));
    CU(ctx, collect_err(d, d.stream));
    CU(ctx, cudaStreamSynchronize(d.stream));
  }
  return check_err_flags(ctx);
}

// =================================================================================================================
// unmet_since
// =================================================================================================================
// Pre-sizes the unmet scratch (index lists + counters, all three buffers of every device) for batches of up to max_n
// Crons per device, so that no b200cron_unmet_since* call allocates or frees device memory afterwards.
extern "C" int b200cron_ctx_reserve_unmet(b200cron_ctx* ctx, size_t max_n) {
  if (!ctx) return B200CRON_EINVAL;
  if (max_n > 0xffffffffULL) return fail(ctx, B200CRON_EINVAL, "reserve_unmet: at most 2^32 - 1 Crons per device");
  std::lock_guard<std::mutex> lk(ctx->mu);
  for (Device& d : ctx->devs) {
    CU(ctx, cudaSetDevice(d.id));
    const size_t need = (B200CRON_UNMET_PLAN == 1 ? 3 : 1) * max_n;
    for (DeferList& b : d.defer) {
      if (!b.done) {
        CU(ctx, cudaEventCreateWithFlags(&b.done, cudaEventDisableTiming));
        CU(ctx, cudaMalloc(&b.count, (2 * (size_t)d.zt.nzones + 3) * sizeof(uint32_t)));
      }
      if (need > b.cap) {
        if (b.idx) {
          CU(ctx, cudaEventSynchronize(b.done));
          CU(ctx, cudaFree(b.idx));
          b.idx = nullptr;
          b.cap = 0;
        }
        CU(ctx, cudaMalloc(&b.idx, need * sizeof(uint32_t)));
        b.cap = need;
      }
    }
  }
  return B200CRON_OK;
}

extern "C" int b200cron_unmet_since_dev(b200cron_ctx* ctx, int dev_index, void* stream,
                                        const uint64_t* const masks[6], const int32_t* zone_id,
                                        const int64_t* earliest, int64_t now, int32_t now_nsec, const uint8_t* policy,
                                        const uint32_t* active, const uint8_t* flags, const int64_t* deadline,
                                        size_t n, int64_t* last, int64_t* next, uint32_t* missed, uint8_t* status,
                                        uint8_t* decision) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!masks || !earliest || !last || !next || !missed || !status || !decision)
    return fail(ctx, B200CRON_EINVAL, "unmet_since: NULL required argument");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "unmet_since: only the seconds plane (masks[0]) may be NULL");
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  UnmetArgs a{};
  for (int p = 0; p < 6; ++p) a.m.p[p] = masks[p];
  a.zone_id = zone_id; a.earliest = earliest; a.now = now; a.now_nsec = now_nsec; a.policy = policy;
  a.active = active; a.flags = flags; a.deadline = deadline; a.n = n;
  a.last = last; a.next = next; a.missed = missed; a.status = status; a.decision = decision;
  CU(ctx, launch_unmet(ctx, *d, stream ? (cudaStream_t)stream : d->stream, a));
  return B200CRON_OK;
}

extern "C" int b200cron_unmet_since(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id,
                                    const int64_t* earliest, int64_t now, int32_t now_nsec, const uint8_t* policy,
                                    const uint32_t* active, const uint8_t* flags, const int64_t* deadline, size_t n,
                                    int64_t* last, int64_t* next, uint32_t* missed, uint8_t* status,
                                    uint8_t* decision) {
  if (!ctx) return B200CRON_EINVAL;
  if (!masks || !earliest || !last || !next || !missed || !status || !decision)
    return fail(ctx, B200CRON_EINVAL, "unmet_since: NULL required argument");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "unmet_since: only the seconds plane (masks[0]) may be NULL");
  std::lock_guard<std::mutex> lk(ctx->mu);
  const int G = (int)ctx->devs.size();
  // Always staged, even from pinned buffers: the unmet kernels gather their inputs through index lists (class-pure and
  // zone-ordered warps) and scatter their results, 8 bytes at a time -- patterns that crawl when every access is a PCIe
  // transaction (measured: 296 ms per 10M Crons in place vs the copies' ~20 ms).  Pinned buffers still pay off: the
  // chunk copies below are then truly asynchronous and run at full PCIe rate.
  // staging: 6 planes | earliest | deadline | last | next  (8 B) | zone | active | missed (4 B) | policy flags status decision (1 B)
  const size_t per_spec = 10 * 8 + 3 * 4 + 4;
  for (int g = 0; g < G; ++g) {
    Device& d = ctx->devs[g];
    const Shard sh = shard_of(n, g, G);
    if (sh.hi <= sh.lo) continue;
    CU(ctx, cudaSetDevice(d.id));
    const size_t len = sh.hi - sh.lo;
    const size_t chunk = std::min(kUnmetChunk, align_up(len, 2));
    const size_t stride = align_up(chunk, 64);
    for (Slot& s : d.slots) {
      int rc = ensure_slot(ctx, s, stride * per_spec + 1024);
      if (rc) return rc;
    }
    int slot = 0;
    for (size_t c = sh.lo; c < sh.hi; c += chunk, slot = (slot + 1) % kSlots) {
      Slot& s = d.slots[slot];
      const size_t cn = std::min(chunk, sh.hi - c);
      uint64_t* q8 = reinterpret_cast<uint64_t*>(s.buf);
      uint64_t* d_masks = q8;
      int64_t* d_earliest = reinterpret_cast<int64_t*>(q8 + 6 * stride);
      int64_t* d_deadline = d_earliest + stride;
      int64_t* d_last = d_deadline + stride;
      int64_t* d_next = d_last + stride;
      int32_t* q4 = reinterpret_cast<int32_t*>(d_next + stride);
      int32_t* d_zone = q4;
      uint32_t* d_active = reinterpret_cast<uint32_t*>(q4 + stride);
      uint32_t* d_missed = d_active + stride;
      uint8_t* q1 = reinterpret_cast<uint8_t*>(d_missed + stride);
      uint8_t* d_policy = q1;
      uint8_t* d_flags = q1 + stride;
      uint8_t* d_status = q1 + 2 * stride;
      uint8_t* d_decision = q1 + 3 * stride;
      UnmetArgs a{};
      for (int p = 0; p < 6; ++p) {
        if (!masks[p]) { a.m.p[p] = nullptr; continue; }  // only plane 0 may be NULL (checked above): second 0
        a.m.p[p] = d_masks + p * stride;
        CU(ctx, cudaMemcpyAsync(d_masks + p * stride, masks[p] + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      }
      CU(ctx, cudaMemcpyAsync(d_earliest, earliest + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      if (deadline) CU(ctx, cudaMemcpyAsync(d_deadline, deadline + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      if (zone_id) CU(ctx, cudaMemcpyAsync(d_zone, zone_id + c, cn * 4, cudaMemcpyHostToDevice, s.stream));
      if (active) CU(ctx, cudaMemcpyAsync(d_active, active + c, cn * 4, cudaMemcpyHostToDevice, s.stream));
      if (policy) CU(ctx, cudaMemcpyAsync(d_policy, policy + c, cn, cudaMemcpyHostToDevice, s.stream));
      if (flags) CU(ctx, cudaMemcpyAsync(d_flags, flags + c, cn, cudaMemcpyHostToDevice, s.stream));
      a.zone_id = zone_id ? d_zone : nullptr;
      a.earliest = d_earliest; a.now = now; a.now_nsec = now_nsec;
      a.policy = policy ? d_policy : nullptr; a.active = active ? d_active : nullptr;
      a.flags = flags ? d_flags : nullptr; a.deadline = deadline ? d_deadline : nullptr;
      a.n = cn; a.last = d_last; a.next = d_next; a.missed = d_missed; a.status = d_status; a.decision = d_decision;
      CU(ctx, launch_unmet(ctx, d, s.stream, a));
      CU(ctx, cudaMemcpyAsync(last + c, d_last, cn * 8, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(next + c, d_next, cn * 8, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(missed + c, d_missed, cn * 4, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(status + c, d_status, cn, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(decision + c, d_decision, cn, cudaMemcpyDeviceToHost, s.stream));
    }
  }
  for (Device& d : ctx->devs) {
    CU(ctx, cudaSetDevice(d.id));
    for (Slot& s : d.slots) CU(ctx, cudaStreamSynchronize(s.stream));
    CU(ctx, collect_err(d, d.stream));
    CU(ctx, cudaStreamSynchronize(d.stream));
  }
  return check_err_flags(ctx);
}


// test support: which Crons did the most recent b200cron_unmet_since[_dev/_set] launch on this device hand to the
// general walk?  Copies up to `cap` of their indices (relative to that launch's arrays) to idx_host, *count = how many
// there were.  Waits for `stream` (NULL = the context's stream) first.
extern "C" int b200cron_debug_last_deferred(b200cron_ctx* ctx, int dev_index, void* stream, uint32_t* idx_host, size_t cap,
                                            size_t* count) {
  Device* d = device_at(ctx, dev_index);
  if (!d || !count) return B200CRON_EINVAL;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  CU(ctx, cudaStreamSynchronize(stream ? (cudaStream_t)stream : d->stream));
  *count = 0;
  if (d->defer_last < 0) return B200CRON_OK;
  const DeferList& b = d->defer[d->defer_last];
  uint32_t c = 0;
  CU(ctx, cudaMemcpy(&c, b.count, sizeof c, cudaMemcpyDeviceToHost));
  *count = c;
  const size_t take = std::min((size_t)c, cap);
  if (take && idx_host) CU(ctx, cudaMemcpy(idx_host, b.idx, take * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  return B200CRON_OK;
}

// =================================================================================================================
// resident spec sets
// =================================================================================================================
// A SpecSchedule is immutable after ParseStandard (cron_controller.go:392) and a Cron's schedule only changes when its
// CR does (api/v1alpha1/cron_types.go:77), yet the host-pointer calls above ship all 48 bytes of masks per spec on
// every tick.  A spec set keeps the mask planes and zone ids resident in device memory, sharded over the context's
// devices exactly as the host-pointer calls shard (shard_of); per tick only the times go in and the results come out.
struct b200cron_specset {
  b200cron_ctx* ctx = nullptr;
  size_t n = 0;
  bool have_sec = false, have_zone = false;
  struct Part {
    size_t lo = 0, hi = 0;
    uint64_t* planes = nullptr;  // 6 x stride
    int32_t* zone = nullptr;
    size_t stride = 0;
  };
  std::vector<Part> parts;  // one per device of the context
};

namespace {

MaskPlanes set_planes(const b200cron_specset* set, const b200cron_specset::Part& p, size_t off) {
  MaskPlanes m;
  for (int q = 0; q < 6; ++q) m.p[q] = (q == 0 && !set->have_sec) ? nullptr : p.planes + (size_t)q * p.stride + off;
  return m;
}

}  // namespace

extern "C" int b200cron_specset_create(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id,
                                       size_t n, b200cron_specset** out) {
  if (!ctx || !out) return B200CRON_EINVAL;
  *out = nullptr;
  if (!masks) return fail(ctx, B200CRON_EINVAL, "specset_create: NULL masks");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "specset_create: only the seconds plane (masks[0]) may be NULL");
  std::lock_guard<std::mutex> lk(ctx->mu);
  auto* set = new b200cron_specset();
  set->ctx = ctx;
  set->n = n;
  set->have_sec = masks[0] != nullptr;
  set->have_zone = zone_id != nullptr;
  const int G = (int)ctx->devs.size();
  set->parts.resize(G);
  for (int g = 0; g < G; ++g) {
    b200cron_specset::Part& p = set->parts[g];
    const Shard sh = shard_of(n, g, G);
    p.lo = sh.lo;
    p.hi = sh.hi;
    const size_t len = sh.hi > sh.lo ? sh.hi - sh.lo : 0;
    if (len == 0) continue;
    p.stride = align_up(len, 64);
    cudaError_t e = cudaSetDevice(ctx->devs[g].id);
    if (e == cudaSuccess) e = cudaMalloc(&p.planes, 6 * p.stride * sizeof(uint64_t));
    if (e == cudaSuccess && zone_id) e = cudaMalloc(&p.zone, p.stride * sizeof(int32_t));
    for (int q = 0; q < 6 && e == cudaSuccess; ++q)
      if (masks[q]) e = cudaMemcpy(p.planes + (size_t)q * p.stride, masks[q] + sh.lo, len * 8, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && zone_id) e = cudaMemcpy(p.zone, zone_id + sh.lo, len * 4, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      for (int h = 0; h <= g; ++h) {
        cudaSetDevice(ctx->devs[h].id);
        if (set->parts[h].planes) cudaFree(set->parts[h].planes);
        if (set->parts[h].zone) cudaFree(set->parts[h].zone);
      }
      delete set;
      return fail(ctx, e == cudaErrorMemoryAllocation ? B200CRON_ENOMEM : B200CRON_ECUDA, "specset_create: %s",
                  cudaGetErrorString(e));
    }
  }
  *out = set;
  return B200CRON_OK;
}

extern "C" void b200cron_specset_destroy(b200cron_specset* set) {
  if (!set) return;
  std::lock_guard<std::mutex> lk(set->ctx->mu);
  for (size_t g = 0; g < set->parts.size(); ++g) {
    cudaSetDevice(set->ctx->devs[g].id);
    if (set->parts[g].planes) cudaFree(set->parts[g].planes);
    if (set->parts[g].zone) cudaFree(set->parts[g].zone);
  }
  delete set;
}

extern "C" size_t b200cron_specset_size(const b200cron_specset* set) { return set ? set->n : 0; }

extern "C" int b200cron_specset_update(b200cron_specset* set, size_t first, size_t count,
                                       const uint64_t* const masks[6], const int32_t* zone_id) {
  if (!set) return B200CRON_EINVAL;
  b200cron_ctx* ctx = set->ctx;
  if (!masks || first > set->n || count > set->n - first) return fail(ctx, B200CRON_EINVAL, "specset_update: range outside the set");
  for (int p = 0; p < 6; ++p)
    if (!masks[p] && (p > 0 || set->have_sec))
      return fail(ctx, B200CRON_EINVAL, "specset_update: plane %d is NULL but the set holds one", p);
  if (set->have_zone != (zone_id != nullptr))
    return fail(ctx, B200CRON_EINVAL, "specset_update: zone ids must be given exactly when the set was created with them");
  std::lock_guard<std::mutex> lk(ctx->mu);
  for (size_t g = 0; g < set->parts.size(); ++g) {
    const b200cron_specset::Part& p = set->parts[g];
    const size_t lo = std::max(first, p.lo), hi = std::min(first + count, p.hi);
    if (hi <= lo) continue;
    Device& d = ctx->devs[g];
    CU(ctx, cudaSetDevice(d.id));
    // in stream order behind whatever still reads the old values on the context's stream
    for (int q = 0; q < 6; ++q)
      if (q > 0 || set->have_sec)
        CU(ctx, cudaMemcpyAsync(p.planes + (size_t)q * p.stride + (lo - p.lo), masks[q] + (lo - first), (hi - lo) * 8,
                                cudaMemcpyHostToDevice, d.stream));
    if (zone_id)
      CU(ctx, cudaMemcpyAsync(p.zone + (lo - p.lo), zone_id + (lo - first), (hi - lo) * 4, cudaMemcpyHostToDevice, d.stream));
    CU(ctx, cudaStreamSynchronize(d.stream));
  }
  return B200CRON_OK;
}

// Next / NextN over a resident set: only `from` goes in and `out` comes back.  Pinned host buffers are used in place
// (the kernel reads the times and writes the results over PCIe while the masks stream from HBM); pageable ones go
// through the staged chunk pipeline.
extern "C" int b200cron_next_n_set(b200cron_ctx* ctx, const b200cron_specset* set, const int64_t* from, int k,
                                   int64_t* out) {
  if (!ctx || !set || set->ctx != ctx) return B200CRON_EINVAL;
  if (!from || !out || k < 1 || k > 64) return fail(ctx, B200CRON_EINVAL, "next_n_set: bad argument (1 <= k <= 64)");
  std::lock_guard<std::mutex> lk(ctx->mu);
  const int G = (int)ctx->devs.size();
  // pinned buffers are used in place: measured 254 us per 1M specs against 322 us through staged chunk copies and 280 us
  // with `from` copied in and the results written in place (profiles/r2o_ab_set_*.log)
  const bool pinned = is_pinned_host(from) && is_pinned_host(out);
  for (int g = 0; g < G; ++g) {
    Device& d = ctx->devs[g];
    const b200cron_specset::Part& p = set->parts[g];
    if (p.hi <= p.lo) continue;
    CU(ctx, cudaSetDevice(d.id));
    if (pinned) {
      CU(ctx, launch_next(ctx, d, d.stream, set_planes(set, p, 0), p.zone, dev_alias(from) + p.lo, p.hi - p.lo, k,
                          dev_alias(out) + p.lo * (size_t)k, false));
      continue;
    }
    const size_t chunk = std::min(kChunk, align_up(p.hi - p.lo, 2));
    const size_t stride = align_up(chunk, 64);
    for (Slot& s : d.slots) {
      int rc = ensure_slot(ctx, s, stride * (8 + (size_t)k * 8) + 1024);
      if (rc) return rc;
    }
    int slot = 0;
    for (size_t c = p.lo; c < p.hi; c += chunk, slot = (slot + 1) % kSlots) {
      Slot& s = d.slots[slot];
      const size_t cn = std::min(chunk, p.hi - c);
      int64_t* d_from = reinterpret_cast<int64_t*>(s.buf);
      int64_t* d_out = d_from + stride;
      CU(ctx, cudaMemcpyAsync(d_from, from + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      CU(ctx, launch_next(ctx, d, s.stream, set_planes(set, p, c - p.lo), p.zone ? p.zone + (c - p.lo) : nullptr, d_from, cn,
                          k, d_out, true));
      CU(ctx, cudaMemcpyAsync(out + c * (size_t)k, d_out, cn * (size_t)k * 8, cudaMemcpyDeviceToHost, s.stream));
    }
  }
  for (Device& d : ctx->devs) {
    CU(ctx, cudaSetDevice(d.id));
    if (!pinned)
      for (Slot& s : d.slots) CU(ctx, cudaStreamSynchronize(s.stream));
    CU(ctx, collect_err(d, d.stream));
    CU(ctx, cudaStreamSynchronize(d.stream));
  }
  return check_err_flags(ctx);
}

// getNextSchedule + gates + decision over a resident set: the per-tick arrays go in, the five result arrays come out
extern "C" int b200cron_unmet_since_set(b200cron_ctx* ctx, const b200cron_specset* set, const int64_t* earliest,
                                        int64_t now, int32_t now_nsec, const uint8_t* policy, const uint32_t* active,
                                        const uint8_t* flags, const int64_t* deadline, int64_t* last, int64_t* next,
                                        uint32_t* missed, uint8_t* status, uint8_t* decision) {
  if (!ctx || !set || set->ctx != ctx) return B200CRON_EINVAL;
  if (!earliest || !last || !next || !missed || !status || !decision)
    return fail(ctx, B200CRON_EINVAL, "unmet_since_set: NULL required argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  const int G = (int)ctx->devs.size();
  // always staged, pinned buffers or not: see b200cron_unmet_since
  // staging: earliest | deadline | last | next (8 B) | active | missed (4 B) | policy flags status decision (1 B)
  const size_t per_spec = 4 * 8 + 2 * 4 + 4;
  for (int g = 0; g < G; ++g) {
    Device& d = ctx->devs[g];
    const b200cron_specset::Part& p = set->parts[g];
    if (p.hi <= p.lo) continue;
    CU(ctx, cudaSetDevice(d.id));
    UnmetArgs a{};
    a.now = now; a.now_nsec = now_nsec;
    const size_t chunk = std::min(kUnmetChunk, align_up(p.hi - p.lo, 2));
    const size_t stride = align_up(chunk, 64);
    for (Slot& s : d.slots) {
      int rc = ensure_slot(ctx, s, stride * per_spec + 1024);
      if (rc) return rc;
    }
    int slot = 0;
    for (size_t c = p.lo; c < p.hi; c += chunk, slot = (slot + 1) % kSlots) {
      Slot& s = d.slots[slot];
      const size_t cn = std::min(chunk, p.hi - c);
      int64_t* d_earliest = reinterpret_cast<int64_t*>(s.buf);
      int64_t* d_deadline = d_earliest + stride;
      int64_t* d_last = d_deadline + stride;
      int64_t* d_next = d_last + stride;
      uint32_t* d_active = reinterpret_cast<uint32_t*>(d_next + stride);
      uint32_t* d_missed = d_active + stride;
      uint8_t* q1 = reinterpret_cast<uint8_t*>(d_missed + stride);
      uint8_t *d_policy = q1, *d_flags = q1 + stride, *d_status = q1 + 2 * stride, *d_decision = q1 + 3 * stride;
      CU(ctx, cudaMemcpyAsync(d_earliest, earliest + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      if (deadline) CU(ctx, cudaMemcpyAsync(d_deadline, deadline + c, cn * 8, cudaMemcpyHostToDevice, s.stream));
      if (active) CU(ctx, cudaMemcpyAsync(d_active, active + c, cn * 4, cudaMemcpyHostToDevice, s.stream));
      if (policy) CU(ctx, cudaMemcpyAsync(d_policy, policy + c, cn, cudaMemcpyHostToDevice, s.stream));
      if (flags) CU(ctx, cudaMemcpyAsync(d_flags, flags + c, cn, cudaMemcpyHostToDevice, s.stream));
      a.m = set_planes(set, p, c - p.lo);
      a.zone_id = p.zone ? p.zone + (c - p.lo) : nullptr;
      a.earliest = d_earliest;
      a.policy = policy ? d_policy : nullptr; a.active = active ? d_active : nullptr;
      a.flags = flags ? d_flags : nullptr; a.deadline = deadline ? d_deadline : nullptr;
      a.n = cn; a.last = d_last; a.next = d_next; a.missed = d_missed; a.status = d_status; a.decision = d_decision;
      CU(ctx, launch_unmet(ctx, d, s.stream, a));
      CU(ctx, cudaMemcpyAsync(last + c, d_last, cn * 8, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(next + c, d_next, cn * 8, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(missed + c, d_missed, cn * 4, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(status + c, d_status, cn, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(decision + c, d_decision, cn, cudaMemcpyDeviceToHost, s.stream));
    }
  }
  for (Device& d : ctx->devs) {
    CU(ctx, cudaSetDevice(d.id));
    for (Slot& s : d.slots) CU(ctx, cudaStreamSynchronize(s.stream));
    CU(ctx, collect_err(d, d.stream));
    CU(ctx, cudaStreamSynchronize(d.stream));
  }
  return check_err_flags(ctx);
}

// =================================================================================================================
// unmet list
// =================================================================================================================
namespace {

// per-Cron list lengths are summed in 32 bits inside a tile of kScanTile Crons: cap <= 2^20 keeps a tile below 2^30
constexpr uint32_t kListCapMax = 1u << 20;

int launch_list_offsets(b200cron_ctx* ctx, Device& d, cudaStream_t st, const uint32_t* missed, const uint8_t* decision,
                        size_t n, uint32_t cap, uint64_t* list_off, uint64_t* tile_sums) {
  const size_t ntiles = (n + kScanTile - 1) / kScanTile;
  ctx->launches.fetch_add(3, std::memory_order_relaxed);
  list_tile_sums_kernel<<<(unsigned)ntiles, kScanThreads, 0, st>>>(missed, decision, n, cap, tile_sums);
  list_scan_sums_kernel<<<1, kScanThreads, 0, st>>>(tile_sums, ntiles);
  list_offsets_kernel<<<(unsigned)ntiles, kScanThreads, 0, st>>>(missed, decision, n, cap, tile_sums, list_off);
  CU(ctx, cudaGetLastError());
  (void)d;
  return 0;
}

int ensure_scratch(b200cron_ctx* ctx, Device& d, size_t entries) {
  if (entries <= d.scratch_cap) return 0;
  if (d.d_scratch) CU(ctx, cudaFree(d.d_scratch));
  d.d_scratch = nullptr;
  d.scratch_cap = 0;
  CU(ctx, cudaMalloc(&d.d_scratch, entries * sizeof(uint64_t)));
  d.scratch_cap = entries;
  return 0;
}

}  // namespace

extern "C" int b200cron_unmet_list_offsets_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint32_t* missed,
                                               const uint8_t* decision, size_t n, uint32_t cap, uint64_t* list_off) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!missed || !decision || !list_off) return fail(ctx, B200CRON_EINVAL, "unmet_list_offsets: NULL argument");
  if (cap > kListCapMax) return fail(ctx, B200CRON_EINVAL, "unmet_list_offsets: cap = %u exceeds %u", cap, kListCapMax);
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
  if (n == 0) {
    CU(ctx, cudaMemsetAsync(list_off, 0, sizeof(uint64_t), st));
    return B200CRON_OK;
  }
  const size_t ntiles = (n + kScanTile - 1) / kScanTile;
  int rc = ensure_scratch(ctx, *d, ntiles + 1);
  if (rc) return rc;
  return launch_list_offsets(ctx, *d, st, missed, decision, n, cap, list_off, d->d_scratch);
}


// -----------------------------------------------------------------------------------------------------------------
// b200cron_unmet_list_offsets_static: the same prefix sum with its per-tile sums in a fixed device array instead of
// the context's cudaMalloc'd scratch.
//
// Who needs it: hosts that cannot let this library allocate device memory at call time -- a service that hands the
// whole GPU to its own pool allocator at start-up (a later cudaMalloc fails or fragments that pool), or a reconcile
// loop that must not run into the device-wide synchronisation cudaMalloc/cudaFree imply while other streams are
// busy.  b200cron_unmet_list_offsets_dev grows the context's scratch with cudaMalloc whenever n outgrows it; this
// variant never allocates or frees anything.
//
// The array below is reserved on every GPU that loads the library, from module load until the process exits
// (kStaticTiles + 1 uint64 = 128 KiB + 8 B per GPU).  In host code `g_static_tile_sums` is only a stand-in for the
// symbol, not a device address -- a kernel handed that name would fault (or, where the GPU can reach host memory,
// silently use the wrong memory) -- so the entry point resolves the real address once per device with
// cudaGetSymbolAddress while that device is current, and passes the resolved pointer to the kernels.
// -----------------------------------------------------------------------------------------------------------------
namespace {
constexpr size_t kStaticTiles = 16384;                        // tiles of kScanTile Crons
constexpr size_t kStaticMaxN = kStaticTiles * (size_t)kScanTile;  // 16,777,216 Crons per call
static __device__ __align__(256) unsigned long long g_static_tile_sums[kStaticTiles + 1];
}  // namespace

extern "C" int b200cron_unmet_list_offsets_static(b200cron_ctx* ctx, int dev_index, void* stream,
                                                  const uint32_t* missed, const uint8_t* decision, size_t n,
                                                  uint32_t cap, uint64_t* list_off) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!missed || !decision || !list_off) return fail(ctx, B200CRON_EINVAL, "unmet_list_offsets_static: NULL argument");
  if (cap > kListCapMax) return fail(ctx, B200CRON_EINVAL, "unmet_list_offsets_static: cap = %u exceeds %u", cap, kListCapMax);
  if (n > kStaticMaxN)
    return fail(ctx, B200CRON_EINVAL,
                "unmet_list_offsets_static: n = %zu exceeds the fixed scratch (at most %zu Crons per call); split the "
                "batch or use b200cron_unmet_list_offsets_dev",
                n, kStaticMaxN);
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  if (!d->d_static_sums) {
    // first call on this device: look the array up while this device is current (no allocation involved)
    void* p = nullptr;
    CU(ctx, cudaGetSymbolAddress(&p, g_static_tile_sums));
    d->d_static_sums = static_cast<uint64_t*>(p);
  }
  cudaStream_t st = stream ? (cudaStream_t)stream : d->stream;
  if (n == 0) {
    CU(ctx, cudaMemsetAsync(list_off, 0, sizeof(uint64_t), st));
    return B200CRON_OK;
  }
  // Nothing in the array has to start at zero: entries [0, ntiles) are written by list_tile_sums_kernel and entry
  // [ntiles] by list_scan_sums_kernel before anything reads them; whatever an earlier (larger) call left behind in
  // the entries above ntiles is never read.
  return launch_list_offsets(ctx, *d, st, missed, decision, n, cap, list_off, d->d_static_sums);
}

extern "C" int b200cron_unmet_list_emit_dev(b200cron_ctx* ctx, int dev_index, void* stream,
                                            const uint64_t* const masks[6], const int32_t* zone_id,
                                            const int64_t* earliest, size_t n, const uint64_t* list_off,
                                            int64_t* list) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!masks || !earliest || !list_off) return fail(ctx, B200CRON_EINVAL, "unmet_list_emit: NULL argument");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "unmet_list_emit: only the seconds plane (masks[0]) may be NULL");
  if (n == 0) return B200CRON_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  MaskPlanes m;
  for (int p = 0; p < 6; ++p) m.p[p] = masks[p];
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  unmet_list_emit_kernel<<<(unsigned)((n + kUnmetThreads - 1) / kUnmetThreads), kUnmetThreads, 0,
                           stream ? (cudaStream_t)stream : d->stream>>>(m, zone_id, earliest, n, list_off, list, d->zt);
  CU(ctx, cudaGetLastError());
  return B200CRON_OK;
}

extern "C" int b200cron_unmet_list(b200cron_ctx* ctx, const uint64_t* const masks[6], const int32_t* zone_id,
                                   const int64_t* earliest, const uint32_t* missed, const uint8_t* decision, size_t n,
                                   uint32_t cap, uint64_t* list_off, int64_t* list, size_t list_cap) {
  if (!ctx) return B200CRON_EINVAL;
  if (!masks || !earliest || !missed || !decision || !list_off)
    return fail(ctx, B200CRON_EINVAL, "unmet_list: NULL required argument");
  for (int p = 1; p < 6; ++p)
    if (!masks[p]) return fail(ctx, B200CRON_EINVAL, "unmet_list: only the seconds plane (masks[0]) may be NULL");
  if (cap > kListCapMax) return fail(ctx, B200CRON_EINVAL, "unmet_list: cap = %u exceeds %u", cap, kListCapMax);
  if (n == 0) { list_off[0] = 0; return B200CRON_OK; }
  std::lock_guard<std::mutex> lk(ctx->mu);
  Device& d = ctx->devs[0];
  CU(ctx, cudaSetDevice(d.id));
  Slot& s = d.slots[0];
  const size_t stride = align_up(n + 1, 64);
  const size_t ntiles = (n + kScanTile - 1) / kScanTile;
  // staging: 6 planes | earliest | list_off (8 B) | zone | missed (4 B) | decision (1 B)
  int rc = ensure_slot(ctx, s, stride * (8 * 8 + 2 * 4 + 1) + 1024);
  if (rc) return rc;
  rc = ensure_scratch(ctx, d, ntiles + 1);
  if (rc) return rc;
  uint64_t* d_masks = reinterpret_cast<uint64_t*>(s.buf);
  int64_t* d_earliest = reinterpret_cast<int64_t*>(d_masks + 6 * stride);
  uint64_t* d_off = reinterpret_cast<uint64_t*>(d_earliest + stride);
  int32_t* d_zone = reinterpret_cast<int32_t*>(d_off + stride);
  uint32_t* d_missed = reinterpret_cast<uint32_t*>(d_zone + stride);
  uint8_t* d_decision = reinterpret_cast<uint8_t*>(d_missed + stride);
  MaskPlanes m;
  for (int p = 0; p < 6; ++p) {
    if (!masks[p]) { m.p[p] = nullptr; continue; }  // only plane 0 may be NULL (checked above): second 0
    m.p[p] = d_masks + p * stride;
    CU(ctx, cudaMemcpyAsync(d_masks + p * stride, masks[p], n * 8, cudaMemcpyHostToDevice, s.stream));
  }
  CU(ctx, cudaMemcpyAsync(d_earliest, earliest, n * 8, cudaMemcpyHostToDevice, s.stream));
  if (zone_id) CU(ctx, cudaMemcpyAsync(d_zone, zone_id, n * 4, cudaMemcpyHostToDevice, s.stream));
  CU(ctx, cudaMemcpyAsync(d_missed, missed, n * 4, cudaMemcpyHostToDevice, s.stream));
  CU(ctx, cudaMemcpyAsync(d_decision, decision, n, cudaMemcpyHostToDevice, s.stream));
  rc = launch_list_offsets(ctx, d, s.stream, d_missed, d_decision, n, cap, d_off, d.d_scratch);
  if (rc) return rc;
  CU(ctx, cudaMemcpyAsync(list_off, d_off, (n + 1) * 8, cudaMemcpyDeviceToHost, s.stream));
  CU(ctx, cudaStreamSynchronize(s.stream));
  const uint64_t total = list_off[n];
  if (total > list_cap || (total && !list))
    return fail(ctx, B200CRON_ENOMEM, "unmet_list: list needs %llu entries, capacity is %llu",
                (unsigned long long)total, (unsigned long long)list_cap);
  if (total == 0) return B200CRON_OK;
  Slot& s1 = d.slots[1];
  rc = ensure_slot(ctx, s1, total * 8 + 1024);
  if (rc) return rc;
  int64_t* d_list = reinterpret_cast<int64_t*>(s1.buf);
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  unmet_list_emit_kernel<<<(unsigned)((n + kUnmetThreads - 1) / kUnmetThreads), kUnmetThreads, 0, s.stream>>>(
      m, zone_id ? d_zone : nullptr, d_earliest, n, d_off, d_list, d.zt);
  CU(ctx, cudaGetLastError());
  CU(ctx, cudaMemcpyAsync(list, d_list, total * 8, cudaMemcpyDeviceToHost, s.stream));
  CU(ctx, cudaStreamSynchronize(s.stream));
  return B200CRON_OK;
}

// =================================================================================================================
// history_prune
// =================================================================================================================
extern "C" int b200cron_history_prune_dev(b200cron_ctx* ctx, int dev_index, void* stream, const uint64_t* seg_off,
                                          const int64_t* key, const int64_t* limit, size_t nseg, size_t nrec,
                                          uint32_t* perm, uint8_t* keep) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!seg_off || (nrec && (!key || !perm || !keep))) return fail(ctx, B200CRON_EINVAL, "history_prune: NULL argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  CU(ctx, launch_history(ctx, *d, stream ? (cudaStream_t)stream : d->stream, seg_off, key, limit, nseg, perm, keep));
  return B200CRON_OK;
}

extern "C" int b200cron_history_prune(b200cron_ctx* ctx, const uint64_t* seg_off, const int64_t* key,
                                      const int64_t* limit, size_t nseg, uint32_t* perm, uint8_t* keep) {
  if (!ctx) return B200CRON_EINVAL;
  if (!seg_off) return fail(ctx, B200CRON_EINVAL, "history_prune: NULL seg_off");
  if (nseg == 0) return B200CRON_OK;
  for (size_t s = 0; s < nseg; ++s) {
    if (seg_off[s + 1] < seg_off[s]) return fail(ctx, B200CRON_EINVAL, "history_prune: seg_off must be non-decreasing");
    if (seg_off[s + 1] - seg_off[s] > kHistMaxSeg)
      return fail(ctx, B200CRON_EINVAL, "history_prune: segment %zu has %llu records (at most %llu per Cron)", s,
                  (unsigned long long)(seg_off[s + 1] - seg_off[s]), (unsigned long long)kHistMaxSeg);
  }
  const uint64_t nrec = seg_off[nseg] - seg_off[0];
  if (nrec && (!key || !perm || !keep)) return fail(ctx, B200CRON_EINVAL, "history_prune: NULL argument");
  std::lock_guard<std::mutex> lk(ctx->mu);
  const int G = (int)ctx->devs.size();
  // Cut on segment boundaries so that records per device balance and no segment straddles two devices.
  std::vector<size_t> cut(G + 1, nseg);
  cut[0] = 0;
  for (int g = 1; g < G; ++g) {
    const uint64_t target = seg_off[0] + nrec * (uint64_t)g / (uint64_t)G;
    cut[g] = (size_t)(std::lower_bound(seg_off, seg_off + nseg + 1, target) - seg_off);
    cut[g] = std::min(std::max(cut[g], cut[g - 1]), nseg);
  }
  for (int g = 0; g < G; ++g) {
    Device& d = ctx->devs[g];
    const size_t s0 = cut[g], s1 = cut[g + 1];
    if (s1 <= s0) continue;
    CU(ctx, cudaSetDevice(d.id));
    const uint64_t r0 = seg_off[s0], r1 = seg_off[s1];
    const size_t ns = s1 - s0, nr = (size_t)(r1 - r0);
    Slot& s = d.slots[0];
    const size_t a_ns = align_up(ns + 1, 64), a_nr = align_up(nr ? nr : 1, 64);
    int rc = ensure_slot(ctx, s, a_ns * 16 + a_nr * 13 + 1024);
    if (rc) return rc;
    uint64_t* d_off = reinterpret_cast<uint64_t*>(s.buf);
    int64_t* d_limit = reinterpret_cast<int64_t*>(d_off + a_ns);
    int64_t* d_key = d_limit + a_ns;
    uint32_t* d_perm = reinterpret_cast<uint32_t*>(d_key + a_nr);
    uint8_t* d_keep = reinterpret_cast<uint8_t*>(d_perm + a_nr);
    CU(ctx, cudaMemcpyAsync(d_off, seg_off + s0, (ns + 1) * 8, cudaMemcpyHostToDevice, s.stream));
    if (limit) CU(ctx, cudaMemcpyAsync(d_limit, limit + s0, ns * 8, cudaMemcpyHostToDevice, s.stream));
    if (nr) CU(ctx, cudaMemcpyAsync(d_key, key + r0, nr * 8, cudaMemcpyHostToDevice, s.stream));
    // the kernel addresses records by absolute offset: bias the record pointers by -r0
    CU(ctx, launch_history(ctx, d, s.stream, d_off, d_key - r0, limit ? d_limit : nullptr, ns, d_perm - r0, d_keep - r0));
    if (nr) {
      CU(ctx, cudaMemcpyAsync(perm + r0, d_perm, nr * 4, cudaMemcpyDeviceToHost, s.stream));
      CU(ctx, cudaMemcpyAsync(keep + r0, d_keep, nr, cudaMemcpyDeviceToHost, s.stream));
    }
  }
  for (Device& d : ctx->devs) {
    CU(ctx, cudaSetDevice(d.id));
    CU(ctx, cudaStreamSynchronize(d.slots[0].stream));
  }
  return B200CRON_OK;
}

// =================================================================================================================
// synthetic workloads (bench / test support)
// =================================================================================================================
extern "C" int b200cron_synth_specs_dev(b200cron_ctx* ctx, int dev_index, void* stream, uint64_t seed, uint64_t start,
                                        size_t n, int with_seconds, double dense_minutes, double plant_unschedulable,
                                        const double thresholds[10], uint64_t* const masks_out[6]) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!thresholds || !masks_out) return fail(ctx, B200CRON_EINVAL, "synth_specs: NULL argument");
  if (n == 0) return B200CRON_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  SynthThresholds th;
  for (int q = 0; q < 5; ++q) { th.t[0][q] = thresholds[q]; th.t[1][q] = thresholds[5 + q]; }
  SynthPlanes out;
  for (int p = 0; p < 6; ++p) out.p[p] = masks_out[p];
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  synth_specs_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream ? (cudaStream_t)stream : d->stream>>>(
      seed, start, n, with_seconds, dense_minutes, plant_unschedulable, th, out);
  CU(ctx, cudaGetLastError());
  return B200CRON_OK;
}

extern "C" int b200cron_synth_times_dev(b200cron_ctx* ctx, int dev_index, void* stream, uint64_t seed, uint64_t start,
                                        size_t n, int64_t lo, int64_t hi, uint64_t stream_id, int64_t* out) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!out || hi <= lo) return fail(ctx, B200CRON_EINVAL, "synth_times: bad argument");
  if (n == 0) return B200CRON_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  synth_times_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream ? (cudaStream_t)stream : d->stream>>>(
      seed, start, n, lo, hi, stream_id, out);
  CU(ctx, cudaGetLastError());
  return B200CRON_OK;
}

extern "C" int b200cron_synth_zones_dev(b200cron_ctx* ctx, int dev_index, void* stream, uint64_t seed, uint64_t start,
                                        size_t n, const int32_t* pool, size_t pool_len, double p_utc, int32_t* out) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!out || !pool || pool_len == 0) return fail(ctx, B200CRON_EINVAL, "synth_zones: bad argument");
  if (n == 0) return B200CRON_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  synth_zones_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream ? (cudaStream_t)stream : d->stream>>>(
      seed, start, n, pool, pool_len, p_utc, out);
  CU(ctx, cudaGetLastError());
  return B200CRON_OK;
}

// =================================================================================================================
// batched ParseStandard on the device
// =================================================================================================================
extern "C" int b200cron_parse_batch_dev(b200cron_ctx* ctx, int dev_index, void* stream, const char* text,
                                        const uint64_t* text_off, size_t n, int with_seconds,
                                        uint64_t* const masks_out[6], uint8_t* status, uint32_t* tz_off,
                                        uint32_t* tz_len) {
  Device* d = device_at(ctx, dev_index);
  if (!d) return B200CRON_EINVAL;
  if (!text_off || !masks_out || !status || !tz_off || !tz_len) return fail(ctx, B200CRON_EINVAL, "parse_batch: NULL argument");
  if (n == 0) return B200CRON_OK;
  std::lock_guard<std::mutex> lk(ctx->mu);
  CU(ctx, cudaSetDevice(d->id));
  ParseOut out;
  for (int p = 0; p < 6; ++p) out.p[p] = masks_out[p];
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  parse_batch_kernel<<<(unsigned)((n + 127) / 128), 128, 0, stream ? (cudaStream_t)stream : d->stream>>>(
      text, text_off, n, with_seconds, out, status, tz_off, tz_len);
  CU(ctx, cudaGetLastError());
  return B200CRON_OK;
}

extern "C" int b200cron_parse_batch(b200cron_ctx* ctx, const char* text, const uint64_t* text_off, size_t n,
                                    int with_seconds, uint64_t* const masks_out[6], uint8_t* status,
                                    uint32_t* tz_off, uint32_t* tz_len) {
  if (!ctx) return B200CRON_EINVAL;
  if (!text_off || !masks_out || !status || !tz_off || !tz_len) return fail(ctx, B200CRON_EINVAL, "parse_batch: NULL argument");
  if (n == 0) return B200CRON_OK;
  for (size_t i = 0; i < n; ++i)
    if (text_off[i + 1] < text_off[i] || text_off[i + 1] - text_off[i] > (1u << 20))
      return fail(ctx, B200CRON_EINVAL, "parse_batch: text_off must be non-decreasing (specs up to 1 MiB)");
  const uint64_t t0 = text_off[0], bytes = text_off[n] - t0;
  if (bytes && !text) return fail(ctx, B200CRON_EINVAL, "parse_batch: NULL text");
  std::lock_guard<std::mutex> lk(ctx->mu);
  Device& d = ctx->devs[0];
  CU(ctx, cudaSetDevice(d.id));
  Slot& s = d.slots[0];
  const size_t stride = align_up(n + 1, 64), tbytes = align_up(bytes + 16, 64);
  int rc = ensure_slot(ctx, s, tbytes + stride * (8 + 6 * 8 + 4 + 4 + 1) + 1024);
  if (rc) return rc;
  char* d_text = reinterpret_cast<char*>(s.buf);
  uint64_t* d_off = reinterpret_cast<uint64_t*>(d_text + tbytes);
  uint64_t* d_masks = d_off + stride;
  uint32_t* d_tzo = reinterpret_cast<uint32_t*>(d_masks + 6 * stride);
  uint32_t* d_tzl = d_tzo + stride;
  uint8_t* d_status = reinterpret_cast<uint8_t*>(d_tzl + stride);
  if (bytes) CU(ctx, cudaMemcpyAsync(d_text, text + t0, bytes, cudaMemcpyHostToDevice, s.stream));
  CU(ctx, cudaMemcpyAsync(d_off, text_off, (n + 1) * 8, cudaMemcpyHostToDevice, s.stream));
  ParseOut out;
  for (int p = 0; p < 6; ++p) out.p[p] = d_masks + p * stride;
  ctx->launches.fetch_add(1, std::memory_order_relaxed);
  // offsets are absolute in the caller's buffer: bias the text pointer by -t0
  parse_batch_kernel<<<(unsigned)((n + 127) / 128), 128, 0, s.stream>>>(d_text - t0, d_off, n, with_seconds, out, d_status,
                                                                      d_tzo, d_tzl);
  CU(ctx, cudaGetLastError());
  for (int p = 0; p < 6; ++p)
    CU(ctx, cudaMemcpyAsync(masks_out[p], d_masks + p * stride, n * 8, cudaMemcpyDeviceToHost, s.stream));
  CU(ctx, cudaMemcpyAsync(status, d_status, n, cudaMemcpyDeviceToHost, s.stream));
  CU(ctx, cudaMemcpyAsync(tz_off, d_tzo, n * 4, cudaMemcpyDeviceToHost, s.stream));
  CU(ctx, cudaMemcpyAsync(tz_len, d_tzl, n * 4, cudaMemcpyDeviceToHost, s.stream));
  CU(ctx, cudaStreamSynchronize(s.stream));
  return B200CRON_OK;
}
