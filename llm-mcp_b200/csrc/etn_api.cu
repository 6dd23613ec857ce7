// etn_api.cu — the device-pointer half of the C-ABI of libembed_truncate_norm.so (declared in
// include/embed_truncate_norm.h): argument checks, kernel selection, metrics, the entry
// point that stages through the built-in array.  The host-buffer runtime (persistent
// per-device workers, chunked prefix-only H2D, kernel, D2H, coalescing, row sharding) is
// etn_host.cu.
//
// Reference behaviour being replaced: truncateEmbeddings, core/internal/api/handlers.go:2062-2078,
// called at handlers.go:1891-1893.  No torch, no NCCL, no CPU compute fallback: if the
// CUDA runtime reports an error the caller gets its code.
#include "../../include/embed_truncate_norm.h"
#include "etn_internal.h"
#include "etn_kernels.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <atomic>
#include <mutex>
#include <vector>

namespace {

using namespace etn;
using etn_rt::aligned_to;
using etn_rt::count;
using etn_rt::finish;

// ---------------------------------------------------------------- metrics ----
// Cumulative, process-wide; the Go side exports them next to the embedding collectors of
// core/internal/metrics/metrics.go:9-26.  Order = ETN_CTR_* in the header.
std::atomic<int64_t> g_ctr[ETN_CTR_COUNT];
inline int counted_launch(cudaError_t e) {
  if (e == cudaSuccess) count(ETN_CTR_KERNEL_LAUNCHES);
  return (int)e;
}

// Fixed-bucket histograms in the shape of prometheus.HistogramOpts (metrics.go:18-26 is the
// one to sit beside): upper bounds, per-bucket counts (not cumulative here), sum of samples.
constexpr int kMaxBuckets = 16;
struct Hist {
  const char* name;
  const char* help;
  int n;                                    // finite upper bounds; bucket n is +Inf
  double le[kMaxBuckets];
  std::atomic<int64_t> cnt[kMaxBuckets + 1];
  std::atomic<int64_t> sum_fixed;           // sum in units of 2^-30 (atomic add on integers)
};
Hist g_hist[ETN_HIST_COUNT] = {
    {"llmcore_embed_gpu_batch_rows", "Rows per host-path call", 11,
     {1, 4, 16, 64, 256, 1024, 4096, 16384, 65536, 262144, 1048576}, {}, {}},
    {"llmcore_embed_gpu_kernel_seconds", "Kernel time per staged chunk (CUDA events)", 12,
     {5e-6, 1e-5, 2e-5, 5e-5, 1e-4, 2e-4, 5e-4, 1e-3, 2e-3, 5e-3, 1e-2, 0.1}, {}, {}},
    {"llmcore_embed_gpu_h2d_seconds", "Host-to-device copy time per staged chunk (CUDA events)", 12,
     {5e-6, 1e-5, 2e-5, 5e-5, 1e-4, 2e-4, 5e-4, 1e-3, 2e-3, 5e-3, 1e-2, 0.1}, {}, {}},
    {"llmcore_embed_gpu_d2h_seconds", "Device-to-host copy time per staged chunk (CUDA events)", 12,
     {5e-6, 1e-5, 2e-5, 5e-5, 1e-4, 2e-4, 5e-4, 1e-3, 2e-3, 5e-3, 1e-2, 0.1}, {}, {}},
    {"llmcore_embed_gpu_call_seconds", "Host-path call latency, submit to completion", 12,
     {5e-5, 1e-4, 2e-4, 5e-4, 1e-3, 2e-3, 5e-3, 1e-2, 2e-2, 5e-2, 0.1, 1.0}, {}, {}},
    {"llmcore_embed_gpu_requests_per_launch", "Host-path requests sharing one staged chunk (coalescing)", 8,
     {1, 2, 4, 8, 16, 32, 64, 128}, {}, {}},
};

// ------------------------------------------------------------ device info ----
struct DeviceInfo {
  std::atomic<int> sms{0};
};
constexpr int kMaxDevices = 64;
constexpr int64_t kWideMaxRowBytes = 112 * 1024;    // two row buffers must fit the 227 KiB of shared memory
DeviceInfo g_info[kMaxDevices];

int sm_count(int dev, int* out) {
  if (dev < 0 || dev >= kMaxDevices) { *out = 148; return 0; }
  int v = g_info[dev].sms.load(std::memory_order_relaxed);
  if (v == 0) {
    cudaError_t e = cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return (int)e;
    g_info[dev].sms.store(v, std::memory_order_relaxed);
  }
  *out = v;
  return 0;
}

// Grid: one warp-iteration per row group (waves == 0, the hardware block scheduler balances
// the SMs), or `waves` full waves of resident blocks looping grid-stride.
template <typename K>
int grid_for(K kernel, int64_t blocks_needed, int waves, int* grid) {
  int64_t cap = (int64_t)1 << 30;
  if (waves > 0) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (int rc = sm_count(dev, &sms)) return rc;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kThreads, 0);
    if (e != cudaSuccess) return (int)e;
    cap = (int64_t)sms * std::max(per_sm, 1) * waves;
  }
  *grid = (int)std::max<int64_t>(1, std::min<int64_t>(blocks_needed, cap));
  return 0;
}

// ------------------------------------------------------------- launchers ----
// Programmatic dependent launch: a launch may be scheduled while the previous kernel of the
// stream drains, and its blocks wait in griddepcontrol.wait (first thing every kernel does)
// until that kernel has completed and its writes are visible.  This hides the 3-5 us between
// back-to-back request-sized launches (C1: 8 us, C2: 20 us of work).  ETN_PDL=0 turns it off.
bool pdl_enabled() {
  static const bool on = [] { const char* v = getenv("ETN_PDL"); return !(v && v[0] == '0'); }();
  return on;
}

template <typename... KArgs, typename... Args>
int launch_kernel(void (*kern)(KArgs...), int grid, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return counted_launch(cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...));
}

// MINB = resident blocks per SM the register allocation is held to.
template <typename T, int VB, int NV, int R, int L, int MODE, int MINB, int ALIAS>
int launch_reg(const void* in, void* out, int64_t batch, int64_t in_pitch, int nvec, int waves, cudaStream_t s) {
  auto kern = rows_reg_kernel<T, VB, NV, R, L, MODE, MINB, VB == 32 ? 1 : 0, ALIAS>;
  int grid = 0;
  const int64_t rows_per_block = (int64_t)kWarpsPerBlock * R * (32 / L);
  if (int rc = grid_for(kern, (batch + rows_per_block - 1) / rows_per_block, waves, &grid)) return rc;
  return launch_kernel(kern, grid, s, (const char*)in, (char*)out, batch, in_pitch, nvec);
}

// Wide rows: one row per block iteration, double-buffered in shared memory.
template <typename T, int MODE>
int launch_wide(const void* in, void* out, int64_t batch, int64_t in_pitch, int nvec16, cudaStream_t s) {
  auto kern = rows_wide_kernel<T, MODE>;
  const size_t stride = ((size_t)nvec16 * 16 + 127) & ~(size_t)127;
  const size_t smem = 2 * stride;
  int dev = 0, sms = 148;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  sm_count(dev, &sms);
  {  // dynamic shared memory beyond 48 KiB is an opt-in per function and device; the largest size is set once
    static std::atomic<uint64_t> done{0};
    const uint64_t bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * (int)kWideMaxRowBytes);
      if (e != cudaSuccess) return (int)e;
      done.fetch_or(bit, std::memory_order_release);
    }
  }
  // two rows per block keep one copy in flight behind the row being reduced; with a single
  // resident block per SM (rows > 56 KiB) four amortise the pipeline fill better
  int rpb = smem > 113 * 1024 ? 4 : 2;
  rpb = (int)std::max<int64_t>(1, std::min<int64_t>(rpb, batch / ((int64_t)sms * 2)));
  const int64_t blocks = std::min<int64_t>((batch + rpb - 1) / rpb, (int64_t)1 << 30);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)std::max<int64_t>(blocks, 1));
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return counted_launch(cudaLaunchKernelEx(&cfg, kern, (const char*)in, (char*)out, batch, in_pitch, nvec16, rpb));
}

// Bulk-async (TMA) ring kernel: 4 warps per block, each with its own S-slot ring of row
// prefixes in shared memory; every warp owns `rpw` consecutive rows.
template <typename T, int NV16, int S, int MODE>
int launch_bulk(const void* in, void* out, int64_t batch, int64_t in_pitch, int nvec16, int rpw, cudaStream_t s) {
  constexpr int NCW = 4;
  constexpr size_t smem = (size_t)NCW * S * NV16 * 512 + (size_t)NCW * S * 8;
  auto kern = rows_bulk_kernel<T, NV16, S, NCW, MODE>;
  if (smem > 48 * 1024) {                              // opt in to large dynamic shared memory once per device
    static std::atomic<uint64_t> done{0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    const uint64_t bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return (int)e;
      done.fetch_or(bit, std::memory_order_release);
    }
  }
  {  // request-sized batches: fewer rows per warp, so that every SM still gets a few blocks
    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) sm_count(dev, &sms);
    rpw = (int)std::max<int64_t>(1, std::min<int64_t>(rpw, batch / ((int64_t)NCW * sms * 2)));
  }
  const int64_t rows_per_block = (int64_t)NCW * rpw;
  const int64_t blocks = std::min<int64_t>((batch + rows_per_block - 1) / rows_per_block, (int64_t)1 << 30);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)std::max<int64_t>(blocks, 1));
  cfg.blockDim = dim3(NCW * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return counted_launch(cudaLaunchKernelEx(&cfg, kern, (const char*)in, (char*)out, batch, in_pitch, nvec16, rpw));
}

template <typename T, int MODE>
int launch_general(const void* in, void* out, int64_t batch, int64_t D, int64_t k,
                   const int32_t* k_per_row, const int64_t* out_offsets, const int64_t* in_offsets, cudaStream_t s) {
  auto kern = rows_general_kernel<T, MODE>;
  int grid = 0;
  if (int rc = grid_for(kern, (batch + kWarpsPerBlock - 1) / kWarpsPerBlock, 0, &grid)) return rc;
  return launch_kernel(kern, grid, s, (const T*)in, (T*)out, batch, D, k, k_per_row, out_offsets, in_offsets);
}

template <typename T, int MODE>
int launch_ring(const void* in, void* out, int64_t batch, int64_t D, int64_t k_dense, const int32_t* k_per_row,
                const int64_t* out_offsets, const int64_t* in_offsets, cudaStream_t s);

// Dense rows.  Kernel choice by prefix size and alignment:
//   16-byte tileable, float64 renorm, 512 B - 8 KiB:  rows_bulk (bulk-async ring)
//   16-byte tileable, 8 KiB - 112 KiB:                rows_wide (row in shared memory, whole block)
//   32-byte tileable, <= 8 KiB:                       rows_reg with 256-bit accesses
//   16-byte tileable, <= 8 KiB:                       rows_reg with 128-bit accesses
//   odd k on aligned rows, or float64 rows off the 16-byte grid, < 8 KiB:
//                                                     rows_ring (aligned span staged, element-wise)
//   anything else (float32 rows off the grid, longer rows of that kind): rows_general
// Truncation moves bytes, so one instantiation (T = float) serves both element types.
// ALIAS = 1 (out == in): the register kernels load through the coherent path (no ld.global.nc).
template <typename T, int MODE, int ALIAS>
int launch_dense(const void* in, void* out, int64_t batch, int64_t D, int64_t kk, cudaStream_t s) {
  using TT = typename std::conditional<MODE == kModeTruncate, float, T>::type;
  // float64 renorm carries the compensated sum and the double-double scale: more registers
  // per thread, so fewer rows per warp and fewer resident blocks (tools/tune.cu, profiles/).
  constexpr bool heavy = MODE == kModeRenorm && sizeof(T) == 8;
  constexpr int R1 = heavy ? 2 : 4, R2 = heavy ? 1 : 2;
  constexpr int MB = 4, MB4 = heavy ? 3 : 4, MB8 = heavy ? 2 : 4;
  const int64_t in_pitch = D * (int64_t)sizeof(T);
  const int64_t k_bytes = kk * (int64_t)sizeof(T);
  const int W = 0;                                                  // one row group per warp
  if constexpr (heavy) {
    // float64 renorm: the reduction is long enough to expose memory latency in the register
    // kernels (5.5-5.8 TB/s sustained); staging rows through the bulk-async ring keeps
    // 6.7-6.8 TB/s at every size from 1 to 8 KiB (profiles/r01_tune_bulk.txt).
    if (aligned_to(in, 16) && aligned_to(out, 16) && in_pitch % 16 == 0 && k_bytes % 16 == 0 &&
        k_bytes > 512 && k_bytes <= 8192) {
      const int nvec16 = (int)(k_bytes / 16);
      const int nv = (nvec16 + 31) / 32;
      if (nv <= 2)  return launch_bulk<T, 2, 4, MODE>(in, out, batch, in_pitch, nvec16, 8, s);
      if (nv <= 4)  return launch_bulk<T, 4, 2, MODE>(in, out, batch, in_pitch, nvec16, 8, s);
      if (nv <= 8)  return launch_bulk<T, 8, 2, MODE>(in, out, batch, in_pitch, nvec16, 4, s);
      if (nv <= 12) return launch_bulk<T, 12, 2, MODE>(in, out, batch, in_pitch, nvec16, 4, s);
      return launch_bulk<T, 16, 2, MODE>(in, out, batch, in_pitch, nvec16, 4, s);
    }
  }
  const bool tile16 = aligned_to(in, 16) && aligned_to(out, 16) && in_pitch % 16 == 0 && k_bytes % 16 == 0;
  if (tile16 && k_bytes > 8192 && k_bytes <= kWideMaxRowBytes)
    return launch_wide<T, MODE>(in, out, batch, in_pitch, (int)(k_bytes / 16), s);
  if (aligned_to(in, 32) && aligned_to(out, 32) && in_pitch % 32 == 0 && k_bytes % 32 == 0 &&
      k_bytes <= 8192) {
    const int nvec = (int)(k_bytes / 32);
    const int nv = (nvec + 31) / 32;
    // short rows: 8 or 16 lanes per row, so a warp finishes 4 or 2 rows per reduction
    constexpr int RS = heavy ? 2 : 8, MBS = heavy ? 4 : 2;
    if (nvec <= 8)  return launch_reg<TT, 32, 1, RS, 8, MODE, MBS, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nvec <= 16) return launch_reg<TT, 32, 1, RS, 16, MODE, MBS, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 1) return launch_reg<TT, 32, 1, R1, 32, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 2) return launch_reg<TT, 32, 2, R2, 32, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 4) return launch_reg<TT, 32, 4, 1, 32, MODE, MB4, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    // 4-8 KiB rows: float32 renorm does as well through the bulk ring (6.79 vs 6.65 TB/s)
    if constexpr (MODE == kModeRenorm && !heavy)
      return launch_bulk<T, 16, 2, MODE>(in, out, batch, in_pitch, (int)(k_bytes / 16), 4, s);
    return launch_reg<TT, 32, 8, 1, 32, MODE, MB8, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
  }
  if (aligned_to(in, 16) && aligned_to(out, 16) && in_pitch % 16 == 0 && k_bytes % 16 == 0 &&
      k_bytes <= 16 * 512) {
    const int nvec = (int)(k_bytes / 16);
    const int nv = (nvec + 31) / 32;
    if (nvec <= 8)  return launch_reg<TT, 16, 1, 2 * R1, 8, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nvec <= 16) return launch_reg<TT, 16, 1, 2 * R1, 16, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 1) return launch_reg<TT, 16, 1, 2 * R1, 32, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 2) return launch_reg<TT, 16, 2, R1, 32, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 4) return launch_reg<TT, 16, 4, R2, 32, MODE, MB, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    if (nv <= 8) return launch_reg<TT, 16, 8, 1, 32, MODE, MB4, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
    return launch_reg<TT, 16, 16, 1, 32, MODE, MB8, ALIAS>(in, out, batch, in_pitch, nvec, W, s);
  }
  // prefix not 16-byte tileable (odd k) or rows off the 16-byte grid (odd D, shifted base): the
  // aligned span around every prefix is staged through the page ring
  // Measured (profiles/r02_offgrid.txt, 1,000,000 x 771 -> 257 and 500,000 x 4097 -> 1024): off the
  // grid the ring wins for float64 (5.6 / 5.0 TB/s against 5.1 / 5.0 in the general kernel) and
  // loses for float32, whose 1 KiB rows are all per-row bookkeeping there (3.7 against 4.8 TB/s).
  {
    const bool on_grid = aligned_to(in, 16) && in_pitch % 16 == 0;
    if (k_bytes <= 8192 - 32 && (on_grid || sizeof(T) == 8))
      return launch_ring<T, MODE>(in, out, batch, D, kk, nullptr, nullptr, nullptr, s);
  }
  return launch_general<T, MODE>(in, out, batch, D, kk, nullptr, nullptr, nullptr, s);
}

// Ragged rows: every row prefix is one bulk-async copy of its own length into a warp's
// page-granular ring (rows_ring_kernel), so mixed k costs no padding traffic and short rows
// still keep the ring full; rows that are not 16-byte tileable or exceed 8 KiB fall back
// inside the kernel.
// in_offsets != nullptr: the caller guarantees every row start is 16-byte aligned.
template <typename T, int MODE>
int launch_ring(const void* in, void* out, int64_t batch, int64_t D, int64_t k_dense, const int32_t* k_per_row,
                const int64_t* out_offsets, const int64_t* in_offsets, cudaStream_t s) {
  const int64_t in_pitch = D * (int64_t)sizeof(T);
  constexpr int NCW = 4;
  constexpr size_t smem = (size_t)NCW * 8192 + (size_t)NCW * 8 * 8 + (size_t)NCW * 8 * 3 * 4;
  // 16 pages per warp, register budget of 4 resident blocks (6 still fit), 4 rows per warp: tools/tune ring,
  // profiles/r02_tune_ring.txt (config-5 mix 6.39 TB/s, 64 merged C1 requests 6.17 TB/s)
  auto kern = rows_ring_kernel<T, NCW, MODE, 16, 4>;
  {
    static std::atomic<uint64_t> done{0};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    const uint64_t bit = 1ull << (dev & 63);
    if (!(done.load(std::memory_order_acquire) & bit)) {
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return (int)e;
      done.fetch_or(bit, std::memory_order_release);
    }
  }
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) sm_count(dev, &sms);
  const int rpw = (int)std::max<int64_t>(1, std::min<int64_t>(4, batch / ((int64_t)NCW * sms * 2)));
  const int64_t rows_per_block = (int64_t)NCW * rpw;
  const int64_t blocks = std::min<int64_t>((batch + rows_per_block - 1) / rows_per_block, (int64_t)1 << 30);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)std::max<int64_t>(blocks, 1));
  cfg.blockDim = dim3(NCW * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return counted_launch(cudaLaunchKernelEx(&cfg, kern, (const char*)in, (char*)out, batch, in_pitch, rpw, D, k_dense,
                                           k_per_row, out_offsets, in_offsets));
}

int check_dense(int64_t batch, int64_t D, int64_t k, int dtype, const void* in, const void* out) {
  if (dtype != ETN_F32 && dtype != ETN_F64) return ETN_ERR_ENUM;
  if (batch < 0 || D <= 0 || k <= 0) return ETN_ERR_SHAPE;
  if (batch == 0) return ETN_OK;
  if (!in || !out) return ETN_ERR_NULL;
  const size_t es = dtype == ETN_F32 ? 4 : 8;
  if (!aligned_to(in, es) || !aligned_to(out, es)) return ETN_ERR_ALIGN;
  if (in == out && k < D) return ETN_ERR_SHAPE;     // in place only when whole rows are kept
  return ETN_OK;
}

}  // namespace

// ------------------------------------------------ shared with etn_host.cu ----
namespace etn_rt {

void count(int i, int64_t v) { g_ctr[i].fetch_add(v, std::memory_order_relaxed); }

int finish(int rc) {
  if (rc == ETN_ERR_JSON || rc == ETN_ERR_JSON_DUP || rc == ETN_ERR_TOO_LARGE || rc == ETN_ERR_NOT_INIT) count(ETN_CTR_FALLBACKS);
  else if (rc != 0) count(ETN_CTR_ERRORS);
  return rc;
}

void observe(int which, double v) {
  if (which < 0 || which >= ETN_HIST_COUNT || !(v >= 0.0)) return;
  Hist& h = g_hist[which];
  int b = 0;
  while (b < h.n && v > h.le[b]) ++b;
  h.cnt[b].fetch_add(1, std::memory_order_relaxed);
  h.sum_fixed.fetch_add((int64_t)std::llround(v * 1073741824.0), std::memory_order_relaxed);
}

int run_dense(int64_t batch, int64_t D, int64_t k, int dtype, int mode, const void* in, void* out,
              cudaStream_t s) {
  const int64_t kk = std::min(k, D);
  if (in == out) {                                   // k >= D: whole rows, normalised where they lie
    if (mode == ETN_TRUNCATE) return ETN_OK;         // copying a row onto itself
    return dtype == ETN_F32 ? launch_dense<float, kModeRenorm, 1>(in, out, batch, D, kk, s)
                            : launch_dense<double, kModeRenorm, 1>(in, out, batch, D, kk, s);
  }
  if (mode == ETN_TRUNCATE) {
    return dtype == ETN_F32 ? launch_dense<float, kModeTruncate, 0>(in, out, batch, D, kk, s)
                            : launch_dense<double, kModeTruncate, 0>(in, out, batch, D, kk, s);
  }
  return dtype == ETN_F32 ? launch_dense<float, kModeRenorm, 0>(in, out, batch, D, kk, s)
                          : launch_dense<double, kModeRenorm, 0>(in, out, batch, D, kk, s);
}

int run_rows(int64_t batch, int64_t D, const int32_t* k_per_row, const int64_t* out_offsets,
             const int64_t* in_offsets, int dtype, int mode, const void* in, void* out, cudaStream_t s) {
  if (mode == ETN_TRUNCATE)
    return dtype == ETN_F32
               ? launch_ring<float, kModeTruncate>(in, out, batch, D, 0, k_per_row, out_offsets, in_offsets, s)
               : launch_ring<double, kModeTruncate>(in, out, batch, D, 0, k_per_row, out_offsets, in_offsets, s);
  return dtype == ETN_F32
             ? launch_ring<float, kModeRenorm>(in, out, batch, D, 0, k_per_row, out_offsets, in_offsets, s)
             : launch_ring<double, kModeRenorm>(in, out, batch, D, 0, k_per_row, out_offsets, in_offsets, s);
}

}  // namespace etn_rt

using etn_rt::run_dense;

// ------------------------------------------------- built-in staging array ----
// Device staging for embed_truncate_norm_host_static: reserved in the module image, so it
// exists from the moment the library is loaded on a GPU until the process exits and no
// allocator is ever involved.  In host code `g_static_stage` is only a handle for the
// symbol, not a device address; the address a kernel or a copy can use is looked up once
// per device with cudaGetSymbolAddress, with that device current.
static __device__ __align__(256) unsigned char g_static_stage[ETN_STATIC_STAGE_BYTES];

struct StaticStage {
  std::mutex mu;            // set-up, and one call at a time on the array
  void* ptr = nullptr;      // device address of g_static_stage on this device
};
static StaticStage g_static[kMaxDevices];

// =============================================================== C-ABI ======
extern "C" {

int etn_version(void) { return ETN_VERSION; }

const char* etn_strerror(int code) {
  switch (code) {
    case ETN_OK: return "ok";
    case ETN_ERR_NULL: return "null pointer";
    case ETN_ERR_SHAPE: return "bad shape (batch < 0, D <= 0, k <= 0, bad pitch or offsets, or out == in with k < D)";
    case ETN_ERR_ENUM: return "unknown dtype or mode";
    case ETN_ERR_ALIGN: return "pointer not aligned to the element size";
    case ETN_ERR_NOT_INIT: return "etn_init has not been called";
    case ETN_ERR_DEVICE: return "bad device list";
    case ETN_ERR_JSON: return "malformed JSON, non-finite value or number outside float64";
    case ETN_ERR_JSON_DUP: return "duplicate \"data\" or \"embedding\" key";
    case ETN_ERR_TOO_LARGE: return "batch does not fit the built-in staging array (ETN_STATIC_STAGE_BYTES)";
    case ETN_ERR_TICKET: return "unknown or already waited-for ticket";
    default: break;
  }
  if (code > 0) return cudaGetErrorString((cudaError_t)code);
  return "unknown error";
}

int embed_truncate_norm(int64_t batch, int64_t D, int64_t k, int dtype, const void* in, void* out,
                        void* stream) {
  if (int rc = check_dense(batch, D, k, dtype, in, out)) return rc;
  if (batch == 0) return ETN_OK;
  count(ETN_CTR_DEVICE_CALLS); count(ETN_CTR_DEVICE_ROWS, batch);
  return finish(run_dense(batch, D, k, dtype, ETN_TRUNCATE_RENORM, in, out, (cudaStream_t)stream));
}

int embed_truncate(int64_t batch, int64_t D, int64_t k, int dtype, const void* in, void* out,
                   void* stream) {
  if (int rc = check_dense(batch, D, k, dtype, in, out)) return rc;
  if (batch == 0) return ETN_OK;
  count(ETN_CTR_DEVICE_CALLS); count(ETN_CTR_DEVICE_ROWS, batch);
  return finish(run_dense(batch, D, k, dtype, ETN_TRUNCATE, in, out, (cudaStream_t)stream));
}

static int check_rows(int64_t batch, int64_t D, const int32_t* k_per_row, const int64_t* out_offsets,
                      int dtype, int mode, const void* in, const void* out) {
  if ((dtype != ETN_F32 && dtype != ETN_F64) || (mode != ETN_TRUNCATE && mode != ETN_TRUNCATE_RENORM))
    return ETN_ERR_ENUM;
  if (batch < 0 || D <= 0) return ETN_ERR_SHAPE;
  if (batch == 0) return ETN_OK;
  if (!in || !out || !k_per_row || !out_offsets) return ETN_ERR_NULL;
  const size_t es = dtype == ETN_F32 ? 4 : 8;
  if (!aligned_to(in, es) || !aligned_to(out, es) || !aligned_to(k_per_row, 4) ||
      !aligned_to(out_offsets, 8))
    return ETN_ERR_ALIGN;
  return ETN_OK;
}

int embed_truncate_norm_ragged(int64_t batch, int64_t D, const int32_t* k_per_row,
                               const int64_t* out_offsets, int dtype, int mode, const void* in,
                               void* out, void* stream) {
  if (int rc = check_rows(batch, D, k_per_row, out_offsets, dtype, mode, in, out)) return rc;
  if (batch == 0) return ETN_OK;
  count(ETN_CTR_DEVICE_CALLS); count(ETN_CTR_DEVICE_ROWS, batch);
  return finish(etn_rt::run_rows(batch, D, k_per_row, out_offsets, nullptr, dtype, mode, in, out,
                                 (cudaStream_t)stream));
}

int embed_truncate_norm_gather(int64_t batch, const int64_t* in_offsets, const int32_t* k_per_row,
                               const int64_t* out_offsets, int dtype, int mode, const void* in_base,
                               void* out_base, void* stream) {
  if (int rc = check_rows(batch, 1, k_per_row, out_offsets, dtype, mode, in_base, out_base)) return rc;
  if (batch == 0) return ETN_OK;
  if (!in_offsets) return ETN_ERR_NULL;
  if (!aligned_to(in_offsets, 8)) return ETN_ERR_ALIGN;
  // rows reach the page ring by bulk-async copies, which need 16-byte aligned sources: the
  // base is checked here, the offsets are the caller's promise (element offsets that are
  // multiples of 16 / element size)
  if (!aligned_to(in_base, 16)) return ETN_ERR_ALIGN;
  count(ETN_CTR_DEVICE_CALLS); count(ETN_CTR_DEVICE_ROWS, batch);
  return finish(etn_rt::run_rows(batch, INT64_MAX / 16, k_per_row, out_offsets, in_offsets, dtype, mode,
                                 in_base, out_base, (cudaStream_t)stream));
}

int embed_truncate_norm_host_static(int64_t batch, int64_t D, int64_t k, int dtype, int mode,
                                    const void* host_in, int64_t in_pitch_bytes, void* host_out,
                                    void* stream) {
  if ((dtype != ETN_F32 && dtype != ETN_F64) || (mode != ETN_TRUNCATE && mode != ETN_TRUNCATE_RENORM))
    return ETN_ERR_ENUM;
  if (batch < 0 || D <= 0 || k <= 0) return ETN_ERR_SHAPE;
  const int64_t es = dtype == ETN_F32 ? 4 : 8;
  if (in_pitch_bytes == 0) in_pitch_bytes = D * es;
  if (in_pitch_bytes < D * es || in_pitch_bytes % es != 0) return ETN_ERR_SHAPE;
  if (batch == 0) return ETN_OK;
  if (!host_in || !host_out) return ETN_ERR_NULL;
  if (!aligned_to(host_in, (size_t)es) || !aligned_to(host_out, (size_t)es)) return ETN_ERR_ALIGN;
  const int64_t kk = std::min(k, D);
  const int64_t k_bytes = kk * es;
  if (k_bytes > (int64_t)ETN_STATIC_STAGE_BYTES || batch > (int64_t)ETN_STATIC_STAGE_BYTES / k_bytes)
    return finish(ETN_ERR_TOO_LARGE);

  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return finish((int)e);
  if (dev < 0 || dev >= kMaxDevices) return finish(ETN_ERR_DEVICE);
  StaticStage& st = g_static[dev];
  // Held until the stream has drained: every call on this device uses the same array.
  std::lock_guard<std::mutex> lock(st.mu);
  if (!st.ptr) {                                    // first call on this device: set-up
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_stage);
    if (e != cudaSuccess) return finish((int)e);
    st.ptr = p;
  }
  count(ETN_CTR_HOST_CALLS); count(ETN_CTR_HOST_ROWS, batch);
  etn_rt::observe(ETN_HIST_BATCH_ROWS, (double)batch);
  cudaStream_t s = (cudaStream_t)stream;
  const size_t bytes = (size_t)(batch * k_bytes);
  // Nothing in the array has to start at zero: the copy below overwrites exactly the
  // [batch x kk] region that the kernel then reads and rewrites in place.
  if (in_pitch_bytes == k_bytes) {
    e = cudaMemcpyAsync(st.ptr, host_in, bytes, cudaMemcpyHostToDevice, s);
  } else {
    e = cudaMemcpy2DAsync(st.ptr, (size_t)k_bytes, host_in, (size_t)in_pitch_bytes, (size_t)k_bytes,
                          (size_t)batch, cudaMemcpyHostToDevice, s);
  }
  if (e != cudaSuccess) return finish((int)e);
  int rc = 0;
  if (mode == ETN_TRUNCATE_RENORM) rc = run_dense(batch, kk, kk, dtype, mode, st.ptr, st.ptr, s);
  if (rc == 0) {
    e = cudaMemcpyAsync(host_out, st.ptr, bytes, cudaMemcpyDeviceToHost, s);
    if (e != cudaSuccess) rc = (int)e;
  }
  e = cudaStreamSynchronize(s);
  if (e != cudaSuccess && rc == 0) rc = (int)e;
  if (rc == 0) { count(ETN_CTR_H2D_BYTES, (int64_t)bytes); count(ETN_CTR_D2H_BYTES, (int64_t)bytes); }
  return finish(rc);
}

int etn_get_counters(int64_t* out, int capacity) {
  if (!out || capacity < 0) return ETN_ERR_NULL;
  const int n = capacity < ETN_CTR_COUNT ? capacity : ETN_CTR_COUNT;
  for (int i = 0; i < n; ++i) out[i] = g_ctr[i].load(std::memory_order_relaxed);
  return n;
}

const char* etn_counter_name(int i) {
  static const char* names[ETN_CTR_COUNT] = {
      "llmcore_embed_gpu_device_calls_total", "llmcore_embed_gpu_device_rows_total",
      "llmcore_embed_gpu_kernel_launches_total", "llmcore_embed_gpu_host_calls_total",
      "llmcore_embed_gpu_host_rows_total", "llmcore_embed_gpu_h2d_bytes_total",
      "llmcore_embed_gpu_d2h_bytes_total", "llmcore_embed_gpu_errors_total",
      "llmcore_embed_gpu_fallback_total", "llmcore_embed_gpu_coalesced_requests_total"};
  return (i >= 0 && i < ETN_CTR_COUNT) ? names[i] : "";
}

void etn_note_fallback(int64_t vectors) {
  if (vectors > 0) count(ETN_CTR_FALLBACKS, vectors);
}

int etn_get_histogram(int which, double* upper_bounds, int64_t* counts, int capacity, double* sum) {
  if (which < 0 || which >= ETN_HIST_COUNT) return ETN_ERR_ENUM;
  const Hist& h = g_hist[which];
  const int n = h.n + 1;
  if (!upper_bounds && !counts) return n;
  if (!upper_bounds || !counts) return ETN_ERR_NULL;
  if (capacity < n) return ETN_ERR_SHAPE;
  for (int i = 0; i < n; ++i) {
    upper_bounds[i] = i < h.n ? h.le[i] : INFINITY;
    counts[i] = h.cnt[i].load(std::memory_order_relaxed);
  }
  if (sum) *sum = (double)h.sum_fixed.load(std::memory_order_relaxed) / 1073741824.0;
  return n;
}

const char* etn_histogram_name(int which) {
  return (which >= 0 && which < ETN_HIST_COUNT) ? g_hist[which].name : "";
}

const char* etn_histogram_help(int which) {
  return (which >= 0 && which < ETN_HIST_COUNT) ? g_hist[which].help : "";
}

int etn_fill_synthetic(uint64_t seed, int64_t row0, int64_t rows, int64_t D, int64_t ncols,
                       int64_t ld, int dtype, void* dev_out, void* stream) {
  if (dtype != ETN_F32 && dtype != ETN_F64) return ETN_ERR_ENUM;
  if (rows < 0 || D <= 0 || ncols < 0 || ncols > D || ld < ncols) return ETN_ERR_SHAPE;
  if (rows == 0 || ncols == 0) return ETN_OK;
  if (!dev_out) return ETN_ERR_NULL;
  const int64_t total = rows * ncols;
  const int grid = (int)std::min<int64_t>((total + 255) / 256, 148 * 16);
  cudaStream_t s = (cudaStream_t)stream;
  if (dtype == ETN_F32) fill_kernel<float><<<grid, 256, 0, s>>>((float*)dev_out, seed, row0, rows, D, ncols, ld);
  else fill_kernel<double><<<grid, 256, 0, s>>>((double*)dev_out, seed, row0, rows, D, ncols, ld);
  return (int)cudaGetLastError();
}

}  // extern "C"
