// tune.cu — kernel-variant timing harness (development tool, not part of the product).
// Builds against llm-mcp_b200/csrc/etn_kernels.cuh directly, times each variant with CUDA
// events on inputs larger than L2, and prints achieved algorithmic GB/s
// (rows * 2 * k * sizeof(T) / t) next to two ceilings measured in the same process:
// cudaMemcpyAsync device-to-device and a plain 16-byte grid-stride copy kernel.
//
//   make -C tools && gpurun -- ./tools/tune [filter]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include <chrono>
#include "../llm-mcp_b200/csrc/etn_kernels.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

using namespace etn;

__global__ void __launch_bounds__(256) copy16_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, size_t n) {
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    out[i] = in[i];
  }
}

struct Timer {
  cudaEvent_t a, b;
  Timer() { CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); }
  template <typename F> double ms(F f, int warm, int iters) {
    for (int i = 0; i < warm; ++i) f(i);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a));
    for (int i = 0; i < iters; ++i) f(i);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float t; CK(cudaEventElapsedTime(&t, a, b));
    return t / iters;
  }
};

struct Shape { const char* name; int64_t rows, D, k; int es; };

static int g_sms = 148;
static int g_iters = 20;
static const char* g_filter = nullptr;
static bool want(const std::string& n) { return !g_filter || n.find(g_filter) != std::string::npos; }

template <typename T, int VB, int NV, int R, int MODE, int MINB, int HINT, int L = 32>
void run_reg(const Shape& s, const void* in, void* out, Timer& tm, int waves = 1) {
  char name[128];
  snprintf(name, sizeof name, "%s reg<VB=%d,NV=%d,R=%d,L=%d,mode=%d,minb=%d,hint=%d,w=%d>", s.name, VB, NV, R, L, MODE, MINB, HINT, waves);
  if (!want(name)) return;
  auto kern = rows_reg_kernel<T, VB, NV, R, L, MODE, MINB, HINT>;
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, 0));
  cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
  const int64_t rpb = kWarpsPerBlock * R * (32 / L);
  const int64_t need = (s.rows + rpb - 1) / rpb;
  const int grid = waves > 0 ? (int)std::min<int64_t>(need, (int64_t)g_sms * per_sm * waves) : (int)need;
  const int nvec = (int)(s.k * s.es / VB);
  double ms = tm.ms([&](int) { kern<<<grid, kThreads>>>((const char*)in, (char*)out, s.rows, s.D * s.es, nvec); }, 3, g_iters);
  CK(cudaGetLastError());
  double gbs = (double)s.rows * 2 * s.k * s.es / ms / 1e6;
  printf("%-78s %8.3f ms %8.1f GB/s  regs=%d occ=%d grid=%d\n", name, ms, gbs, fa.numRegs, per_sm, grid);
}

// Request-sized launches (C2: 65,536 x 768 -> 256 f32, 128 MiB per launch): rotate over NSET
// buffer pairs so nothing is served from L2, launch with programmatic dependent launch.
template <typename T, int VB, int NV, int R, int MODE, int MINB, int L>
void run_reg_small(const char* label, int64_t rows, int64_t D, int64_t k, int es, Timer& tm, bool pdl, int threads = kThreads,
                   int waves = 0) {
  constexpr int NSET = 6;
  static void* in[NSET]; static void* out[NSET]; static int64_t have = 0;
  if (have != rows * D) {
    for (int i = 0; i < NSET; ++i) {
      if (in[i]) { cudaFree(in[i]); cudaFree(out[i]); }
      CK(cudaMalloc(&in[i], (size_t)rows * D * es)); CK(cudaMalloc(&out[i], (size_t)rows * k * es));
      fill_kernel<float><<<g_sms * 16, 256>>>((float*)in[i], 3 + i, 0, rows, D * es / 4, D * es / 4, D * es / 4);
    }
    have = rows * D;
  }
  auto kern = rows_reg_kernel<T, VB, NV, R, L, MODE, MINB, 1>;
  const int64_t rpb = (threads / 32) * R * (32 / L);
  int grid = (int)((rows + rpb - 1) / rpb);
  if (waves > 0) {                                      // persistent: whole waves of resident blocks, grid-stride loop
    int per_sm = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, threads, 0));
    grid = std::min(grid, g_sms * per_sm * waves);
  }
  const int nvec = (int)(k * es / VB);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
  double ms = tm.ms([&](int i) {
    CK(cudaLaunchKernelEx(&cfg, kern, (const char*)in[i % NSET], (char*)out[i % NSET], rows, D * es, nvec));
  }, 10, 200);
  printf("%-20s reg<VB=%d,NV=%d,R=%d,L=%d,minb=%d,thr=%d> pdl=%d grid=%d: %7.2f us  %7.1f GB/s\n", label, VB, NV, R, L, MINB, threads, (int)pdl, grid,
         ms * 1e3, (double)rows * 2 * k * es / ms / 1e6);
}

template <typename T, int NV16, int S, int NCW, int MODE>
void run_bulk(const Shape& s, const void* in, void* out, Timer& tm, const void* ref_out = nullptr, int rpw = 0) {
  char name[128];
  snprintf(name, sizeof name, "%s bulk<NV16=%d,S=%d,NCW=%d,mode=%d,rpw=%d>", s.name, NV16, S, NCW, MODE, rpw);
  if (!want(name)) return;
  auto kern = rows_bulk_kernel<T, NV16, S, NCW, MODE>;
  const size_t smem = (size_t)NCW * S * NV16 * 512 + (size_t)NCW * S * 8;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
  const int nvec16 = (int)(s.k * s.es / 16);
  const size_t out_bytes = (size_t)s.rows * s.k * s.es;
  const int grid = rpw > 0 ? (int)((s.rows + (int64_t)NCW * rpw - 1) / ((int64_t)NCW * rpw)) : g_sms;
  CK(cudaMemset(out, 0, out_bytes));
  double ms = tm.ms([&](int) { kern<<<grid, NCW * 32, smem>>>((const char*)in, (char*)out, s.rows, s.D * s.es, nvec16, rpw); }, 3, g_iters);
  CK(cudaGetLastError());
  double gbs = (double)s.rows * 2 * s.k * s.es / ms / 1e6;
  int same = -1;
  if (ref_out) {
    std::vector<char> a(out_bytes), b(out_bytes);
    CK(cudaMemcpy(a.data(), out, out_bytes, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(b.data(), ref_out, out_bytes, cudaMemcpyDeviceToHost));
    same = memcmp(a.data(), b.data(), out_bytes) == 0;
  }
  printf("%-78s %8.3f ms %8.1f GB/s  regs=%d smem=%zuK same_as_reg=%d\n", name, ms, gbs, fa.numRegs, smem >> 10, same);
}

template <typename T, int MODE>
void run_wide(const Shape& s, const void* in, void* out, Timer& tm, int rpb) {
  char name[128];
  snprintf(name, sizeof name, "%s wide<mode=%d,rpb=%d>", s.name, MODE, rpb);
  if (!want(name)) return;
  auto kern = rows_wide_kernel<T, MODE>;
  const int nvec16 = (int)(s.k * s.es / 16);
  const size_t stride = ((size_t)nvec16 * 16 + 127) & ~(size_t)127;
  const size_t smem = 2 * stride;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem));
  cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
  const int grid = (int)((s.rows + rpb - 1) / rpb);
  double ms = tm.ms([&](int) { kern<<<grid, kThreads, smem>>>((const char*)in, (char*)out, s.rows, s.D * s.es, nvec16, rpb); }, 3, g_iters);
  CK(cudaGetLastError());
  double gbs = (double)s.rows * 2 * s.k * s.es / ms / 1e6;
  printf("%-78s %8.3f ms %8.1f GB/s  regs=%d occ=%d smem=%zuK\n", name, ms, gbs, fa.numRegs, per_sm, smem >> 10);
}

// max relative difference between two outputs (different summation orders differ in the last bit)
template <typename T> double max_rel_diff(const void* a, const void* b, size_t n) {
  std::vector<T> ha(n), hb(n);
  CK(cudaMemcpy(ha.data(), a, n * sizeof(T), cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hb.data(), b, n * sizeof(T), cudaMemcpyDeviceToHost));
  double m = 0;
  for (size_t i = 0; i < n; ++i) { double d = fabs((double)ha[i] - (double)hb[i]) / (fabs((double)hb[i]) + 1e-300); if (d > m) m = d; }
  return m;
}

// C1-sized launches (1,024 x 4096 f64 -> 1024, 16 MiB per launch): bulk ring (one row per warp)
// against the wide-row kernel (a block per row), rotating buffers, programmatic dependent launch.
static void c1_study(Timer& tm) {
  const int64_t rows = 1024, D = 4096, k = 1024; const int es = 8, NSET = 8;
  void* in[NSET]; void* out[NSET];
  for (int i = 0; i < NSET; ++i) {
    CK(cudaMalloc(&in[i], (size_t)rows * D * es)); CK(cudaMalloc(&out[i], (size_t)rows * k * es));
    fill_kernel<double><<<g_sms * 16, 256>>>((double*)in[i], 3 + i, 0, rows, D, D, D);
  }
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  const int nvec16 = (int)(k * es / 16);
  auto report = [&](const char* nm, double ms) { printf("%-60s %7.2f us  %7.1f GB/s\n", nm, ms * 1e3, (double)rows * 2 * k * es / ms / 1e6); };
  {
    auto kern = rows_bulk_kernel<double, 16, 2, 4, 1>;
    const size_t smem = (size_t)4 * 2 * 16 * 512 + 4 * 2 * 8;
    CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int rpw : {1, 2}) {
      cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3((unsigned)((rows + 4 * rpw - 1) / (4 * rpw))); cfg.blockDim = dim3(128);
      cfg.dynamicSmemBytes = smem; cfg.attrs = attr; cfg.numAttrs = 1;
      double ms = tm.ms([&](int i) { CK(cudaLaunchKernelEx(&cfg, kern, (const char*)in[i % NSET], (char*)out[i % NSET], rows, D * es, nvec16, rpw)); }, 10, 400);
      char nm[96]; snprintf(nm, sizeof nm, "C1 bulk<double,16,2,4> rpw=%d grid=%u", rpw, cfg.gridDim.x); report(nm, ms);
    }
  }
  {
    auto kern = rows_wide_kernel<double, 1>;
    const size_t smem = 2 * (size_t)8192;
    for (int rpb : {1, 2}) {
      cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3((unsigned)((rows + rpb - 1) / rpb)); cfg.blockDim = dim3(kThreads);
      cfg.dynamicSmemBytes = smem; cfg.attrs = attr; cfg.numAttrs = 1;
      double ms = tm.ms([&](int i) { CK(cudaLaunchKernelEx(&cfg, kern, (const char*)in[i % NSET], (char*)out[i % NSET], rows, D * es, nvec16, rpb)); }, 10, 400);
      char nm[96]; snprintf(nm, sizeof nm, "C1 wide<double> rpb=%d grid=%u", rpb, cfg.gridDim.x); report(nm, ms);
    }
  }
  {
    auto kern = rows_reg_kernel<double, 32, 8, 1, 32, 1, 2, 1>;
    cudaLaunchConfig_t cfg = {}; cfg.gridDim = dim3((unsigned)(rows / 8)); cfg.blockDim = dim3(kThreads); cfg.attrs = attr; cfg.numAttrs = 1;
    double ms = tm.ms([&](int i) { CK(cudaLaunchKernelEx(&cfg, kern, (const char*)in[i % NSET], (char*)out[i % NSET], rows, D * es, (int)(k * es / 32))); }, 10, 400);
    report("C1 reg<double,32,8,1> grid=128", ms);
  }
  for (int i = 0; i < NSET; ++i) { cudaFree(in[i]); cudaFree(out[i]); }
}

// PCIe study for the host path: DMA ceilings, prefix-only 2-D DMA, and the kernels reading and
// writing pinned host memory directly (zero copy).  C3 slice: rows x 4096 f32 -> 1024.
static void pcie_study(Timer& tm) {
  const int64_t rows = 262144, D = 4096, k = 1024, es = 4;
  const size_t in_bytes = (size_t)rows * D * es, pre_bytes = (size_t)rows * k * es;
  char *hin, *hout, *din, *dout;
  CK(cudaHostAlloc(&hin, in_bytes, cudaHostAllocPortable));
  CK(cudaHostAlloc(&hout, pre_bytes, cudaHostAllocPortable));
  CK(cudaMalloc(&din, in_bytes)); CK(cudaMalloc(&dout, pre_bytes));
  fill_kernel<float><<<g_sms * 16, 256>>>((float*)din, 3, 0, rows, D, D, D);
  CK(cudaMemcpy(hin, din, in_bytes, cudaMemcpyDeviceToHost));
  cudaStream_t s1, s2; CK(cudaStreamCreate(&s1)); CK(cudaStreamCreate(&s2));
  auto rep = [&](const char* nm, double ms, double bytes) { printf("%-64s %8.3f ms %8.1f GB/s\n", nm, ms, bytes / ms / 1e6); };
  double ms;
  ms = tm.ms([&](int) { CK(cudaMemcpyAsync(din, hin, pre_bytes, cudaMemcpyHostToDevice)); }, 1, 5);
  rep("pcie H2D 1 GiB contiguous", ms, pre_bytes);
  ms = tm.ms([&](int) { CK(cudaMemcpyAsync(hout, dout, pre_bytes, cudaMemcpyDeviceToHost)); }, 1, 5);
  rep("pcie D2H 1 GiB contiguous", ms, pre_bytes);
  ms = tm.ms([&](int) { CK(cudaMemcpy2DAsync(din, k * es, hin, D * es, k * es, rows, cudaMemcpyHostToDevice)); }, 1, 5);
  rep("pcie H2D 2-D prefix (4 KiB of every 16 KiB)", ms, pre_bytes);
  {
    cudaEvent_t a, b, c; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); CK(cudaEventCreate(&c));
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a, s1)); CK(cudaStreamWaitEvent(s2, a));
    for (int i = 0; i < 5; ++i) {
      CK(cudaMemcpyAsync(din, hin, pre_bytes, cudaMemcpyHostToDevice, s1));
      CK(cudaMemcpyAsync(hout, dout, pre_bytes, cudaMemcpyDeviceToHost, s2));
    }
    CK(cudaEventRecord(c, s2)); CK(cudaStreamWaitEvent(s1, c)); CK(cudaEventRecord(b, s1));
    CK(cudaEventSynchronize(b));
    float t; CK(cudaEventElapsedTime(&t, a, b));
    rep("pcie H2D + D2H concurrently (each direction)", t / 5, pre_bytes);
  }
  const int nvec = (int)(k * es / 32);
  auto kern = rows_reg_kernel<float, 32, 4, 1, 32, 1, 4, 1>;
  const int grid = (int)((rows + 7) / 8);
  ms = tm.ms([&](int) { kern<<<grid, kThreads>>>(hin, hout, rows, D * es, nvec); }, 1, 5);
  rep("zero-copy kernel: host in -> host out (per direction)", ms, pre_bytes);
  ms = tm.ms([&](int) { kern<<<grid, kThreads>>>(hin, dout, rows, D * es, nvec); }, 1, 5);
  rep("zero-copy kernel: host in -> device out", ms, pre_bytes);
  ms = tm.ms([&](int) { kern<<<grid, kThreads>>>(din, hout, rows, D * es, nvec); }, 1, 5);
  rep("zero-copy kernel: device in -> host out", ms, pre_bytes);
  auto kern16 = rows_reg_kernel<float, 16, 8, 1, 32, 1, 4, 0>;
  ms = tm.ms([&](int) { kern16<<<grid, kThreads>>>(hin, hout, rows, D * es, (int)(k * es / 16)); }, 1, 5);
  rep("zero-copy kernel 128-bit: host in -> host out", ms, pre_bytes);
  for (int g : {148, 296, 592, 2368}) {
    char nm[96]; snprintf(nm, sizeof nm, "zero-copy kernel grid=%d: host in -> host out", g);
    ms = tm.ms([&](int) { kern<<<g, kThreads>>>(hin, hout, rows, D * es, nvec); }, 1, 5);
    rep(nm, ms, pre_bytes);
  }
  // hybrid: the kernel reads the host input itself (no H2D DMA) while the copy engine drains
  // the previous chunk's output D2H
  {
    cudaEvent_t a, b, c; CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); CK(cudaEventCreate(&c));
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(a, s1)); CK(cudaStreamWaitEvent(s2, a));
    for (int i = 0; i < 5; ++i) {
      kern<<<grid, kThreads, 0, s1>>>(hin, dout, rows, D * es, nvec);
      CK(cudaMemcpyAsync(hout, din, pre_bytes, cudaMemcpyDeviceToHost, s2));
    }
    CK(cudaEventRecord(c, s2)); CK(cudaStreamWaitEvent(s1, c)); CK(cudaEventRecord(b, s1));
    CK(cudaEventSynchronize(b));
    float t; CK(cudaEventElapsedTime(&t, a, b));
    rep("hybrid: zero-copy read kernel || D2H DMA (each direction)", t / 5, pre_bytes);
    CK(cudaEventRecord(a, s1)); CK(cudaStreamWaitEvent(s2, a));
    for (int i = 0; i < 5; ++i) {
      CK(cudaMemcpy2DAsync(din, k * es, hin, D * es, k * es, rows, cudaMemcpyHostToDevice, s1));
      kern<<<grid, kThreads, 0, s2>>>(dout, hout, rows, k * es, nvec);
    }
    CK(cudaEventRecord(c, s2)); CK(cudaStreamWaitEvent(s1, c)); CK(cudaEventRecord(b, s1));
    CK(cudaEventSynchronize(b));
    CK(cudaEventElapsedTime(&t, a, b));
    rep("hybrid: H2D DMA || zero-copy write kernel (each direction)", t / 5, pre_bytes);
  }
  // small request latency: C1 1024 x 4096 f64 -> 1024 (8 MiB each way)
  {
    auto k64 = rows_reg_kernel<double, 32, 8, 1, 32, 1, 2, 1>;
    ms = tm.ms([&](int) { k64<<<128, kThreads>>>(hin, hout, 1024, 4096 * 8, 256); }, 2, 20);
    rep("C1 zero-copy kernel host->host (8 MiB each way)", ms, 8.0 * 1048576);
    ms = tm.ms([&](int) {
      CK(cudaMemcpy2DAsync(din, 8192, hin, 32768, 8192, 1024, cudaMemcpyHostToDevice));
      k64<<<128, kThreads>>>(din, dout, 1024, 8192, 256);
      CK(cudaMemcpyAsync(hout, dout, 8388608, cudaMemcpyDeviceToHost)); }, 2, 20);
    rep("C1 staged H2D(2-D) + kernel + D2H (8 MiB each way)", ms, 8.0 * 1048576);
  }
  // C1 through chunked pipelines: wall-clock per call, n chunks round-robin over 4 streams
  {
    auto k64 = rows_reg_kernel<double, 32, 8, 1, 32, 1, 2, 1>;
    cudaStream_t st[4];
    for (auto& x : st) CK(cudaStreamCreateWithFlags(&x, cudaStreamNonBlocking));
    for (int nch : {1, 2, 4, 8}) {
      for (int twod : {1, 0}) {
        const int rows_c = 1024 / nch;
        auto once = [&]() {
          for (int c = 0; c < nch; ++c) {
            cudaStream_t s = st[c % 4];
            char* d = din + (size_t)c * rows_c * 8192;
            if (twod) CK(cudaMemcpy2DAsync(d, 8192, hin + (size_t)c * rows_c * 32768, 32768, 8192, rows_c, cudaMemcpyHostToDevice, s));
            else CK(cudaMemcpyAsync(d, hin + (size_t)c * rows_c * 8192, (size_t)rows_c * 8192, cudaMemcpyHostToDevice, s));
            k64<<<(rows_c + 7) / 8, kThreads, 0, s>>>(d, d, rows_c, 8192, 256);
            CK(cudaMemcpyAsync(hout + (size_t)c * rows_c * 8192, d, (size_t)rows_c * 8192, cudaMemcpyDeviceToHost, s));
          }
          for (auto& x : st) CK(cudaStreamSynchronize(x));
        };
        for (int i = 0; i < 5; ++i) once();
        auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < 50; ++i) once();
        double us = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count() / 50;
        printf("C1 pipeline %d chunk(s), %s H2D: %.1f us per call\n", nch, twod ? "2-D" : "1-D", us);
      }
    }
  }
  CK(cudaFreeHost(hin)); CK(cudaFreeHost(hout)); CK(cudaFree(din)); CK(cudaFree(dout));
}

// Page-ring kernel (ragged / gathered rows): ring pages per warp x resident blocks, on the
// config-5 length mix and on 64 merged C1 requests (uniform 8 KiB float64 rows).
template <int P, int MINB>
void ring_variant(const char* label, const char* in, char* out, int64_t rows, int64_t D, const int32_t* d_k,
                  const int64_t* d_off, double algo_bytes, Timer& tm) {
  constexpr int NCW = 4;
  auto kern = rows_ring_kernel<double, NCW, kModeRenorm, P, MINB>;
  const size_t smem = (size_t)NCW * P * 512 + (size_t)NCW * 8 * 8 + (size_t)NCW * 8 * 3 * 4;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, NCW * 32, smem));
  for (int rpw : {4, 8, 16, 32}) {
    const int64_t blocks = (rows + (int64_t)NCW * rpw - 1) / ((int64_t)NCW * rpw);
    double ms = tm.ms([&](int) {
      kern<<<(unsigned)blocks, NCW * 32, smem>>>(in, out, rows, D * 8, rpw, D, 0, d_k, d_off, nullptr);
    }, 3, g_iters);
    CK(cudaGetLastError());
    printf("%-28s ring P=%2d (%2d KiB/warp) minb=%d occ=%d rpw=%2d %8.3f ms %8.1f GB/s\n", label, P, P / 2, MINB, per_sm, rpw, ms,
           algo_bytes / ms / 1e6);
  }
}

void ring_study(Timer& tm) {
  const int64_t D = 4096;
  for (int which = 0; which < 2; ++which) {
    const int64_t rows = which == 0 ? 1048576 : 65536;
    std::vector<int32_t> k(rows);
    std::vector<int64_t> off(rows);
    uint64_t st = 5;
    auto rnd = [&] { st = st * 6364136223846793005ull + 1442695040888963407ull; return (uint32_t)(st >> 33); };
    const int ks[5] = {64, 128, 256, 512, 1024};
    int64_t r = 0, acc = 0;
    while (r < rows) {
      int64_t m = which == 0 ? 1 + rnd() % 2048 : rows;
      const int kk = which == 0 ? ks[rnd() % 5] : 1024;
      for (; m > 0 && r < rows; --m, ++r) { k[r] = kk; off[r] = acc; acc += kk; }
    }
    char *in, *out; int32_t* d_k; int64_t* d_off;
    CK(cudaMalloc(&in, rows * D * 8)); CK(cudaMalloc(&out, acc * 8));
    CK(cudaMalloc(&d_k, rows * 4)); CK(cudaMalloc(&d_off, rows * 8));
    fill_kernel<double><<<148 * 16, 256>>>((double*)in, 5, 0, rows, D, D, D);
    CK(cudaMemcpy(d_k, k.data(), rows * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(d_off, off.data(), rows * 8, cudaMemcpyHostToDevice));
    CK(cudaDeviceSynchronize());
    const char* label = which == 0 ? "C5 mix 1Mi rows" : "64 x C1 (8 KiB rows)";
    const double algo = 16.0 * acc;
    ring_variant<16, 6>(label, in, out, rows, D, d_k, d_off, algo, tm);
    ring_variant<16, 4>(label, in, out, rows, D, d_k, d_off, algo, tm);
    ring_variant<32, 3>(label, in, out, rows, D, d_k, d_off, algo, tm);
    ring_variant<24, 4>(label, in, out, rows, D, d_k, d_off, algo, tm);
    ring_variant<48, 2>(label, in, out, rows, D, d_k, d_off, algo, tm);
    CK(cudaFree(in)); CK(cudaFree(out)); CK(cudaFree(d_k)); CK(cudaFree(d_off));
  }
}

int main(int argc, char** argv) {
  if (argc > 1) g_filter = argv[1];
  if (getenv("ITERS")) g_iters = atoi(getenv("ITERS"));
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  g_sms = p.multiProcessorCount;
  printf("device %s, %d SMs, L2 %d MiB\n", p.name, g_sms, p.l2CacheSize >> 20);
  Timer tm;
  if (g_filter && !strcmp(g_filter, "pcie")) { pcie_study(tm); return 0; }
  if (g_filter && !strcmp(g_filter, "c1")) { c1_study(tm); return 0; }
  if (g_filter && !strcmp(g_filter, "ring")) { ring_study(tm); return 0; }
  if (g_filter && !strcmp(g_filter, "small")) {
    for (bool pdl : {false, true}) {
      run_reg_small<float, 32, 1, 4, 1, 4, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl);
      run_reg_small<float, 32, 1, 2, 1, 4, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl);
      run_reg_small<float, 32, 1, 2, 1, 6, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl);
      run_reg_small<float, 32, 1, 8, 1, 2, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl);
      run_reg_small<float, 32, 1, 4, 1, 4, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl, 128);
      run_reg_small<float, 32, 1, 4, 1, 4, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl, 64);
      run_reg_small<float, 16, 2, 4, 1, 4, 32>("C2 64Kx768->256", 65536, 768, 256, 4, tm, pdl);
      // whole waves of resident blocks, rows dealt grid-stride (finer quanta with fewer rows per warp)
      run_reg_small<float, 32, 1, 1, 1, 8, 32>("C2 persistent", 65536, 768, 256, 4, tm, pdl, kThreads, 1);
      run_reg_small<float, 32, 1, 2, 1, 6, 32>("C2 persistent", 65536, 768, 256, 4, tm, pdl, kThreads, 1);
      run_reg_small<float, 32, 1, 4, 1, 4, 32>("C2 persistent", 65536, 768, 256, 4, tm, pdl, kThreads, 1);
      run_reg_small<float, 32, 1, 1, 1, 8, 32>("C2 one-shot R=1", 65536, 768, 256, 4, tm, pdl);
      run_reg_small<float, 32, 1, 2, 1, 8, 32>("C2 one-shot R=2 minb8", 65536, 768, 256, 4, tm, pdl);
    }
    return 0;
  }

  // ceilings: 4 GiB read + 4 GiB written
  {
    size_t bytes = 4ull << 30;
    void *a, *b; CK(cudaMalloc(&a, bytes)); CK(cudaMalloc(&b, bytes));
    CK(cudaMemset(a, 1, bytes));
    double ms = tm.ms([&](int) { CK(cudaMemcpyAsync(b, a, bytes, cudaMemcpyDeviceToDevice)); }, 2, 10);
    printf("%-64s %8.3f ms %8.1f GB/s\n", "ceiling cudaMemcpyAsync D2D 4GiB", ms, 2.0 * bytes / ms / 1e6);
    for (int mult : {8, 16, 32}) {
      char nm[64]; snprintf(nm, sizeof nm, "ceiling copy16 kernel grid=%dxSMs", mult);
      ms = tm.ms([&](int) { copy16_kernel<<<g_sms * mult, 256>>>((const uint4*)a, (uint4*)b, bytes / 16); }, 2, 10);
      printf("%-64s %8.3f ms %8.1f GB/s\n", nm, ms, 2.0 * bytes / ms / 1e6);
    }
    CK(cudaFree(a)); CK(cudaFree(b));
  }

  std::vector<Shape> shapes = {
      {"C3f32 1Mx4096->1024", 1048576, 4096, 1024, 4},
      {"C4f32 2Mx3072->256", 2097152, 3072, 256, 4},
  };
  for (const Shape& s : shapes) {
    void *in, *out;
    CK(cudaMalloc(&in, (size_t)s.rows * s.D * s.es));
    CK(cudaMalloc(&out, (size_t)s.rows * s.k * s.es));
    if (s.es == 4) fill_kernel<float><<<g_sms * 16, 256>>>((float*)in, 3, 0, s.rows, s.D, s.D, s.D);
    else fill_kernel<double><<<g_sms * 16, 256>>>((double*)in, 3, 0, s.rows, s.D, s.D, s.D);
    CK(cudaDeviceSynchronize());
    const int64_t kb = s.k * s.es;
    void* out2; CK(cudaMalloc(&out2, (size_t)s.rows * s.k * s.es));
    if (s.es == 4 && kb == 4096) {
      run_reg<float, 32, 4, 1, 1, 4, 1>(s, in, out, tm, 0);
      run_reg<float, 32, 4, 2, 1, 2, 1>(s, in, out, tm, 0);
      run_bulk<float, 8, 2, 8, 1>(s, in, out2, tm, out, 8);
      run_bulk<float, 8, 2, 8, 1>(s, in, out2, tm, out, 4);
      run_bulk<float, 8, 3, 8, 1>(s, in, out2, tm, out, 8);
      run_bulk<float, 8, 2, 4, 1>(s, in, out2, tm, out, 8);
      run_bulk<float, 8, 2, 4, 1>(s, in, out2, tm, out, 4);
      run_bulk<float, 8, 4, 4, 1>(s, in, out2, tm, out, 8);
      run_bulk<float, 8, 2, 8, 0>(s, in, out2, tm, nullptr, 8);
      run_reg<float, 32, 4, 1, 0, 4, 1>(s, in, out, tm, 0);
    } else if (s.es == 4 && kb == 1024) {
      run_reg<float, 32, 1, 4, 1, 4, 1>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 8, 1, 2, 1>(s, in, out, tm, 0);
      run_bulk<float, 2, 4, 4, 1>(s, in, out2, tm, out, 16);
      run_bulk<float, 2, 4, 8, 1>(s, in, out2, tm, out, 8);
      run_bulk<float, 2, 8, 4, 1>(s, in, out2, tm, out, 16);
      run_bulk<float, 2, 4, 4, 1>(s, in, out2, tm, out, 8);
    }
    CK(cudaFree(out2));
    CK(cudaFree(in)); CK(cudaFree(out));
  }
  return 0;
}
