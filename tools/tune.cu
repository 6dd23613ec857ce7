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
static const char* g_filter = nullptr;
static bool want(const std::string& n) { return !g_filter || n.find(g_filter) != std::string::npos; }

template <typename T, int VB, int NV, int R, int MODE, int MINB, int HINT, int L = 32>
void run_reg(const Shape& s, const void* in, void* out, Timer& tm, int waves = 1) {
  char name[128];
  snprintf(name, sizeof name, "%s reg<VB=%d,NV=%d,R=%d,L=%d,mode=%d,minb=%d,hint=%d,w=%d>", s.name, VB, NV, R, L, MODE, MINB, HINT, waves);
  if (!want(name)) return;
  auto kern = rows_reg_kernel<T, VB, NV, R, L, MODE, MINB, HINT, false>;
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, 0));
  cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
  const int64_t rpb = kWarpsPerBlock * R * (32 / L);
  const int64_t need = (s.rows + rpb - 1) / rpb;
  const int grid = waves > 0 ? (int)std::min<int64_t>(need, (int64_t)g_sms * per_sm * waves) : (int)need;
  const int nvec = (int)(s.k * s.es / VB);
  double ms = tm.ms([&](int) { kern<<<grid, kThreads>>>((const char*)in, (char*)out, s.rows, s.D * s.es, nvec, s.D, nullptr, nullptr); }, 3, 20);
  CK(cudaGetLastError());
  double gbs = (double)s.rows * 2 * s.k * s.es / ms / 1e6;
  printf("%-78s %8.3f ms %8.1f GB/s  regs=%d occ=%d grid=%d\n", name, ms, gbs, fa.numRegs, per_sm, grid);
}

template <typename T, int VB, int NV, int G, int MODE, int MINB, int HINT>
void run_coop(const Shape& s, const void* in, void* out, Timer& tm, int waves = 0) {
  char name[128];
  snprintf(name, sizeof name, "%s coop<VB=%d,NV=%d,G=%d,mode=%d,minb=%d,hint=%d,w=%d>", s.name, VB, NV, G, MODE, MINB, HINT, waves);
  if (!want(name)) return;
  auto kern = rows_coop_kernel<T, VB, NV, G, MODE, MINB, HINT>;
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, 0));
  cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
  const int64_t need = (s.rows + kWarpsPerBlock / G - 1) / (kWarpsPerBlock / G);
  const int grid = waves > 0 ? (int)std::min<int64_t>(need, (int64_t)g_sms * per_sm * waves) : (int)need;
  const int nvec = (int)(s.k * s.es / VB);
  double ms = tm.ms([&](int) { kern<<<grid, kThreads>>>((const char*)in, (char*)out, s.rows, s.D * s.es, nvec); }, 3, 20);
  CK(cudaGetLastError());
  double gbs = (double)s.rows * 2 * s.k * s.es / ms / 1e6;
  printf("%-78s %8.3f ms %8.1f GB/s  regs=%d occ=%d grid=%d\n", name, ms, gbs, fa.numRegs, per_sm, grid);
}

int main(int argc, char** argv) {
  if (argc > 1) g_filter = argv[1];
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  g_sms = p.multiProcessorCount;
  printf("device %s, %d SMs, L2 %d MiB\n", p.name, g_sms, p.l2CacheSize >> 20);
  Timer tm;

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
      {"K64f64 2Mx4096->64", 2097152, 4096, 64, 8},
      {"K32f64 2Mx4096->32", 2097152, 4096, 32, 8},
      {"K128f32 2Mx768->128", 2097152, 768, 128, 4},
      {"K64f32 4Mx768->64", 4194304, 768, 64, 4},
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
    if (s.es == 8 && kb == 512) {
      run_reg<double, 32, 1, 2, 1, 4, 1, 32>(s, in, out, tm, 0);
      run_reg<double, 32, 1, 2, 1, 4, 1, 16>(s, in, out, tm, 0);
      run_reg<double, 32, 1, 4, 1, 4, 1, 16>(s, in, out, tm, 0);
      run_reg<double, 32, 1, 4, 1, 3, 1, 16>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 4, 0, 4, 1, 16>(s, in, out, tm, 0);
    } else if (s.es == 8 && kb == 256) {
      run_reg<double, 32, 1, 2, 1, 4, 1, 16>(s, in, out, tm, 0);
      run_reg<double, 32, 1, 2, 1, 4, 1, 8>(s, in, out, tm, 0);
      run_reg<double, 32, 1, 4, 1, 4, 1, 8>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 4, 0, 4, 1, 8>(s, in, out, tm, 0);
    } else if (s.es == 4 && kb == 512) {
      run_reg<float, 32, 1, 4, 1, 4, 1, 32>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 4, 1, 4, 1, 16>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 8, 1, 2, 1, 16>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 4, 0, 4, 1, 16>(s, in, out, tm, 0);
    } else if (s.es == 4 && kb == 256) {
      run_reg<float, 32, 1, 4, 1, 4, 1, 16>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 4, 1, 4, 1, 8>(s, in, out, tm, 0);
      run_reg<float, 32, 1, 4, 0, 4, 1, 8>(s, in, out, tm, 0);
    } else if (s.es == 4 && kb == 1024) {
      run_reg<float, 32, 1, 4, 1, 4, 1, 32>(s, in, out, tm, 0);
      run_reg<float, 32, 2, 4, 1, 4, 1, 16>(s, in, out, tm, 0);
      run_reg<float, 32, 2, 2, 1, 4, 1, 16>(s, in, out, tm, 0);
    }
    CK(cudaFree(in)); CK(cudaFree(out));
  }
  return 0;
}
