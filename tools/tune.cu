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

template <typename T, int NV, int R, int MODE, int MINB>
void run_reg(const Shape& s, const void* in, void* out, Timer& tm, int waves = 1) {
  char name[128];
  snprintf(name, sizeof name, "%s reg<NV=%d,R=%d,mode=%d,minb=%d,w=%d>", s.name, NV, R, MODE, MINB, waves);
  if (!want(name)) return;
  auto kern = rows_reg_kernel<T, NV, R, MODE, MINB>;
  int per_sm = 0;
  CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, 0));
  cudaFuncAttributes fa; CK(cudaFuncGetAttributes(&fa, kern));
  const int64_t need = (s.rows + kWarpsPerBlock * R - 1) / (kWarpsPerBlock * R);
  const int grid = (int)std::min<int64_t>(need, (int64_t)g_sms * per_sm * waves);
  const int nvec = (int)(s.k * s.es / 16);
  double ms = tm.ms([&](int) { kern<<<grid, kThreads>>>((const char*)in, (char*)out, s.rows, s.D * s.es, nvec); }, 3, 20);
  CK(cudaGetLastError());
  double gbs = (double)s.rows * 2 * s.k * s.es / ms / 1e6;
  printf("%-64s %8.3f ms %8.1f GB/s  regs=%d occ=%d grid=%d\n", name, ms, gbs, fa.numRegs, per_sm, grid);
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
      {"C3f32 1Mx4096->1024", 1048576, 4096, 1024, 4},
      {"C4f32 2Mx3072->256", 2097152, 3072, 256, 4},
      {"C2x16 1Mx768->256", 1048576, 768, 256, 4},
      {"C3f64 512Kx4096->1024", 524288, 4096, 1024, 8},
      {"C5f64 1Mx4096->256", 1048576, 4096, 256, 8},
  };
  for (const Shape& s : shapes) {
    if (g_filter && !strstr(s.name, g_filter) && !strchr(g_filter, '<') && !strstr(g_filter, "reg")) { /* filter may target variants */ }
    void *in, *out;
    CK(cudaMalloc(&in, (size_t)s.rows * s.D * s.es));
    CK(cudaMalloc(&out, (size_t)s.rows * s.k * s.es));
    if (s.es == 4) fill_kernel<float><<<g_sms * 16, 256>>>((float*)in, 3, 0, s.rows, s.D, s.D, s.D);
    else fill_kernel<double><<<g_sms * 16, 256>>>((double*)in, 3, 0, s.rows, s.D, s.D, s.D);
    CK(cudaDeviceSynchronize());
    const int nvl = (int)((s.k * s.es / 16 + 31) / 32);
    if (s.es == 4) {
      if (nvl == 8) {
        run_reg<float, 8, 1, 1, 4>(s, in, out, tm);
        run_reg<float, 8, 1, 1, 3>(s, in, out, tm);
        run_reg<float, 8, 1, 1, 2>(s, in, out, tm);
        run_reg<float, 8, 1, 1, 4>(s, in, out, tm, 4);
        run_reg<float, 8, 1, 0, 4>(s, in, out, tm);
      } else if (nvl == 2) {
        run_reg<float, 2, 4, 1, 4>(s, in, out, tm);
        run_reg<float, 2, 4, 1, 3>(s, in, out, tm);
        run_reg<float, 2, 2, 1, 4>(s, in, out, tm);
        run_reg<float, 2, 2, 1, 6>(s, in, out, tm);
        run_reg<float, 2, 1, 1, 8>(s, in, out, tm);
        run_reg<float, 2, 8, 1, 2>(s, in, out, tm);
        run_reg<float, 2, 4, 1, 4>(s, in, out, tm, 4);
        run_reg<float, 2, 4, 0, 4>(s, in, out, tm);
      }
    } else {
      if (nvl == 16) {
        run_reg<double, 16, 1, 1, 1>(s, in, out, tm);
        run_reg<double, 16, 1, 1, 2>(s, in, out, tm);
        run_reg<float, 16, 1, 0, 2>(s, in, out, tm);
      } else if (nvl == 4) {
        run_reg<double, 4, 2, 1, 2>(s, in, out, tm);
        run_reg<double, 4, 2, 1, 3>(s, in, out, tm);
        run_reg<double, 4, 1, 1, 4>(s, in, out, tm);
        run_reg<double, 4, 4, 1, 1>(s, in, out, tm);
        run_reg<float, 4, 2, 0, 4>(s, in, out, tm);
      }
    }
    CK(cudaFree(in)); CK(cudaFree(out));
  }
  return 0;
}
