// etn_json_gpu.cu — the number work of the JSON packer on the GPU.
//
// on the reference's path the CPU time is not the truncation (a reslice) but the
// two loops around it — json.Decode turning every number of the provider's body into a boxed
// float64 (core/internal/api/handlers.go:2047-2050) and json.Encode writing every number back
// out (core/internal/api/helpers.go:11-15).  The host packer (etn_json.cpp) already skips all
// numbers that are cut away; for truncate+renormalise it still had to parse and re-format the
// kept ones on the host cores (std::from_chars / std::to_chars, ~30 ms for a C1-sized response
// wrapped around a 7 us kernel).  Here the kept TEXT goes to the device instead:
//
//   host    skeleton scan (unchanged), kept spans copied into pinned memory      ~23 MB for C1
//   H2D     the spans
//   kernel  parse_vectors    one block per vector: span -> shared memory, commas counted and
//                            scanned into token starts, one Eisel-Lemire parse per thread and
//                            number (etn_numfmt.cuh); a number it cannot decide is listed for
//                            the host (correctly rounded from_chars) instead of being guessed
//   kernel  the library's own renormalisation kernels on the [n x dims] matrix (run_dense)
//   kernel  format_vectors   one block per vector: shortest round-trip digits per number
//                            (Schubfach), encoding/json's layout, lengths scanned across the
//                            block, "[a,b,c]" assembled in shared memory and stored coalesced
//   D2H     the texts (fixed stride per vector) and their lengths
//   host    splice (unchanged)
//
// Results are bit-identical to the host path by construction of etn_numfmt.cuh (checked on the
// CPU against std::from_chars/to_chars by tests/test_numfmt.py, on the GPU against the host
// path by tests/test_gpu_json.py).  Hand-written sm_100a kernels, no library.
#include "../../include/embed_truncate_norm.h"
#include "etn_internal.h"
#include "etn_numfmt.cuh"

#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

using etn_num::Pow10Row;

__device__ const Pow10Row d_pow10[] = {
#include "etn_pow10_table.inc"
};

constexpr int kThreads = 256;
constexpr int kNumStride = 26;                       // bytes of text one number can need, comma included
constexpr size_t kMaxSmem = 200 * 1024;
constexpr int kMaxFallback = 4096;                   // numbers per call the host may be asked to redo

__device__ __forceinline__ bool is_ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r'; }

// Exclusive scan of one int per thread over the block; total returned to every thread.
__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_tot /*[kThreads/32 + 1]*/, int& total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int y = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += y;
  }
  if (lane == 31) warp_tot[w] = x;
  __syncthreads();
  if (w == 0) {
    int t = lane < kThreads / 32 ? warp_tot[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, t, o);
      if (lane >= o) t += y;
    }
    if (lane < kThreads / 32) warp_tot[lane] = t;        // inclusive totals of the warps
  }
  __syncthreads();
  total = warp_tot[kThreads / 32 - 1];
  const int before = w ? warp_tot[w - 1] : 0;
  __syncthreads();                                        // warp_tot may be reused by the caller
  return before + x - v;
}

// text: the kept spans back to back, span g at [toff[g], toff[g+1]) (16-byte aligned starts; the
// span itself is tlen[g] bytes).  vals[g * dims + i] <- i-th number of span g.
__global__ void __launch_bounds__(kThreads)
parse_vectors(const char* __restrict__ text, const int64_t* __restrict__ toff, const int32_t* __restrict__ tlen, int dims,
              double* __restrict__ vals, int* __restrict__ fb_count, int64_t* __restrict__ fb_list, int* __restrict__ bad) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int warp_tot[kThreads / 32 + 1];
  const int g = blockIdx.x, tid = threadIdx.x;
  const int L = tlen[g];
  const int Lpad = (L + 15) & ~15;
  char* s = reinterpret_cast<char*>(smem);
  int* tok = reinterpret_cast<int*>(smem + Lpad);        // dims + 1 token starts
  {
    const uint4* src = reinterpret_cast<const uint4*>(text + toff[g]);
    uint4* dst = reinterpret_cast<uint4*>(smem);
    for (int i = tid; i < Lpad / 16; i += kThreads) dst[i] = src[i];
  }
  __syncthreads();
  // commas per thread segment -> token starts
  const int seg = (L + kThreads - 1) / kThreads;
  const int b0 = min(L, tid * seg), b1 = min(L, b0 + seg);
  int mine = 0;
  for (int i = b0; i < b1; ++i) mine += s[i] == ',';
  int total;
  int at = block_exclusive_scan(mine, warp_tot, total);
  if (total != dims - 1) {                               // not dims numbers after all: the host redoes this call
    if (tid == 0) atomicExch(bad, 1);
    return;
  }
  if (tid == 0) { tok[0] = 0; tok[dims] = L + 1; }
  for (int i = b0; i < b1; ++i)
    if (s[i] == ',') tok[++at] = i + 1;
  __syncthreads();
  for (int i = tid; i < dims; i += kThreads) {
    int a = tok[i], e = tok[i + 1] - 1;                  // [a, e): the token with its white space
    while (a < e && is_ws(s[a])) ++a;
    while (e > a && is_ws(s[e - 1])) --e;
    double v = 0.0;
    if (!etn_num::parse_float64(s + a, e - a, d_pow10, v)) {
      const int slot = atomicAdd(fb_count, 1);
      if (slot < kMaxFallback) fb_list[slot] = (int64_t)g * dims + i;
      v = 0.0;
    }
    vals[(int64_t)g * dims + i] = v;
  }
}

// vals[g * dims + i] -> "[a,b,c]" at out + g * ostride, its length in olen[g].
__global__ void __launch_bounds__(kThreads)
format_vectors(const double* __restrict__ vals, int dims, char* __restrict__ out, int64_t ostride, int32_t* __restrict__ olen,
               int* __restrict__ bad) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ int warp_tot[kThreads / 32 + 1];
  const int g = blockIdx.x, tid = threadIdx.x;
  char* text = reinterpret_cast<char*>(smem);            // dims * kNumStride + 2 bytes, rounded up to 16
  const int per = (dims + kThreads - 1) / kThreads;      // consecutive numbers per thread
  const int i0 = min(dims, tid * per), i1 = min(dims, i0 + per);
  // pass 1: lengths (format into registers/local, count)
  int mine = 0;
  for (int i = i0; i < i1; ++i) {
    const double v = vals[(int64_t)g * dims + i];
    if (!(fabs(v) <= 1.7976931348623157e308)) { atomicExch(bad, 2); }     // NaN/Inf: Go's encoder refuses them
    char tmp[32];
    mine += etn_num::format_go_float64(v, d_pow10, tmp) + 1;               // + ',' (or the closing ']')
  }
  int total;
  int at = 1 + block_exclusive_scan(mine, warp_tot, total);
  if (tid == 0) text[0] = '[';
  // pass 2: write (formatting twice is cheaper than keeping `per` strings per thread around)
  for (int i = i0; i < i1; ++i) {
    const double v = vals[(int64_t)g * dims + i];
    at += etn_num::format_go_float64(v, d_pow10, text + at);
    text[at++] = i + 1 == dims ? ']' : ',';
  }
  __syncthreads();
  const int len = total + 1;
  if (tid == 0) olen[g] = len;
  uint4* dst = reinterpret_cast<uint4*>(out + (int64_t)g * ostride);
  const uint4* src = reinterpret_cast<const uint4*>(smem);
  for (int i = tid; i < (len + 15) / 16; i += kThreads) dst[i] = src[i];
}

// ------------------------------------------------------------------ workspace ----
struct Buf {
  void* p = nullptr;
  size_t cap = 0;
  bool host = false;
  int ensure(size_t n) {
    if (n <= cap) return 0;
    if (p) { if (host) cudaFreeHost(p); else cudaFree(p); p = nullptr; cap = 0; }
    const size_t want = n + n / 4 + 4096;
    cudaError_t e = host ? cudaHostAlloc(&p, want, cudaHostAllocPortable) : cudaMalloc(&p, want);
    if (e != cudaSuccess) { cudaGetLastError(); p = nullptr; return (int)e; }
    cap = want;
    return 0;
  }
  void release() {
    if (p) { if (host) cudaFreeHost(p); else cudaFree(p); }
    p = nullptr; cap = 0;
  }
};

struct Workspace {
  std::mutex mu;
  int dev = -1;
  cudaStream_t stream = nullptr;
  Buf d_text, d_toff, d_tlen, d_vals, d_norm, d_out, d_olen, d_flags, d_fb;
  Buf h_text, h_toff, h_tlen, h_out, h_olen, h_flags, h_fb;
  Workspace() { h_text.host = h_toff.host = h_tlen.host = h_out.host = h_olen.host = h_flags.host = h_fb.host = true; }
};
constexpr int kMaxDevices = 64;
Workspace g_ws[kMaxDevices];

template <typename F>
void parallel_for(size_t n, size_t grain, F f) {
  size_t nt = std::min<size_t>(std::min<size_t>(std::thread::hardware_concurrency(), 16), n / std::max<size_t>(grain, 1));
  if (nt <= 1) { for (size_t i = 0; i < n; ++i) f(i); return; }
  std::vector<std::thread> th;
  for (size_t t = 0; t < nt; ++t)
    th.emplace_back([=] { for (size_t i = t; i < n; i += nt) f(i); });
  for (auto& x : th) x.join();
}

}  // namespace

namespace etn_rt {

#define JG_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cudaGetLastError(); return (int)e_; } } while (0)

// See etn_internal.h.  Returns 0 (result valid until json_gpu_release), kJsonGpuUnavailable when
// the caller should use the host path (no device bound, workspace busy, shapes that do not fit
// shared memory, too many numbers the fast parser could not decide), or a CUDA error.
int json_renorm_gpu(const char* json, const int64_t* begin, const int64_t* end, size_t n, int64_t dims, JsonGpuResult* res) {
  res->ws = nullptr;
  if (n == 0 || dims <= 0 || dims > 100000) return kJsonGpuUnavailable;
  const int dev = least_loaded_device();
  if (dev < 0 || dev >= kMaxDevices) return kJsonGpuUnavailable;
  int64_t max_len = 0, total = 0;
  std::vector<int64_t> toff(n + 1);
  for (size_t g = 0; g < n; ++g) {
    const int64_t len = end[g] - begin[g];
    if (len <= 0 || len > (int64_t)kMaxSmem) return kJsonGpuUnavailable;
    max_len = std::max(max_len, len);
    toff[g] = total;
    total += (len + 15) & ~(int64_t)15;
  }
  toff[n] = total;
  const size_t parse_smem = (size_t)((max_len + 15) & ~(int64_t)15) + sizeof(int) * (size_t)(dims + 1);
  const int64_t ostride = (dims * kNumStride + 2 + 15) & ~(int64_t)15;
  if (parse_smem > kMaxSmem || (size_t)ostride > kMaxSmem) return kJsonGpuUnavailable;

  Workspace& w = g_ws[dev];
  std::unique_lock<std::mutex> lock(w.mu, std::try_to_lock);
  if (!lock.owns_lock()) return kJsonGpuUnavailable;        // another handler is using it: host path, no queueing
  int prev = 0;
  const bool have_prev = cudaGetDevice(&prev) == cudaSuccess;
  struct Restore { bool on; int dev; ~Restore() { if (on) cudaSetDevice(dev); } } restore{have_prev, prev};
  JG_TRY(cudaSetDevice(dev));
  if (!w.stream) { JG_TRY(cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking)); w.dev = dev; }
  static std::once_flag attr_once[kMaxDevices];
  int attr_rc = 0;
  std::call_once(attr_once[dev], [&] {
    cudaError_t e = cudaFuncSetAttribute(parse_vectors, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(format_vectors, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMaxSmem);
    if (e != cudaSuccess) { cudaGetLastError(); attr_rc = (int)e; }
  });
  if (attr_rc) return attr_rc;
  const size_t nvals = n * (size_t)dims;
  int rc = 0;
  if ((rc = w.h_text.ensure((size_t)total + 16)) || (rc = w.d_text.ensure((size_t)total + 16)) || (rc = w.h_toff.ensure((n + 1) * 8)) ||
      (rc = w.d_toff.ensure((n + 1) * 8)) || (rc = w.h_tlen.ensure(n * 4)) || (rc = w.d_tlen.ensure(n * 4)) ||
      (rc = w.d_vals.ensure(nvals * 8)) || (rc = w.d_norm.ensure(nvals * 8)) || (rc = w.d_out.ensure(n * (size_t)ostride)) ||
      (rc = w.h_out.ensure(n * (size_t)ostride)) || (rc = w.d_olen.ensure(n * 4)) || (rc = w.h_olen.ensure(n * 4)) ||
      (rc = w.d_flags.ensure(16)) || (rc = w.h_flags.ensure(16)) || (rc = w.d_fb.ensure(kMaxFallback * 8)) ||
      (rc = w.h_fb.ensure(kMaxFallback * 8)))
    return rc;
  // kept spans into pinned memory, on the host cores
  char* ht = (char*)w.h_text.p;
  int32_t* hl = (int32_t*)w.h_tlen.p;
  memcpy(w.h_toff.p, toff.data(), (n + 1) * 8);
  parallel_for(n, 16, [&](size_t g) {
    const int64_t len = end[g] - begin[g];
    memcpy(ht + toff[g], json + begin[g], (size_t)len);
    memset(ht + toff[g] + len, ' ', (size_t)(((len + 15) & ~(int64_t)15) - len));
    hl[g] = (int32_t)len;
  });
  cudaStream_t s = w.stream;
  int* d_flags = (int*)w.d_flags.p;                          // [0] fallback count, [1] bad
  int* h_flags = (int*)w.h_flags.p;
  JG_TRY(cudaMemsetAsync(d_flags, 0, 16, s));
  JG_TRY(cudaMemcpyAsync(w.d_text.p, ht, (size_t)total, cudaMemcpyHostToDevice, s));
  JG_TRY(cudaMemcpyAsync(w.d_toff.p, w.h_toff.p, (n + 1) * 8, cudaMemcpyHostToDevice, s));
  JG_TRY(cudaMemcpyAsync(w.d_tlen.p, hl, n * 4, cudaMemcpyHostToDevice, s));
  parse_vectors<<<(unsigned)n, kThreads, parse_smem, s>>>((const char*)w.d_text.p, (const int64_t*)w.d_toff.p, (const int32_t*)w.d_tlen.p,
                                                          (int)dims, (double*)w.d_vals.p, d_flags, (int64_t*)w.d_fb.p, d_flags + 1);
  JG_TRY(cudaGetLastError());
  count(ETN_CTR_KERNEL_LAUNCHES);
  JG_TRY(cudaMemcpyAsync(h_flags, d_flags, 16, cudaMemcpyDeviceToHost, s));
  JG_TRY(cudaMemcpyAsync(w.h_fb.p, w.d_fb.p, kMaxFallback * 8, cudaMemcpyDeviceToHost, s));
  JG_TRY(cudaStreamSynchronize(s));
  if (h_flags[1] != 0 || h_flags[0] > kMaxFallback) return kJsonGpuUnavailable;
  if (h_flags[0] > 0) {
    // numbers the fast parser would not decide: the host's correctly rounded parser redoes the
    // vectors they sit in (etn_json_parse_array on the span), and the rows are patched
    std::vector<int64_t> rows;
    const int64_t* fb = (const int64_t*)w.h_fb.p;
    for (int i = 0; i < h_flags[0]; ++i) rows.push_back(fb[i] / dims);
    std::sort(rows.begin(), rows.end());
    rows.erase(std::unique(rows.begin(), rows.end()), rows.end());
    std::vector<double> row((size_t)dims);
    for (int64_t g : rows) {
      std::string arr = "[" + std::string(json + begin[g], (size_t)(end[g] - begin[g])) + "]";
      const int64_t got = etn_json_parse_array(arr.c_str(), arr.size(), row.data(), dims);
      if (got != dims) return got < 0 ? (int)got : ETN_ERR_JSON;
      JG_TRY(cudaMemcpyAsync((double*)w.d_vals.p + g * dims, row.data(), (size_t)dims * 8, cudaMemcpyHostToDevice, s));
      JG_TRY(cudaStreamSynchronize(s));                      // `row` is reused
    }
  }
  if ((rc = run_dense((int64_t)n, dims, dims, ETN_F64, ETN_TRUNCATE_RENORM, w.d_vals.p, w.d_norm.p, s))) return rc;
  format_vectors<<<(unsigned)n, kThreads, (size_t)ostride, s>>>((const double*)w.d_norm.p, (int)dims, (char*)w.d_out.p, ostride,
                                                                (int32_t*)w.d_olen.p, d_flags + 1);
  JG_TRY(cudaGetLastError());
  count(ETN_CTR_KERNEL_LAUNCHES);
  JG_TRY(cudaMemcpyAsync(w.h_out.p, w.d_out.p, n * (size_t)ostride, cudaMemcpyDeviceToHost, s));
  JG_TRY(cudaMemcpyAsync(w.h_olen.p, w.d_olen.p, n * 4, cudaMemcpyDeviceToHost, s));
  JG_TRY(cudaMemcpyAsync(h_flags, d_flags, 16, cudaMemcpyDeviceToHost, s));
  JG_TRY(cudaStreamSynchronize(s));
  if (h_flags[1] != 0) return ETN_ERR_JSON;                  // a NaN/Inf came out: Go's encoder would refuse the body
  count(ETN_CTR_HOST_CALLS); count(ETN_CTR_HOST_ROWS, (int64_t)n);
  count(ETN_CTR_H2D_BYTES, total); count(ETN_CTR_D2H_BYTES, (int64_t)(n * (size_t)ostride));
  observe(ETN_HIST_BATCH_ROWS, (double)n);
  res->text = (const char*)w.h_out.p;
  res->len = (const int32_t*)w.h_olen.p;
  res->stride = ostride;
  res->ws = &w;
  lock.release();                                            // stays locked until json_gpu_release
  return 0;
}

void json_gpu_release(JsonGpuResult* res) {
  if (res && res->ws) { static_cast<Workspace*>(res->ws)->mu.unlock(); res->ws = nullptr; }
}

void json_gpu_shutdown() {
  int prev = 0;
  const bool have_prev = cudaGetDevice(&prev) == cudaSuccess;
  for (Workspace& w : g_ws) {
    std::lock_guard<std::mutex> lk(w.mu);
    if (w.dev < 0) continue;
    if (cudaSetDevice(w.dev) != cudaSuccess) { cudaGetLastError(); continue; }
    if (w.stream) { cudaStreamSynchronize(w.stream); cudaStreamDestroy(w.stream); w.stream = nullptr; }
    for (Buf* b : {&w.d_text, &w.d_toff, &w.d_tlen, &w.d_vals, &w.d_norm, &w.d_out, &w.d_olen, &w.d_flags, &w.d_fb, &w.h_text, &w.h_toff,
                   &w.h_tlen, &w.h_out, &w.h_olen, &w.h_flags, &w.h_fb})
      b->release();
    w.dev = -1;
  }
  if (have_prev) cudaSetDevice(prev);
}

}  // namespace etn_rt

// ---- device twins of etn_json_parse_array / etn_json_format_array, for tests and the bench
namespace {

__global__ void parse_tokens(const char* __restrict__ text, const int64_t* __restrict__ beg, const int32_t* __restrict__ len, int64_t n,
                             double* __restrict__ out, uint8_t* __restrict__ decided) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    double v = 0.0;
    const bool ok = etn_num::parse_float64(text + beg[i], len[i], d_pow10, v);
    out[i] = v;
    decided[i] = ok ? 1 : 0;
  }
}

__global__ void format_numbers(const double* __restrict__ v, int64_t n, char* __restrict__ out, int32_t* __restrict__ len) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    len[i] = etn_num::format_go_float64(v[i], d_pow10, out + i * 32);
}

}  // namespace

extern "C" {

int etn_json_parse_tokens_device(const char* text, const int64_t* tok_begin, const int32_t* tok_len, int64_t n, double* out,
                                 uint8_t* decided, void* stream) {
  if (n < 0) return ETN_ERR_SHAPE;
  if (n == 0) return ETN_OK;
  if (!text || !tok_begin || !tok_len || !out || !decided) return ETN_ERR_NULL;
  parse_tokens<<<(unsigned)std::min<int64_t>((n + 255) / 256, 148 * 8), 256, 0, (cudaStream_t)stream>>>(text, tok_begin, tok_len, n, out, decided);
  const cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) etn_rt::count(ETN_CTR_KERNEL_LAUNCHES);
  return (int)e;
}

int etn_json_format_numbers_device(const double* v, int64_t n, char* out, int32_t* len, void* stream) {
  if (n < 0) return ETN_ERR_SHAPE;
  if (n == 0) return ETN_OK;
  if (!v || !out || !len) return ETN_ERR_NULL;
  format_numbers<<<(unsigned)std::min<int64_t>((n + 255) / 256, 148 * 8), 256, 0, (cudaStream_t)stream>>>(v, n, out, len);
  const cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) etn_rt::count(ETN_CTR_KERNEL_LAUNCHES);
  return (int)e;
}

}  // extern "C"
