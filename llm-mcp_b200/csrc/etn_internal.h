// etn_internal.h — what the translation units of libembed_truncate_norm.so share with each
// other (not part of the C-ABI): kernel dispatch (etn_api.cu), metrics (etn_api.cu) and the
// host runtime (etn_host.cu).
#pragma once
#ifdef ETN_NO_CUDA                        // the sanitizer harness builds etn_json.cpp with plain g++
typedef struct CUstream_st* cudaStream_t;
#else
#include <cuda_runtime.h>
#endif
#include <stddef.h>
#include <stdint.h>

namespace etn_rt {

// ---- metrics (indices are ETN_CTR_* / ETN_HIST_* of the public header)
void count(int ctr, int64_t v = 1);
int finish(int rc);                       // counts a non-zero return, passes it through
void observe(int hist, double value);     // one sample into a fixed-bucket histogram

// ---- kernel dispatch on the calling thread's current device
// Dense rows: out[r, :kk] = op(in[r, :kk]), kk = min(k, D).  out == in is legal for k >= D.
int run_dense(int64_t batch, int64_t D, int64_t k, int dtype, int mode, const void* in, void* out,
              cudaStream_t s);
// Rows that share nothing but the element type: row r keeps min(k_per_row[r], D) elements and
// is written at element out_offsets[r] of `out`; it starts at element in_offsets[r] of `in`
// when in_offsets != nullptr, else at r * D.
int run_rows(int64_t batch, int64_t D, const int32_t* k_per_row, const int64_t* out_offsets,
             const int64_t* in_offsets, int dtype, int mode, const void* in, void* out, cudaStream_t s);

// ---- etn_json.cpp
void release_json_staging();              // frees the packer's reusable pinned buffer

// ---- etn_host.cu
int least_loaded_device();                // CUDA ordinal of the bound device with the least queued bytes, -1 before etn_init

// ---- etn_json_gpu.cu: the packer's number work on the device
constexpr int kJsonGpuUnavailable = 1 << 30;   // "use the host path" (not an error)
struct JsonGpuResult {
  const char* text = nullptr;             // pinned: vector g's "[a,b,c]" starts at text + g * stride
  const int32_t* len = nullptr;           // pinned: its length in bytes
  int64_t stride = 0;
  void* ws = nullptr;                     // the workspace that stays locked until json_gpu_release
};
// vector g's kept numbers are the text json[begin[g], end[g]) ("a, b, c" without brackets), `dims` of them
int json_renorm_gpu(const char* json, const int64_t* begin, const int64_t* end, size_t n, int64_t dims, JsonGpuResult* res);
void json_gpu_release(JsonGpuResult* res);
void json_gpu_shutdown();

inline bool aligned_to(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

}  // namespace etn_rt
