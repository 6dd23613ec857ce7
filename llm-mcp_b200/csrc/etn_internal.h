// etn_internal.h — what the translation units of libembed_truncate_norm.so share with each
// other (not part of the C-ABI): kernel dispatch (etn_api.cu), metrics (etn_api.cu) and the
// host runtime (etn_host.cu).
#pragma once
#include <cuda_runtime.h>
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

inline bool aligned_to(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

}  // namespace etn_rt
