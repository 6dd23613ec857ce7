// etn_kernels.cuh — device code for libembed_truncate_norm.so (sm_100a).
//
// The op replaces the per-vector loop of truncateEmbeddings
// (/root/reference/core/internal/api/handlers.go:2068-2076) on a packed row-major matrix:
// keep the first k of D components of every row and, in RENORM mode, divide them by their
// L2 norm.  It is a bandwidth-bound streaming reduction (3 flops per 2*sizeof(T) bytes), so
// there are no tensor cores here: the kernels are built around 16-byte coalesced loads of
// the row prefix only, whole rows held in registers between the load and the store, and a
// warp-shuffle reduction.
//
// Numerics (bar: <= 4 ulp of the exactly rounded quotient, see tests/):
//   float32  squares are exact in float64 and are summed there; the scale 1/sqrt(ss) is
//            rounded to float32 once and applied with one multiply  -> <= 1.5 ulp.
//   float64  the sum of squares is carried as an error-free (sum, compensation) pair
//            (TwoProduct by FMA + TwoSum), the reciprocal norm as a double-double
//            (r, r_lo) from one Newton correction, and y = fma(x, r, x*r_lo) -> ~0.5 ulp.
//   Rows whose squares would overflow/underflow are rescaled by a power of two first;
//   all-zero rows are copied.  Both are warp-uniform slow paths.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <float.h>

namespace etn {

constexpr int kModeTruncate = 0;
constexpr int kModeRenorm = 1;
constexpr int kWarpsPerBlock = 8;
constexpr int kThreads = kWarpsPerBlock * 32;
constexpr unsigned kFull = 0xffffffffu;

// ------------------------------------------------------------ memory access ----
// Input is read exactly once and output written exactly once: keep both out of L1.
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};"
               :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// -------------------------------------------------------------- reductions ----
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFull, v, o);
  return v;
}
__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(kFull, v, o));
  return v;
}

// Error-free accumulator of squares: value = s + c exactly up to O(eps^2).
struct Acc2 {
  double s, c;
  __device__ __forceinline__ void clear() { s = 0.0; c = 0.0; }
  __device__ __forceinline__ void add_square(double x) {
    double p = x * x;
    double e = fma(x, x, -p);            // TwoProduct: x*x = p + e exactly
    double t = s + p;                    // TwoSum (Knuth): s + p = t + err exactly
    double bp = t - s;
    double err = (s - (t - bp)) + (p - bp);
    s = t;
    c += err + e;
  }
  __device__ __forceinline__ void merge(double os, double oc) {
    double t = s + os;
    double bp = t - s;
    double err = (s - (t - bp)) + (os - bp);
    s = t;
    c += oc + err;
  }
  __device__ __forceinline__ void warp_reduce() {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      double os = __shfl_xor_sync(kFull, s, o);
      double oc = __shfl_xor_sync(kFull, c, o);
      merge(os, oc);
    }
  }
};

// ------------------------------------------------------------ row scaling ----
enum : int { kCopy = 0, kFast = 1, kSlow = 2 };

// float32 rows.  ss = exact-ish float64 sum of squares (never overflows for float input).
struct Scale32 {
  int kind;
  float inv_f;
  double inv_d;
  __device__ __forceinline__ void from_ss(double ss) {
    inv_f = 0.f; inv_d = 0.0;
    if (!(ss > 0.0)) { kind = kCopy; return; }          // all zeros (or NaN in: pass through)
    inv_d = rsqrt(ss);
    inv_f = (float)inv_d;
    // a subnormal or infinite float scale would lose bits: do the product in float64 then
    kind = (inv_f >= FLT_MIN && inv_f <= FLT_MAX) ? kFast : kSlow;
  }
  __device__ __forceinline__ float apply(float x) const {
    return kind == kFast ? x * inv_f : (float)((double)x * inv_d);
  }
};

// float64 rows.  (r, r_lo) is 1/sqrt(ss) as a double-double; pre is a power-of-two
// exponent applied to x first (slow path only).
struct Scale64 {
  int kind;
  int pre;           // x is multiplied by 2^pre before use when kind == kSlow
  double r, r_lo;
  // ss_hi + ss_lo = sum of squares (of the possibly pre-scaled row), ss_hi normal and finite.
  __device__ __forceinline__ void from_ss(double ss_hi, double ss_lo) {
    r = rsqrt(ss_hi);                              // <= 1 ulp
    double t = ss_hi * r;
    double te = fma(ss_hi, r, -t);                 // ss_hi*r = t + te
    double h = fma(t, r, -1.0);                    // ss*r^2 - 1, leading part
    h += fma(te, r, ss_lo * r * r);
    r_lo = -0.5 * h * r;                           // Newton: 1/sqrt(ss) = r*(1 - h/2)
  }
  __device__ __forceinline__ double apply(double x) const {
    if (kind == kSlow) x = scalbn(x, pre);
    return fma(x, r, x * r_lo);
  }
};

// Range in which squares neither overflow nor lose bits to underflow in a way that matters:
// 2^-900 <= ss < inf.  Elements below 2^-511 then contribute < 2^-120 relative.
__device__ __forceinline__ bool ss_in_range(double ss) {
  return ss >= 0x1p-900 && ss <= DBL_MAX;
}

// ------------------------------------------------------- register-row kernel ----
// One warp owns R rows at a time; each lane holds NV 16-byte vectors of every row, i.e. the
// whole k-prefix of R rows sits in registers between the load and the store.  All R*NV
// loads of an iteration are issued before anything is consumed (memory-level parallelism).
// Requires 16-byte aligned rows on both sides; nvec = 16-byte vectors per row prefix.
template <typename T> struct VecTraits;
template <> struct VecTraits<float>  { static constexpr int kPer = 4; };
template <> struct VecTraits<double> { static constexpr int kPer = 2; };

template <typename T, int NV, int R, int MODE, int MINB>
__global__ void __launch_bounds__(kThreads, MINB)
rows_reg_kernel(const char* __restrict__ in, char* __restrict__ out, int64_t batch,
                int64_t in_pitch /*bytes*/, int nvec) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * kThreads) >> 5;
  const int64_t out_pitch = (int64_t)nvec * 16;

  for (int64_t row0 = warp * R; row0 < batch; row0 += nwarps * R) {
    uint4 v[R][NV];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const bool row_ok = row0 + r < batch;
      const char* src = in + (row0 + r) * in_pitch;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int idx = j * 32 + lane;
        if (row_ok && idx < nvec) v[r][j] = ld_stream16(src + (int64_t)idx * 16);
        else v[r][j] = make_uint4(0u, 0u, 0u, 0u);
      }
    }

    if constexpr (MODE == kModeRenorm) {
      if constexpr (sizeof(T) == 4) {
        double ss[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          double a0 = 0.0, a1 = 0.0;
#pragma unroll
          for (int j = 0; j < NV; ++j) {
            const double x0 = (double)__uint_as_float(v[r][j].x);
            const double x1 = (double)__uint_as_float(v[r][j].y);
            const double x2 = (double)__uint_as_float(v[r][j].z);
            const double x3 = (double)__uint_as_float(v[r][j].w);
            a0 = fma(x0, x0, a0); a1 = fma(x1, x1, a1);
            a0 = fma(x2, x2, a0); a1 = fma(x3, x3, a1);
          }
          ss[r] = a0 + a1;
        }
        // R independent butterflies, interleaved
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int r = 0; r < R; ++r) ss[r] += __shfl_xor_sync(kFull, ss[r], o);
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
          Scale32 sc; sc.from_ss(ss[r]);
          if (sc.kind == kFast) {                    // warp-uniform
            const float f = sc.inv_f;
#pragma unroll
            for (int j = 0; j < NV; ++j) {
              v[r][j].x = __float_as_uint(__uint_as_float(v[r][j].x) * f);
              v[r][j].y = __float_as_uint(__uint_as_float(v[r][j].y) * f);
              v[r][j].z = __float_as_uint(__uint_as_float(v[r][j].z) * f);
              v[r][j].w = __float_as_uint(__uint_as_float(v[r][j].w) * f);
            }
          } else if (sc.kind == kSlow) {
            const double d = sc.inv_d;
#pragma unroll
            for (int j = 0; j < NV; ++j) {
              v[r][j].x = __float_as_uint((float)((double)__uint_as_float(v[r][j].x) * d));
              v[r][j].y = __float_as_uint((float)((double)__uint_as_float(v[r][j].y) * d));
              v[r][j].z = __float_as_uint((float)((double)__uint_as_float(v[r][j].z) * d));
              v[r][j].w = __float_as_uint((float)((double)__uint_as_float(v[r][j].w) * d));
            }
          }
        }
      } else {
#pragma unroll
        for (int r = 0; r < R; ++r) {
          Acc2 acc; acc.clear();
#pragma unroll
          for (int j = 0; j < NV; ++j) {
            acc.add_square(__hiloint2double((int)v[r][j].y, (int)v[r][j].x));
            acc.add_square(__hiloint2double((int)v[r][j].w, (int)v[r][j].z));
          }
          acc.warp_reduce();
          Scale64 sc;
          sc.pre = 0;
          const double ss_hi = acc.s + acc.c;
          if (ss_in_range(ss_hi)) {
            sc.kind = kFast;
            sc.from_ss(ss_hi, (acc.s - ss_hi) + acc.c);
          } else {
            // zero row, or squares out of range: rescale by the row's largest magnitude
            double m = 0.0;
#pragma unroll
            for (int j = 0; j < NV; ++j) {
              m = fmax(m, fabs(__hiloint2double((int)v[r][j].y, (int)v[r][j].x)));
              m = fmax(m, fabs(__hiloint2double((int)v[r][j].w, (int)v[r][j].z)));
            }
            m = warp_max(m);
            if (!(m > 0.0) || !(m <= DBL_MAX)) {
              sc.kind = kCopy;                       // all zeros (or non-finite in: pass through)
            } else {
              sc.kind = kSlow;
              sc.pre = -ilogb(m);
              Acc2 a2; a2.clear();
#pragma unroll
              for (int j = 0; j < NV; ++j) {
                a2.add_square(scalbn(__hiloint2double((int)v[r][j].y, (int)v[r][j].x), sc.pre));
                a2.add_square(scalbn(__hiloint2double((int)v[r][j].w, (int)v[r][j].z), sc.pre));
              }
              a2.warp_reduce();
              const double h2 = a2.s + a2.c;
              sc.from_ss(h2, (a2.s - h2) + a2.c);
            }
          }
          if (sc.kind != kCopy) {
#pragma unroll
            for (int j = 0; j < NV; ++j) {
              const double y0 = sc.apply(__hiloint2double((int)v[r][j].y, (int)v[r][j].x));
              const double y1 = sc.apply(__hiloint2double((int)v[r][j].w, (int)v[r][j].z));
              v[r][j].x = (unsigned)__double2loint(y0); v[r][j].y = (unsigned)__double2hiint(y0);
              v[r][j].z = (unsigned)__double2loint(y1); v[r][j].w = (unsigned)__double2hiint(y1);
            }
          }
        }
      }
    }

#pragma unroll
    for (int r = 0; r < R; ++r) {
      const bool row_ok = row0 + r < batch;
      char* dst = out + (row0 + r) * out_pitch;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int idx = j * 32 + lane;
        if (row_ok && idx < nvec) st_stream16(dst + (int64_t)idx * 16, v[r][j]);
      }
    }
  }
}

// ------------------------------------------------------------ general kernel ----
// Any k, any D, any element-aligned pointers; also the ragged path.  One warp per row, two
// passes over the row prefix (the second pass re-reads from L1/L2), element-wise coalesced
// accesses.  k_per_row == nullptr -> every row keeps k; out_offsets == nullptr -> dense.
template <typename T, int MODE>
__global__ void __launch_bounds__(kThreads)
rows_general_kernel(const T* __restrict__ in, T* __restrict__ out, int64_t batch, int64_t D,
                    int64_t k, const int32_t* __restrict__ k_per_row,
                    const int64_t* __restrict__ out_offsets) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * kThreads + threadIdx.x) >> 5;
  const int64_t nwarps = ((int64_t)gridDim.x * kThreads) >> 5;

  for (int64_t row = warp; row < batch; row += nwarps) {
    int64_t kr = k_per_row ? (int64_t)k_per_row[row] : k;
    if (kr > D) kr = D;
    const T* x = in + row * D;
    T* y = out + (out_offsets ? out_offsets[row] : row * kr);

    if constexpr (MODE == kModeTruncate) {
      for (int64_t i = lane; i < kr; i += 32) y[i] = x[i];
    } else if constexpr (sizeof(T) == 4) {
      double a = 0.0;
      for (int64_t i = lane; i < kr; i += 32) { const double xv = (double)x[i]; a = fma(xv, xv, a); }
      Scale32 sc; sc.from_ss(warp_sum(a));
      if (sc.kind == kCopy) { for (int64_t i = lane; i < kr; i += 32) y[i] = x[i]; }
      else { for (int64_t i = lane; i < kr; i += 32) y[i] = sc.apply(x[i]); }
    } else {
      Acc2 acc; acc.clear();
      for (int64_t i = lane; i < kr; i += 32) acc.add_square(x[i]);
      acc.warp_reduce();
      Scale64 sc; sc.pre = 0;
      const double ss_hi = acc.s + acc.c;
      if (ss_in_range(ss_hi)) {
        sc.kind = kFast;
        sc.from_ss(ss_hi, (acc.s - ss_hi) + acc.c);
      } else {
        double m = 0.0;
        for (int64_t i = lane; i < kr; i += 32) m = fmax(m, fabs(x[i]));
        m = warp_max(m);
        if (!(m > 0.0) || !(m <= DBL_MAX)) {
          sc.kind = kCopy;
        } else {
          sc.kind = kSlow;
          sc.pre = -ilogb(m);
          Acc2 a2; a2.clear();
          for (int64_t i = lane; i < kr; i += 32) a2.add_square(scalbn(x[i], sc.pre));
          a2.warp_reduce();
          const double h2 = a2.s + a2.c;
          sc.from_ss(h2, (a2.s - h2) + a2.c);
        }
      }
      if (sc.kind == kCopy) { for (int64_t i = lane; i < kr; i += 32) y[i] = x[i]; }
      else { for (int64_t i = lane; i < kr; i += 32) y[i] = sc.apply(x[i]); }
    }
  }
}

// ---------------------------------------------------------- synthetic fill ----
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

template <typename T>
__global__ void __launch_bounds__(256)
fill_kernel(T* __restrict__ out, uint64_t seed, int64_t row0, int64_t rows, int64_t D,
            int64_t ncols, int64_t ld) {
  const int64_t total = rows * ncols;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / ncols, c = i - r * ncols;
    const uint64_t h = splitmix64(seed ^ ((uint64_t)(row0 + r) * (uint64_t)D + (uint64_t)c));
    const int32_t q = (int32_t)(uint32_t)(h >> 32) >> 8;   // signed 24-bit
    out[r * ld + c] = (T)((double)q * 0x1p-28);
  }
}

}  // namespace etn
