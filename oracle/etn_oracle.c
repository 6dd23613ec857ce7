/*
 * etn_oracle.c — CPU oracle for the Matryoshka truncate(+renormalise) path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (llm-mcp_b200/) may call,
 * link or import this file.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or as the
 * timed CPU baseline — never as the thing shipped.
 *
 * PARITY UNPINNED: the reference (plagness/LLM-MCP @ VERSION 2026.03.11) has no test,
 * golden vector or fixture for this path, and its Go sources cannot be
 * compiled here (no Go toolchain).  This file is therefore a restatement,
 * pinned only against golden vectors authored in this repo with an independent
 * arbitrary-precision computation (tests/golden/make_golden.py, Python decimal).
 *
 * What it restates:
 *   O-trunc  core/internal/api/handlers.go:2062-2078 (truncateEmbeddings): keep the first
 *            `dims` components of each embedding, original order, values bit-identical,
 *            rows with len <= dims untouched.  On the dense [batch x D] view that is a
 *            strided copy of the k = min(D, dims) prefix.
 *   O-exact  L2 renormalisation of that prefix.  The reference does NOT renormalise
 *            (SURVEY.md F1); this is the behaviour BASELINE.json's north_star defines.
 *            Sum of squares, sqrt and the quotient are evaluated in IEEE binary128 and the
 *            quotient is rounded once to the element type: the oracle of record for the
 *            "<= 4 ulp" bar.  A row whose prefix is all zeros is copied unchanged.
 *   O-go     the loop a Go author would add after handlers.go:2075 — sequential float64
 *            sum of squares, math.Sqrt, per-element divide — compiled scalar and without
 *            FMA contraction to model the Go compiler at GOAMD64=v1 (core/Dockerfile:6).
 *            Used as the timed CPU baseline and to report a naive loop's distance
 *            from exact.
 *   fill     the counter-based synthetic generator of SURVEY.md §8(d).
 *
 * Element type: the reference's is float64 (handlers.go:2047-2050 decode into
 * map[string]any → boxed float64); float32 is this project's addition.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -fno-tree-vectorize -fopenmp).
 */
#include <math.h>
#include <quadmath.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ETN_F32 0
#define ETN_F64 1

/* ------------------------------------------------------------------ fill ---- */

static inline uint64_t splitmix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ULL;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}

/* x(row, col) = ((int32)(splitmix64(seed ^ (row*D + col)) >> 32) >> 8) * 2^-28:
 * a signed 24-bit integer scaled into [-1/32, 1/32), so it is exactly representable in
 * float32 and float64.  (SURVEY.md §8(d) proposed the full int32 * 2^-36, which has 31
 * significant bits and is NOT exact in float32; the low 8 bits are dropped here.) */
static inline double synth_value(uint64_t seed, uint64_t row, uint64_t D, uint64_t col) {
  uint64_t h = splitmix64(seed ^ (row * D + col));
  int32_t v = (int32_t)(uint32_t)(h >> 32) >> 8;   /* arithmetic shift */
  return (double)v * 0x1p-28;
}

/* Fills rows [row0, row0+rows) x cols [0, ncols) of the logical [.. x D] matrix into
 * `out`, laid out row-major with leading dimension ld elements. */
int etn_oracle_fill(uint64_t seed, int64_t row0, int64_t rows, int64_t D, int64_t ncols,
                    int64_t ld, int dtype, void *out) {
  if (!out || rows < 0 || D <= 0 || ncols < 0 || ncols > D || ld < ncols) return -2;
  if (dtype != ETN_F32 && dtype != ETN_F64) return -3;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < rows; ++r) {
    for (int64_t c = 0; c < ncols; ++c) {
      double v = synth_value(seed, (uint64_t)(row0 + r), (uint64_t)D, (uint64_t)c);
      if (dtype == ETN_F32) ((float *)out)[r * ld + c] = (float)v;
      else ((double *)out)[r * ld + c] = v;
    }
  }
  return 0;
}

/* --------------------------------------------------------------- O-trunc ---- */

/* handlers.go:2073-2076: m["embedding"] = emb[:dims] when len(emb) > dims. */
int etn_oracle_truncate(int64_t batch, int64_t D, int64_t k, int dtype, const void *in,
                        void *out) {
  if (batch < 0 || D <= 0 || k <= 0) return -2;
  if (dtype != ETN_F32 && dtype != ETN_F64) return -3;
  if (batch == 0) return 0;
  if (!in || !out) return -1;
  size_t es = dtype == ETN_F32 ? 4 : 8;
  int64_t kk = k < D ? k : D;
  for (int64_t r = 0; r < batch; ++r)
    memcpy((char *)out + (size_t)r * kk * es, (const char *)in + (size_t)r * D * es,
           (size_t)kk * es);
  return 0;
}

/* --------------------------------------------------------------- O-exact ---- */

static void exact_row_f64(const double *x, int64_t k, double *y) {
  __float128 ss = 0;
  for (int64_t i = 0; i < k; ++i) ss += (__float128)x[i] * (__float128)x[i];
  if (ss == 0) { memcpy(y, x, (size_t)k * sizeof(double)); return; }
  __float128 n = sqrtq(ss);
  for (int64_t i = 0; i < k; ++i) y[i] = (double)((__float128)x[i] / n);
}

static void exact_row_f32(const float *x, int64_t k, float *y) {
  __float128 ss = 0;
  for (int64_t i = 0; i < k; ++i) ss += (__float128)x[i] * (__float128)x[i];
  if (ss == 0) { memcpy(y, x, (size_t)k * sizeof(float)); return; }
  __float128 n = sqrtq(ss);
  for (int64_t i = 0; i < k; ++i) y[i] = (float)((__float128)x[i] / n);
}

int etn_oracle_renorm_exact(int64_t batch, int64_t D, int64_t k, int dtype, const void *in,
                            void *out) {
  if (batch < 0 || D <= 0 || k <= 0) return -2;
  if (dtype != ETN_F32 && dtype != ETN_F64) return -3;
  if (batch == 0) return 0;
  if (!in || !out) return -1;
  int64_t kk = k < D ? k : D;
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < batch; ++r) {
    if (dtype == ETN_F32) exact_row_f32((const float *)in + r * D, kk, (float *)out + r * kk);
    else exact_row_f64((const double *)in + r * D, kk, (double *)out + r * kk);
  }
  return 0;
}

/* Ragged: row r keeps k_per_row[r] (clamped to D) components, written at element offset
 * out_offsets[r] of the packed output.  mode 0 = truncate, 1 = exact renorm. */
int etn_oracle_ragged(int64_t batch, int64_t D, const int32_t *k_per_row,
                      const int64_t *out_offsets, int dtype, int mode, const void *in,
                      void *out) {
  if (batch < 0 || D <= 0) return -2;
  if ((dtype != ETN_F32 && dtype != ETN_F64) || (mode != 0 && mode != 1)) return -3;
  if (batch == 0) return 0;
  if (!in || !out || !k_per_row || !out_offsets) return -1;
  for (int64_t r = 0; r < batch; ++r)
    if (k_per_row[r] <= 0) return -2;
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t r = 0; r < batch; ++r) {
    int64_t kk = k_per_row[r] < D ? k_per_row[r] : D;
    if (dtype == ETN_F32) {
      const float *x = (const float *)in + r * D;
      float *y = (float *)out + out_offsets[r];
      if (mode == 0) memcpy(y, x, (size_t)kk * 4); else exact_row_f32(x, kk, y);
    } else {
      const double *x = (const double *)in + r * D;
      double *y = (double *)out + out_offsets[r];
      if (mode == 0) memcpy(y, x, (size_t)kk * 8); else exact_row_f64(x, kk, y);
    }
  }
  return 0;
}

/* ------------------------------------------------------------------ O-go ---- */

/* The Go loop being modelled (would sit after handlers.go:2075):
 *   ss := 0.0
 *   for _, v := range e[:k] { ss += v * v }
 *   n := math.Sqrt(ss)
 *   if n > 0 { for i := range out { out[i] = e[i] / n } }
 * float32 flavour: the same with float64(v) widening and a float32() narrowing store. */
static void go_row_f64(const double *x, int64_t k, double *y) {
  double ss = 0.0;
  for (int64_t i = 0; i < k; ++i) ss += x[i] * x[i];
  double n = sqrt(ss);
  if (n > 0) { for (int64_t i = 0; i < k; ++i) y[i] = x[i] / n; }
  else memcpy(y, x, (size_t)k * sizeof(double));
}

static void go_row_f32(const float *x, int64_t k, float *y) {
  double ss = 0.0;
  for (int64_t i = 0; i < k; ++i) { double v = (double)x[i]; ss += v * v; }
  double n = sqrt(ss);
  if (n > 0) { for (int64_t i = 0; i < k; ++i) y[i] = (float)((double)x[i] / n); }
  else memcpy(y, x, (size_t)k * sizeof(float));
}

/* nthreads <= 1: one thread (a single handler goroutine); otherwise rows are split over
 * nthreads OS threads, standing in for GOMAXPROCS = nthreads. Returns threads used. */
int etn_oracle_renorm_go(int64_t batch, int64_t D, int64_t k, int dtype, const void *in,
                         void *out, int nthreads) {
  if (batch < 0 || D <= 0 || k <= 0) return -2;
  if (dtype != ETN_F32 && dtype != ETN_F64) return -3;
  if (batch == 0) return 0;
  if (!in || !out) return -1;
  int64_t kk = k < D ? k : D;
  if (nthreads < 1) nthreads = 1;
  if ((int64_t)nthreads > batch) nthreads = (int)batch;
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (int64_t r = 0; r < batch; ++r) {
    if (dtype == ETN_F32) go_row_f32((const float *)in + r * D, kk, (float *)out + r * kk);
    else go_row_f64((const double *)in + r * D, kk, (double *)out + r * kk);
  }
  return nthreads;
}

/* The reference's actual work (handlers.go:2068-2076) on a dense stand-in for its
 * []any-of-[]any operand: one slice-header update per vector, no element is touched.
 * `lens` plays len(emb); returns how many headers changed.  O(batch). */
int64_t etn_oracle_reslice(int64_t batch, int64_t dims, int64_t *lens) {
  int64_t changed = 0;
  for (int64_t r = 0; r < batch; ++r)
    if (lens[r] > dims) { lens[r] = dims; ++changed; }
  return changed;
}

/* ------------------------------------------------------------- ulp tools ---- */

static inline int64_t ord_f64(double v) {
  int64_t b; memcpy(&b, &v, 8);
  return b < 0 ? (int64_t)0x8000000000000000ULL - b : b;   /* monotone, -0 == +0 */
}
static inline int64_t ord_f32(float v) {
  int32_t b; memcpy(&b, &v, 4);
  return b < 0 ? (int64_t)(int32_t)0x80000000 - (int64_t)b : (int64_t)b;
}

/* Largest distance in units-in-the-last-place between a[i] and b[i]; -1 if a NaN is
 * present on either side.  *argmax receives the index. */
int64_t etn_oracle_max_ulp(int64_t n, int dtype, const void *a, const void *b,
                           int64_t *argmax) {
  int64_t worst = 0, wi = 0;
  for (int64_t i = 0; i < n; ++i) {
    int64_t d;
    if (dtype == ETN_F32) {
      float x = ((const float *)a)[i], y = ((const float *)b)[i];
      if (x != x || y != y) { if (argmax) *argmax = i; return -1; }
      d = ord_f32(x) - ord_f32(y);
    } else {
      double x = ((const double *)a)[i], y = ((const double *)b)[i];
      if (x != x || y != y) { if (argmax) *argmax = i; return -1; }
      /* ord difference can exceed int64 only for opposite-sign huge values */
      __int128 dd = (__int128)ord_f64(x) - (__int128)ord_f64(y);
      if (dd < 0) dd = -dd;
      d = dd > (__int128)INT64_MAX ? INT64_MAX : (int64_t)dd;
    }
    if (d < 0) d = -d;
    if (d > worst) { worst = d; wi = i; }
  }
  if (argmax) *argmax = wi;
  return worst;
}

int etn_oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
