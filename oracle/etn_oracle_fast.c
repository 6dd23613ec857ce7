/*
 * etn_oracle_fast.c — "best CPU" flavour of the renormalise loop (baseline 3):
 * the same mathematics as O-go, written so that gcc vectorises it (eight independent partial
 * sums, multiply by the reciprocal norm), built -O3 with AVX2+FMA and AVX-512 clones chosen
 * at load time (target_clones), rows split over OpenMP threads.
 *
 * TEST/BENCH INFRASTRUCTURE ONLY, like the rest of oracle/: a timed CPU baseline, never a
 * product path.  It is NOT the oracle of record (results may differ from O-exact by 1-2 ulp).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define CLONES __attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))

CLONES static void fast_row_f32(const float *restrict x, int64_t k, float *restrict y) {
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int64_t i = 0;
  for (; i + 8 <= k; i += 8)
    for (int j = 0; j < 8; ++j) { double v = (double)x[i + j]; acc[j] += v * v; }
  double ss = 0;
  for (int j = 0; j < 8; ++j) ss += acc[j];
  for (; i < k; ++i) ss += (double)x[i] * (double)x[i];
  if (!(ss > 0)) { memcpy(y, x, (size_t)k * sizeof(float)); return; }
  const float inv = (float)(1.0 / sqrt(ss));
  for (i = 0; i < k; ++i) y[i] = x[i] * inv;
}

CLONES static void fast_row_f64(const double *restrict x, int64_t k, double *restrict y) {
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int64_t i = 0;
  for (; i + 8 <= k; i += 8)
    for (int j = 0; j < 8; ++j) acc[j] += x[i + j] * x[i + j];
  double ss = 0;
  for (int j = 0; j < 8; ++j) ss += acc[j];
  for (; i < k; ++i) ss += x[i] * x[i];
  if (!(ss > 0)) { memcpy(y, x, (size_t)k * sizeof(double)); return; }
  const double inv = 1.0 / sqrt(ss);
  for (i = 0; i < k; ++i) y[i] = x[i] * inv;
}

int etn_oracle_renorm_fast(int64_t batch, int64_t D, int64_t k, int dtype, const void *in, void *out,
                           int nthreads) {
  if (batch < 0 || D <= 0 || k <= 0) return -2;
  if (dtype != 0 && dtype != 1) return -3;
  if (batch == 0) return 0;
  if (!in || !out) return -1;
  const int64_t kk = k < D ? k : D;
  if (nthreads < 1) nthreads = 1;
  if ((int64_t)nthreads > batch) nthreads = (int)batch;
#pragma omp parallel for schedule(static) num_threads(nthreads)
  for (int64_t r = 0; r < batch; ++r) {
    if (dtype == 0) fast_row_f32((const float *)in + r * D, kk, (float *)out + r * kk);
    else fast_row_f64((const double *)in + r * D, kk, (double *)out + r * kk);
  }
  return nthreads;
}
