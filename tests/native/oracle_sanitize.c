/* ASan + UBSan run of the CPU oracle (sanitizers on the CPU side).  Includes the
 * oracle's source directly and drives every entry point on small and edge-case inputs. */
#include "../../oracle/etn_oracle.c"
#include <stdio.h>

int main(void) {
  enum { ROWS = 37, D = 100, K = 63 };
  float *xf = malloc(sizeof(float) * ROWS * D), *yf = malloc(sizeof(float) * ROWS * K);
  double *xd = malloc(sizeof(double) * ROWS * D), *yd = malloc(sizeof(double) * ROWS * K), *zd = malloc(sizeof(double) * ROWS * K);
  if (etn_oracle_fill(3, 5, ROWS, D, D, D, ETN_F32, xf) || etn_oracle_fill(3, 5, ROWS, D, D, D, ETN_F64, xd)) return 1;
  for (int i = 0; i < D; ++i) xd[2 * D + i] = 0.0;                 /* zero row */
  xd[3 * D] = 1e200; xd[3 * D + 1] = -1e-200; xd[4 * D] = 4.9e-324;  /* range edges */
  if (etn_oracle_truncate(ROWS, D, K, ETN_F32, xf, yf) || etn_oracle_truncate(ROWS, D, K, ETN_F64, xd, yd)) return 2;
  if (etn_oracle_renorm_exact(ROWS, D, K, ETN_F32, xf, yf) || etn_oracle_renorm_exact(ROWS, D, K, ETN_F64, xd, yd)) return 3;
  if (etn_oracle_renorm_go(ROWS, D, K, ETN_F64, xd, zd, 3) != 3 || etn_oracle_renorm_go(ROWS, D, K, ETN_F32, xf, yf, 1) != 1) return 4;
  int64_t arg = -1;
  const int64_t d = etn_oracle_max_ulp((int64_t)ROWS * K, ETN_F64, yd, zd, &arg);
  if (d < 0 || arg < 0) return 5;
  int32_t kpr[ROWS]; int64_t off[ROWS]; int64_t tot = 0;
  for (int r = 0; r < ROWS; ++r) { kpr[r] = 1 + (r * 7) % (D + 20); off[r] = tot; tot += kpr[r] < D ? kpr[r] : D; }
  double *pk = malloc(sizeof(double) * (size_t)tot);
  if (etn_oracle_ragged(ROWS, D, kpr, off, ETN_F64, 1, xd, pk) || etn_oracle_ragged(ROWS, D, kpr, off, ETN_F64, 0, xd, pk)) return 6;
  int64_t lens[4] = {4096, 10, 1024, 1025};
  if (etn_oracle_reslice(4, 1024, lens) != 2) return 7;
  if (etn_oracle_truncate(0, D, K, ETN_F64, NULL, NULL) != 0 || etn_oracle_renorm_exact(2, D, K, 9, xd, yd) != -3) return 8;
  printf("oracle sanitize ok, go-vs-exact max ulp %ld\n", (long)d);
  free(xf); free(yf); free(xd); free(yd); free(zd); free(pk);
  return 0;
}
