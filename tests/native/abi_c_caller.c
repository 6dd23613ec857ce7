/* A plain C caller of libembed_truncate_norm.so — what the cgo package sees: nothing but
 * include/embed_truncate_norm.h, no CUDA headers, no Python, no torch.  Built with gcc and run
 * by tests/test_zy_c_caller.py on the GPU box.  Exits 0 when every check passes. */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/embed_truncate_norm.h"

#define CHECK(cond, ...) do { if (!(cond)) { fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); return 1; } } while (0)

int main(void) {
  CHECK(etn_version() == ETN_VERSION, "version");
  CHECK(strcmp(etn_strerror(ETN_ERR_NOT_INIT), "etn_init has not been called") == 0, "strerror");

  /* before init: the host path refuses, the allocation-free one works */
  double probe_in[4] = {3, 4, 0, 0}, probe_out[2] = {0, 0};
  CHECK(embed_truncate_norm_host(1, 4, 2, ETN_F64, ETN_TRUNCATE_RENORM, probe_in, 0, probe_out, 0) == ETN_ERR_NOT_INIT, "not init");
  int rc = embed_truncate_norm_host_static(1, 4, 2, ETN_F64, ETN_TRUNCATE_RENORM, probe_in, 0, probe_out, NULL);
  CHECK(rc == 0, "static: %s", etn_strerror(rc));
  CHECK(fabs(probe_out[0] - 0.6) < 1e-15 && fabs(probe_out[1] - 0.8) < 1e-15, "static values %g %g", probe_out[0], probe_out[1]);

  rc = etn_init(NULL, 0);
  CHECK(rc == 0, "etn_init: %s", etn_strerror(rc));
  CHECK(etn_device_count() >= 1, "device count");

  /* C1-shaped request through pinned buffers: 1024 x 4096 float64 -> 1024 */
  const int64_t batch = 1024, D = 4096, k = 1024;
  double *in = (double *)etn_alloc_pinned((size_t)(batch * D) * sizeof(double));
  double *out = (double *)etn_alloc_pinned((size_t)(batch * k) * sizeof(double));
  CHECK(in && out, "pinned allocation");
  for (int64_t i = 0; i < batch * D; ++i) in[i] = (double)((i * 2654435761u) % 2001) / 32000.0 - 0.03125;
  rc = embed_truncate_norm_host(batch, D, k, ETN_F64, ETN_TRUNCATE_RENORM, in, 0, out, 0);
  CHECK(rc == 0, "host path: %s", etn_strerror(rc));
  for (int64_t r = 0; r < batch; r += 97) {
    long double ss = 0, ss_in = 0;
    for (int64_t c = 0; c < k; ++c) { ss += (long double)out[r * k + c] * out[r * k + c]; ss_in += (long double)in[r * D + c] * in[r * D + c]; }
    CHECK(fabsl(ss - 1.0L) < 1e-14L, "row %ld norm^2 = %.18Lg", (long)r, ss);
    const double want = (double)((long double)in[r * D + 5] / sqrtl(ss_in));
    CHECK(fabs(out[r * k + 5] - want) <= 4 * fabs(nextafter(want, INFINITY) - want), "row %ld value", (long)r);
  }
  rc = embed_truncate_norm_host(batch, D, k, ETN_F64, ETN_TRUNCATE, in, 0, out, 0);
  CHECK(rc == 0, "truncate: %s", etn_strerror(rc));
  for (int64_t r = 0; r < batch; r += 131) CHECK(memcmp(out + r * k, in + r * D, (size_t)k * 8) == 0, "truncate row %ld", (long)r);

  /* two responses with different `dimensions` in flight at once (tickets), then one ragged call
   * that carries both: the batcher's view of handlers.go:1874-1877 */
  {
    const int64_t n1 = 300, n2 = 200, k1 = 256, k2 = 100;
    double *o1 = (double *)etn_alloc_pinned((size_t)(n1 * k1) * 8), *o2 = (double *)etn_alloc_pinned((size_t)(n2 * k2) * 8);
    double *packed = (double *)etn_alloc_pinned((size_t)(n1 * k1 + n2 * k2) * 8);
    int32_t *kpr = (int32_t *)malloc((size_t)(n1 + n2) * sizeof(int32_t));
    CHECK(o1 && o2 && packed && kpr, "allocation");
    int64_t t1 = 0, t2 = 0;
    rc = etn_submit_host(n1, D, k1, NULL, ETN_F64, ETN_TRUNCATE_RENORM, in, 0, o1, 0, &t1);
    CHECK(rc == 0 && t1 > 0, "submit 1: %s", etn_strerror(rc));
    rc = etn_submit_host(n2, D, k2, NULL, ETN_F64, ETN_TRUNCATE_RENORM, in + n1 * D, 0, o2, 0, &t2);
    CHECK(rc == 0 && t2 > t1, "submit 2: %s", etn_strerror(rc));
    CHECK(etn_wait(t2) == 0 && etn_wait(t1) == 0, "wait");
    CHECK(etn_wait(t1) == ETN_ERR_TICKET, "a ticket is waited for once");
    for (int64_t r = 0; r < n1 + n2; ++r) kpr[r] = r < n1 ? (int32_t)k1 : (int32_t)k2;
    rc = embed_truncate_norm_ragged_host(n1 + n2, D, kpr, ETN_F64, ETN_TRUNCATE_RENORM, in, 0, packed, 0);
    CHECK(rc == 0, "ragged host: %s", etn_strerror(rc));
    CHECK(memcmp(packed, o1, (size_t)(n1 * k1) * 8) == 0, "ragged rows of request 1 differ from its own call");
    CHECK(memcmp(packed + n1 * k1, o2, (size_t)(n2 * k2) * 8) == 0, "ragged rows of request 2 differ from its own call");
    int64_t cuts[3];
    CHECK(etn_shard_ragged(n1 + n2, D, kpr, 2, cuts) == 0 && cuts[0] == 0 && cuts[2] == n1 + n2 && cuts[1] > 0 && cuts[1] < n1, "shard cut %ld", (long)cuts[1]);
    kpr[5] = 0;
    CHECK(embed_truncate_norm_ragged_host(n1 + n2, D, kpr, ETN_F64, ETN_TRUNCATE_RENORM, in, 0, packed, 0) == ETN_ERR_SHAPE, "k <= 0 must be refused");
    etn_free_pinned(o1); etn_free_pinned(o2); etn_free_pinned(packed); free(kpr);
  }

  /* the JSON packer, the way proxyOpenAIEmbed would hand it a provider body */
  const char body[] = "{\"object\":\"list\",\"data\":[{\"index\":0,\"embedding\":[3,4,0,0]},{\"index\":1,\"embedding\":[1,2]}],\"usage\":{\"prompt_tokens\":2}}";
  char rewritten[256];
  int64_t stats[3];
  int64_t n = etn_json_truncate_response(body, sizeof body - 1, 2, ETN_TRUNCATE_RENORM, rewritten, sizeof rewritten, 0, stats);
  CHECK(n > 0, "packer: %ld", (long)n);
  rewritten[n] = 0;
  CHECK(strstr(rewritten, "\"embedding\":[0.6,0.8]") && strstr(rewritten, "\"embedding\":[1,2]"), "packer text: %s", rewritten);
  CHECK(stats[0] == 1 && stats[1] == 0 && stats[2] == 2, "packer stats");

  int64_t ctr[ETN_CTR_COUNT];
  CHECK(etn_get_counters(ctr, ETN_CTR_COUNT) == ETN_CTR_COUNT, "counters");
  /* host calls: the static probe, two C1 calls, the packer's one vector; the refused call before etn_init is a fallback, not an error */
  /* + two tickets (300 + 200 rows) and one ragged call (500 rows); its refused twin (k <= 0) is an argument error */
  CHECK(ctr[ETN_CTR_HOST_CALLS] == 7 && ctr[ETN_CTR_HOST_ROWS] == 2 * batch + 2 + 1000 && ctr[ETN_CTR_ERRORS] == 1 && ctr[ETN_CTR_FALLBACKS] == 1,
        "counter values %ld %ld %ld %ld", (long)ctr[ETN_CTR_HOST_CALLS], (long)ctr[ETN_CTR_HOST_ROWS], (long)ctr[ETN_CTR_ERRORS],
        (long)ctr[ETN_CTR_FALLBACKS]);
  double le[32], hsum = 0;
  int64_t hc[32];
  int nb = etn_get_histogram(ETN_HIST_BATCH_ROWS, le, hc, 32, &hsum);
  CHECK(nb > 2 && hsum == (double)(2 * batch + 2 + 1000), "batch-rows histogram: %d buckets, sum %g", nb, hsum);
  CHECK(strcmp(etn_counter_name(ETN_CTR_KERNEL_LAUNCHES), "llmcore_embed_gpu_kernel_launches_total") == 0, "counter name");

  etn_free_pinned(in);
  etn_free_pinned(out);
  etn_shutdown();
  CHECK(etn_device_count() == 0, "shutdown");
  printf("c caller ok\n");
  return 0;
}
