/*
 * embed_truncate_norm.h — C-ABI of libembed_truncate_norm.so (CUDA, sm_100a / B200).
 *
 * One op: Matryoshka truncation of embedding vectors, optionally followed by L2
 * renormalisation, batched and device-resident.  It replaces the per-vector loop of
 *     truncateEmbeddings(result map[string]any, dims int)
 *         /root/reference/core/internal/api/handlers.go:2062-2078
 * called from Server.HandleEmbeddings at handlers.go:1891-1893 when dims > 0.
 *
 * The reference has no FFI boundary (CGO_ENABLED=0, one private function,
 * one call site), so every entry point below is the function a new cgo package
 * `core/internal/embedgpu` would bind; INTEGRATION.md shows that binding.  The reference
 * interface each one stands in for is cited per function.
 *
 * Conventions
 *   - Matrices are row-major.  `in` is [batch x D]; dense `out` is [batch x kk] with
 *     kk = min(k, D) (handlers.go:2074: vectors with len <= dims are left whole, never
 *     padded).  Components keep their order; TRUNCATE keeps their bits.
 *   - dtype: ETN_F32 = IEEE binary32, ETN_F64 = IEEE binary64 (the reference's element
 *     type: encoding/json decodes into float64, handlers.go:2047-2050).  Little endian.
 *   - mode:  ETN_TRUNCATE        out = in[:, :kk]                       (the reference)
 *            ETN_TRUNCATE_RENORM out = in[:, :kk] / ||in[:, :kk]||_2    (BASELINE.json;
 *            not in the reference).  A row whose prefix is all zeros is
 *            copied unchanged; finite input always gives finite output (Go's encoder
 *            drops the whole body on NaN/Inf, helpers.go:13-14).  Results are within
 *            2 ulp of the exactly rounded quotient for every finite input, subnormals and
 *            values whose squares overflow included.
 *   - Return value: 0 = ok; negative = ETN_ERR_* argument errors; positive = the
 *     cudaError_t reported by the CUDA runtime.  Nothing aborts, throws or prints.
 *   - Device entry points work on the calling thread's current CUDA device and enqueue
 *     on `stream` (a cudaStream_t; NULL = the legacy default stream) without
 *     synchronising.  They allocate nothing and keep no pointer.
 *   - `in` and `out` must not overlap, except out == in with k >= D (normalise in place;
 *     out == in with k < D is refused with ETN_ERR_SHAPE).
 *   - Every entry point is re-entrant and may be called from any OS thread (Go handlers
 *     run on arbitrary threads).
 */
#ifndef EMBED_TRUNCATE_NORM_H_
#define EMBED_TRUNCATE_NORM_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define ETN_API __attribute__((visibility("default")))
#else
#define ETN_API
#endif

#define ETN_VERSION 200 /* 0.2.0 */

#define ETN_F32 0
#define ETN_F64 1

#define ETN_TRUNCATE 0
#define ETN_TRUNCATE_RENORM 1
#define ETN_JSON_BASE64 0x10 /* etn_json_truncate_response only: cut base64 float32 embeddings too */

#define ETN_OK 0
#define ETN_ERR_NULL (-1)     /* a required pointer is NULL */
#define ETN_ERR_SHAPE (-2)    /* batch < 0, D <= 0, k <= 0, bad pitch or offsets */
#define ETN_ERR_ENUM (-3)     /* unknown dtype or mode */
#define ETN_ERR_ALIGN (-4)    /* pointer not aligned to the element size */
#define ETN_ERR_NOT_INIT (-5) /* host path used before etn_init */
#define ETN_ERR_DEVICE (-6)   /* bad device list / n_devices */
#define ETN_ERR_JSON (-8)      /* malformed JSON, a non-finite value, or a number outside float64 */
#define ETN_ERR_JSON_DUP (-9)  /* "data" or an item's "embedding" occurs twice: take the reference path */
#define ETN_ERR_TOO_LARGE (-7) /* more than ETN_STATIC_STAGE_BYTES asked of the _static entry point */
#define ETN_ERR_TICKET (-10)   /* etn_wait: unknown ticket, or one that was already waited for */

/* Library version (ETN_VERSION of the build).  No reference counterpart. */
ETN_API int etn_version(void);

/* Text for any code returned by this library (static storage; cudaGetErrorString for
 * positive codes).  Stands in for the err.Error() strings the handler logs with
 * slog.Error (handlers.go:1882). */
ETN_API const char *etn_strerror(int code);

/* The op, device-resident — north_star signature.  TRUNCATE_RENORM of the first k of D
 * components of each of `batch` rows.  Replaces the loop body handlers.go:2068-2076 for a
 * whole packed response at once.  k >= D is legal (normalise whole rows). */
ETN_API int embed_truncate_norm(int64_t batch, int64_t D, int64_t k, int dtype, const void *in,
                        void *out, void *stream);

/* Same shape contract, ETN_TRUNCATE only: bit-exact with the reference's
 * m["embedding"] = emb[:dims] (handlers.go:2075) on the packed matrix. */
ETN_API int embed_truncate(int64_t batch, int64_t D, int64_t k, int dtype, const void *in, void *out,
                   void *stream);

/* Ragged batch: row r keeps kr = min(k_per_row[r], D) components, written at element
 * offset out_offsets[r] of the packed output (CSR style; out_offsets has `batch` entries,
 * device memory, as has k_per_row).  One response may mix vector lengths — each vector is
 * handled on its own at handlers.go:2068-2076 — and one launch may carry many requests
 * with different `dimensions` (handlers.go:1874-1877). */
ETN_API int embed_truncate_norm_ragged(int64_t batch, int64_t D, const int32_t *k_per_row,
                               const int64_t *out_offsets, int dtype, int mode, const void *in,
                               void *out, void *stream);

/* Many requests, one launch, device-resident: row r is the k_per_row[r] elements that start at
 * element in_offsets[r] of in_base and is written at element out_offsets[r] of out_base (all
 * three arrays in device memory, `batch` entries; k_per_row already holds min(dims, len) —
 * handlers.go:2074 — because rows of different requests have different lengths).  This is the
 * launch the host runtime issues when it coalesces concurrent handlers (north_star item 2,
 * "one launch per batch"; handlers run one goroutine per connection, server.go:54), exposed
 * for callers whose vectors already live on the GPU.  Row starts must be 16-byte aligned. */
ETN_API int embed_truncate_norm_gather(int64_t batch, const int64_t *in_offsets,
                                       const int32_t *k_per_row, const int64_t *out_offsets,
                                       int dtype, int mode, const void *in_base, void *out_base,
                                       void *stream);

/* ---- host path: what the handler calls between proxyOpenAIEmbed (handlers.go:1879) and
 * WriteJSON (handlers.go:1894) once the response vectors are packed in host memory. ---- */

/* Creates, per device, a worker thread, four streams with their events and four staging slots
 * (64 MiB in + 64 MiB out each).  device_ids == NULL or n_devices == 0 selects every visible
 * device.  Call once from main (the reference wires its components at
 * core/cmd/core/main.go:49-51).  Calling it again is a no-op.  Environment, read here the way
 * config.Getenv reads its knobs (config.go:9-24):
 *   EMBED_GPU_MAX_WAIT_US        how long a worker that holds one small request waits for
 *                                others to share its launch (default 0: only requests that
 *                                are already queued are merged)
 *   EMBED_GPU_MIN_DEVICE_BYTES   kept-prefix bytes a device must get before a call is split
 *                                over one more device (default 4 MiB)
 *   EMBED_GPU_TIMING             every n-th staged chunk is timed with CUDA events for the
 *                                histograms (default 16; 1 = every chunk, 0 = none: the four
 *                                timestamped records cost a small request a quarter of its time)
 *   EMBED_GPU_PIECES             chunks a lone request is cut into so that its D2H overlaps
 *                                its own H2D (default 16, but no chunk under 2 MiB or over a slot) */
ETN_API int etn_init(const int *device_ids, int n_devices);

/* Releases what etn_init created (main.go:118-122, graceful shutdown). */
ETN_API void etn_shutdown(void);

/* Number of devices etn_init bound (0 before init).  No reference counterpart (the
 * reference never enumerates GPUs; its device table is Ollama hosts, router.go:286-322). */
ETN_API int etn_device_count(void);

/* Page-locked host memory for the packed request/response buffers (never Go memory, so
 * no Go pointer is retained by C and async DMA is legal).  NULL on failure.  Takes the place
 * of the []any backing arrays json.Decode allocates at handlers.go:2047-2050. */
ETN_API void *etn_alloc_pinned(size_t bytes);
ETN_API void etn_free_pinned(void *p);

/* End to end on host buffers, synchronous — the call that replaces
 * truncateEmbeddings(result, dims) at handlers.go:1891-1893.  The rows are split into
 * contiguous blocks over at most n_devices of the devices etn_init bound (0 = all of them;
 * the least loaded ones are taken, and a call brings in one more device only for every
 * EMBED_GPU_MIN_DEVICE_BYTES of kept prefixes, so a request-sized batch stays on one GPU and
 * concurrent requests spread over the box).  Each block is queued to its device's worker, which
 * copies it prefix-only (kk of D columns) in chunks, processes it and copies it back, the
 * copies of one chunk overlapping the kernel of another; small requests of several callers
 * that are queued together share one chunk and one launch.  host_in is [batch x D] with a row
 * pitch of in_pitch_bytes (>= D * elem size, or 0 for dense); host_out is dense [batch x kk]
 * and may be host_in itself when in_pitch_bytes == kk * elem size.  Pinned buffers
 * (etn_alloc_pinned) give full PCIe speed; pageable memory works but is slower.  No
 * collective is involved: rows are independent. */
ETN_API int embed_truncate_norm_host(int64_t batch, int64_t D, int64_t k, int dtype, int mode,
                             const void *host_in, int64_t in_pitch_bytes, void *host_out,
                             int n_devices);

/* Ragged batch end to end on host buffers, synchronous: row r keeps kr = min(k_per_row[r], D)
 * components and the results are packed back to back in row order (row r at element
 * sum_{i<r} k_i of host_out_packed).  This is what a handler-side batcher hands over when it
 * has merged responses of requests that resolved different `dimensions`
 * (handlers.go:1874-1877 resolves dims per request; handlers.go:2068-2076 treats every vector
 * on its own).  The rows are cut into contiguous ranges balanced by kept bytes, not by row
 * count (the cost of a row is its k), over at most n_devices devices; runs of rows with one k
 * travel as one prefix-only 2-D copy each.  k_per_row is host memory; every k must be
 * positive (ETN_ERR_SHAPE otherwise, like k <= 0 in the dense calls: the handler only calls
 * with dims > 0, handlers.go:1891). */
ETN_API int embed_truncate_norm_ragged_host(int64_t batch, int64_t D, const int32_t *k_per_row,
                                            int dtype, int mode, const void *host_in,
                                            int64_t in_pitch_bytes, void *host_out_packed,
                                            int n_devices);

/* The cut points embed_truncate_norm_ragged_host uses: cuts[0..n_shards] with shard g =
 * rows [cuts[g], cuts[g+1]), balanced by sum of min(k, D).  For callers that shard a
 * ragged batch themselves (one process per GPU).  No reference counterpart. */
ETN_API int etn_shard_ragged(int64_t batch, int64_t D, const int32_t *k_per_row, int n_shards,
                             int64_t *cuts);

/* The two host calls above without the wait: *ticket receives a handle, the buffers must stay
 * untouched until etn_wait(ticket) returns the call's result (each ticket is waited for
 * exactly once).  k_per_row == NULL selects the dense form with `k`, otherwise `k` is
 * ignored.  A handler that has several responses in hand — or a batcher goroutine serving
 * many connections — submits them all and then waits, so the device workers see the whole
 * queue and merge it into few launches (north_star item 2).  No reference counterpart: the
 * reference handles one response per goroutine (handlers.go:1891-1893). */
ETN_API int etn_submit_host(int64_t batch, int64_t D, int64_t k, const int32_t *k_per_row, int dtype,
                            int mode, const void *host_in, int64_t in_pitch_bytes, void *host_out,
                            int n_devices, int64_t *ticket);
ETN_API int etn_wait(int64_t ticket);

/* Size of the staging array built into the library image: 16 MiB of device memory on every
 * GPU that loads the library, held until the process exits. */
#define ETN_STATIC_STAGE_BYTES (16u << 20)

/* embed_truncate_norm_host for callers that cannot allocate device memory.  It has no
 * counterpart in the reference.
 *
 * Who needs it: embed_truncate_norm_host stages through buffers that etn_init obtains with
 * cudaMalloc.  A llmcore process that shares its GPU with a model server (Ollama, vLLM) often
 * finds the device already filled by that server's up-front reservation, so a 256 MiB
 * cudaMalloc at start-up or request time fails; a short-lived helper (a health probe such as
 * POST /v1/debug/test, a CLI check) does not want etn_init's set-up at all; and a handler on a
 * latency budget should not be exposed to an allocator call that may block behind other
 * work on the device.  This entry point asks no allocator for anything: not cudaMalloc, not
 * cudaFree, not a stream or event, nothing from any library.  Its device staging is an array
 * compiled into the library.  It needs no etn_init.
 *
 * Same arguments and results as embed_truncate_norm_host, except: it runs on the calling
 * thread's current device only, on `stream` (NULL = the legacy default stream), and the
 * kept prefixes must fit the array: batch * min(k, D) * element size <=
 * ETN_STATIC_STAGE_BYTES, otherwise ETN_ERR_TOO_LARGE and nothing is done (C1, 1024 x 1024
 * float64, is 8 MiB).  Synchronous: it returns after `stream` has drained.
 *
 * The first call on each device looks up the array's device address (set-up, once per
 * device); make that first call before starting any stream capture on that device.
 * All calls on a device share the one array, which keeps whatever the previous call left
 * in it; nothing in it has to be zero, since each call first overwrites the region it then
 * uses.  Two calls overlapping on the GPU would race on it, so the library holds a per-device
 * lock for the length of a call and concurrent callers queue.  Calls and rows are counted
 * with the host path's counters (ETN_CTR_HOST_*), a refused size with ETN_CTR_FALLBACKS. */
ETN_API int embed_truncate_norm_host_static(int64_t batch, int64_t D, int64_t k, int dtype, int mode,
                                            const void *host_in, int64_t in_pitch_bytes,
                                            void *host_out, void *stream);

/* ---- JSON packer: embedding arrays go from the provider's response
 * text into pinned memory and back without the map[string]any round trip of
 * json.NewDecoder(resp.Body).Decode(&result) (handlers.go:2047-2050) and
 * json.NewEncoder(w).Encode(payload) (helpers.go:11-15). ---- */

/* Rewrites an OpenAI-format embeddings response: every result["data"][i]["embedding"] that is
 * an array longer than `dims` is cut to its first `dims` elements (handlers.go:2068-2076:
 * non-object items, non-array embeddings and shorter vectors are left alone) and, with
 * ETN_TRUNCATE_RENORM, L2-renormalised on the GPU (needs etn_init; vectors whose kept
 * prefix holds anything but numbers are only cut, like the reference would).  In
 * ETN_TRUNCATE mode no number is parsed: the kept tokens are copied as text.  All bytes
 * outside those arrays pass through unchanged; formatted numbers follow encoding/json
 * (shortest round trip, 'e' form below 1e-6 and from 1e21).  dims <= 0 copies the input
 * (handlers.go:1891).  Returns the number of bytes written to `out`, or a negative code:
 * ETN_ERR_SHAPE = out_cap too small (len + 2 always suffices for TRUNCATE; allow 26 bytes per
 * kept number for RENORM), ETN_ERR_JSON / ETN_ERR_JSON_DUP = fall back to the reference path,
 * -1000 - c = CUDA error c.  mode | ETN_JSON_BASE64 also treats an
 * "embedding" that is a base64 string of little-endian float32 (OpenAI's
 * encoding_format "base64", parsed at handlers.go:1831 but never forwarded) as a vector:
 * its bytes go into pinned memory without any number parsing and come back re-encoded.  The
 * reference leaves such strings untouched (handlers.go:2073), hence the opt-in flag.
 * stats (may be NULL) receives {vectors renormalised on the GPU,
 * vectors cut as text, embedding arrays seen}. */
ETN_API int64_t etn_json_truncate_response(const char *json, size_t len, int64_t dims, int mode,
                                           char *out, size_t out_cap, int n_devices,
                                           int64_t *stats);

/* result["usage"]["prompt_tokens"] as proxyOpenAIEmbed extracts it for the token metric
 * (handlers.go:2052-2057: a float64 converted with int()); 0 when absent or not a number,
 * ETN_ERR_JSON for malformed text.  Lets the handler skip the map decode altogether. */
ETN_API int64_t etn_json_prompt_tokens(const char *json, size_t len);

/* "[a,b,c]" -> up to max_n float64 values, correctly rounded like strconv.ParseFloat (the
 * decoder's number path at handlers.go:2048); numbers past max_n are validated and skipped.
 * Returns the element count of the array or a negative code. */
ETN_API int64_t etn_json_parse_array(const char *text, size_t len, double *dst, int64_t max_n);

/* n float64 values -> "[a,b,c]" exactly as encoding/json writes them (helpers.go:14).
 * Returns bytes written, ETN_ERR_SHAPE if cap is too small (34 bytes per value suffice),
 * ETN_ERR_JSON for NaN/Inf (which Go's encoder refuses). */
ETN_API int64_t etn_json_format_array(const double *v, int64_t n, char *buf, size_t cap);

/* The per-number work of the two calls above on the GPU, one thread per number — the kernels
 * etn_json_truncate_response uses in ETN_TRUNCATE_RENORM mode, which ships the kept TEXT of the
 * embedding arrays to the device, parses, renormalises and formats there, and brings text back
 * (set ETN_JSON_GPU=0 to keep that work on the host cores).  All pointers are device memory.
 * parse: token i is text[tok_begin[i] .. +tok_len[i]); out[i] receives its float64 and
 * decided[i] = 1, or decided[i] = 0 when the fast algorithm (Eisel-Lemire) will not commit
 * — then the caller asks etn_json_parse_array; it never answers wrongly.  Stands in for
 * strconv.ParseFloat at handlers.go:2048.  format: v[i] is written at out + 32 * i exactly as
 * encoding/json spells it (helpers.go:14), len[i] bytes (<= 25); v[i] must be finite. */
ETN_API int etn_json_parse_tokens_device(const char *text, const int64_t *tok_begin,
                                         const int32_t *tok_len, int64_t n, double *out,
                                         uint8_t *decided, void *stream);
ETN_API int etn_json_format_numbers_device(const double *v, int64_t n, char *out, int32_t *len,
                                           void *stream);

/* ---- observability ---- */

/* Cumulative process-wide counters, to be exported as Prometheus counters beside
 * llmcore_embedding_requests_total and friends (core/internal/metrics/metrics.go:9-26,
 * registered at :63-70).  Index = ETN_CTR_*; etn_counter_name gives the metric name.
 * etn_get_counters writes min(capacity, ETN_CTR_COUNT) values and returns that number. */
#define ETN_CTR_DEVICE_CALLS 0    /* embed_truncate_norm / embed_truncate / _ragged calls */
#define ETN_CTR_DEVICE_ROWS 1     /* rows those calls covered */
#define ETN_CTR_KERNEL_LAUNCHES 2 /* kernels this library launched */
#define ETN_CTR_HOST_CALLS 3      /* embed_truncate_norm_host calls */
#define ETN_CTR_HOST_ROWS 4
#define ETN_CTR_H2D_BYTES 5       /* bytes copied host -> device by the host path */
#define ETN_CTR_D2H_BYTES 6
#define ETN_CTR_ERRORS 7          /* failed calls: CUDA errors and bad arguments */
#define ETN_CTR_FALLBACKS 8       /* times the GPU path stepped aside and the reference path is to be
                                     taken instead: ETN_ERR_JSON / _JSON_DUP / _TOO_LARGE / _NOT_INIT returns, and
                                     vectors reported through etn_note_fallback */
#define ETN_CTR_COALESCED 9       /* host-path requests that shared their launch with another request */
#define ETN_CTR_COUNT 10
ETN_API int etn_get_counters(int64_t *out, int capacity);
ETN_API const char *etn_counter_name(int index);

/* The wrapper's side of ETN_CTR_FALLBACKS: vectors it served with the reference reslice
 * (handlers.go:2075) because they could not be packed (non-numeric elements) or because a
 * call failed.  No reference counterpart. */
ETN_API void etn_note_fallback(int64_t vectors);

/* Fixed-bucket histograms, to be registered beside llmcore_embedding_duration_seconds
 * (metrics.go:18-26; its buckets are at :23).  etn_get_histogram fills upper_bounds[i]
 * (the last one is +Inf) and counts[i] = samples in bucket i (NOT cumulative; Prometheus'
 * `le` series is the running sum) and returns the number of buckets; *sum receives the sum of
 * all samples.  With both arrays NULL it only returns the number of buckets.  Kernel and
 * copy times come from CUDA events the host path records around every n-th staged chunk
 * (EMBED_GPU_TIMING, default 16: a sample, not a census) and reads when the chunk retires, so
 * nothing synchronises for the sake of a metric. */
#define ETN_HIST_BATCH_ROWS 0      /* rows per host-path call */
#define ETN_HIST_KERNEL_SECONDS 1  /* kernel time per staged chunk */
#define ETN_HIST_H2D_SECONDS 2     /* host-to-device copy time per staged chunk */
#define ETN_HIST_D2H_SECONDS 3     /* device-to-host copy time per staged chunk */
#define ETN_HIST_CALL_SECONDS 4    /* host-path call latency, submit to completion */
#define ETN_HIST_REQS_PER_LAUNCH 5 /* host-path requests sharing one staged chunk */
#define ETN_HIST_COUNT 6
ETN_API int etn_get_histogram(int which, double *upper_bounds, int64_t *counts, int capacity,
                              double *sum);
ETN_API const char *etn_histogram_name(int which);
ETN_API const char *etn_histogram_help(int which);

/* ---- bench/test utility ---- */

/* Writes the synthetic matrix of SURVEY.md §8(d) (as amended in DESIGN.md: signed 24-bit
 * integers * 2^-28) for rows [row0, row0+rows) x cols [0, ncols) of a logical
 * [.. x D] matrix into device memory with leading dimension ld.  No reference
 * counterpart (the reference has no benchmark for this path). */
ETN_API int etn_fill_synthetic(uint64_t seed, int64_t row0, int64_t rows, int64_t D, int64_t ncols,
                       int64_t ld, int dtype, void *dev_out, void *stream);

/* Raw copy ceiling of the calling thread's current device, measured the way the host path
 * copies: `reps` x `bytes` H2D from host_in (directions & 1) and/or D2H into host_out
 * (directions & 2) on two streams, wall-clock seconds into *seconds.  bench.py runs it on
 * every rank at once to state end-to-end throughput as a fraction of what the box's copy
 * path can do (profiles/r02_dma_matrix.txt).  Allocates `bytes` of device memory per
 * direction for the duration of the call.  No reference counterpart. */
ETN_API int etn_raw_copy_seconds(const void *host_in, void *host_out, size_t bytes, int reps,
                                 int directions, double *seconds);

#ifdef __cplusplus
}
#endif
#endif /* EMBED_TRUNCATE_NORM_H_ */
