// etn_host.cu — the host-buffer runtime of libembed_truncate_norm.so: what a handler calls
// between proxyOpenAIEmbed (core/internal/api/handlers.go:1879) and WriteJSON (:1894) once
// the response vectors sit packed in host memory.  It stands where the reference runs
// truncateEmbeddings(result, dims) (handlers.go:1891-1893, 2062-2078).
//
// Shape of the runtime (B200-first, not a translation of anything in the reference, which
// has no device code at all):
//   * etn_init starts ONE worker thread per bound GPU.  A worker owns four staging slots
//     (64 MiB in + 64 MiB out of device memory, a stream, four events, a pinned metadata
//     block each) and is the only thread that ever touches them, so there is no per-device
//     lock on the data path and no thread is created per call.
//   * A call (embed_truncate_norm_host, _ragged_host, etn_submit_host) becomes one job per
//     device it is spread over; jobs are queued to the workers and the caller sleeps on a
//     condition variable.  Small calls stay on one device — the least loaded one — and only
//     bring in another device for every EMBED_GPU_MIN_DEVICE_BYTES of kept prefixes; ragged
//     calls are cut by kept bytes, not by row count.
//   * A worker turns its queue into chunks: a piece of one large job, or SEVERAL small jobs
//     of different callers at once (north_star item 2: "one launch per batch" — handlers run
//     one goroutine per connection, core/internal/api/server.go:54, so concurrent requests
//     are the normal case).  A chunk is: prefix-only H2D (one 2-D copy per run of rows with
//     one k), one kernel, D2H.  Chunks rotate over the four slots, so the H2D of one overlaps
//     the kernel and the D2H of others; a slot is retired (stream synchronised, callers woken,
//     event times fed to the histograms) right before it is reused or when the queue is empty.
//   * Staging in and out are separate buffers: no kernel ever runs in place here.
// No CPU compute path: if CUDA reports an error the caller gets its code.
#include "../../include/embed_truncate_norm.h"
#include "etn_internal.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <climits>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <numeric>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

using etn_rt::aligned_to;
using etn_rt::count;
using etn_rt::finish;
using etn_rt::observe;
using Clock = std::chrono::steady_clock;

constexpr int kSlots = 4;                           // chunks in flight per device
constexpr int64_t kStageBytes = 64ll << 20;         // staging per slot, each direction
constexpr int64_t kMaxMetaRows = 256 << 10;         // rows of a chunk that needs per-row metadata
constexpr int64_t kMetaBytes = kMaxMetaRows * (8 + 8 + 4);
constexpr int64_t kMinChunkBytes = 2 << 20;         // smaller chunks cost more in submission than their overlap buys
constexpr int64_t kSmallJobBytes = 16ll << 20;      // jobs up to this size may share a chunk
constexpr size_t kMaxJobsPerChunk = 256;
constexpr int64_t kShortRun = 16;                   // ragged: runs shorter than this absorb their neighbours
// Waking a sleeping thread costs tens of microseconds (more inside a VM) — as much as a small
// request itself.  So an idle worker polls its queue for a moment before it sleeps, and a
// caller polls for completion for a moment before it sleeps; only a few callers may poll at
// once, so that pollers never crowd the workers off the cores.
constexpr int64_t kWorkerSpinUs = 200;
constexpr int64_t kCallerSpinUs = 500;
constexpr int kMaxSpinningCallers = 2;
std::atomic<int> g_spinning_callers{0};

inline void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
  __builtin_ia32_pause();
#endif
}

inline int64_t up16(int64_t b) { return (b + 15) & ~(int64_t)15; }

struct Runtime;
struct Call;

// One device's share of a call.
struct Job {
  Call* call = nullptr;
  int64_t row0 = 0, rows = 0;      // rows [row0, row0 + rows) of the call
  int64_t bytes = 0;               // kept-prefix bytes (what the device has to move each way)
  // worker-thread state
  int64_t cursor = 0;              // rows handed to chunks so far
  int pending = 0;                 // chunks in flight
  bool issued_all = false;
  int rc = 0;
};

struct Call {
  std::shared_ptr<Runtime> rt;     // keeps workers alive until the call has been waited for
  int dtype = 0, mode = 0;
  int64_t es = 4, batch = 0, D = 0, kk = 0;
  const int32_t* kpr = nullptr;    // ragged: per-row k (host memory of the caller)
  std::vector<int64_t> csum;       // ragged: csum[r] = kept elements before row r (batch + 1 entries)
  const char* hin = nullptr;
  int64_t pitch = 0;
  char* hout = nullptr;
  std::vector<Job> jobs;
  std::mutex mu;
  std::condition_variable cv;
  int remaining = 0;
  std::atomic<bool> done{false};
  int rc = 0;
  Clock::time_point t0;

  int64_t krow(int64_t r) const {
    if (!kpr) return kk;
    const int64_t k = kpr[r];
    return k < 0 ? 0 : (k > D ? D : k);
  }
};

struct Slot {
  cudaStream_t stream = nullptr;   // everything that touches this slot's buffers runs on this stream, in order
  char* d_in = nullptr;
  char* d_out = nullptr;
  char* d_meta = nullptr;          // device: in_off[int64] | out_off[int64] | k[int32], kMaxMetaRows each
  char* h_meta = nullptr;          // its pinned twin, written by the worker
  int64_t meta_seq = -1;           // sequence number of the last chunk that wrote h_meta
};

// One staged chunk in flight.  Chunks retire in the order they were issued.
struct Chunk {
  cudaEvent_t ev[4] = {};          // before H2D, after H2D, after the kernel, after D2H (timed chunks only)
  cudaEvent_t done = nullptr;      // after D2H, every chunk; no timestamp, so recording it is cheap
  int slot = 0;
  std::vector<Job*> jobs;          // jobs with a piece in this chunk
  bool has_kernel = false;
  bool timed = true;               // ev[0..3] recorded (kernel and copy times for the histograms)
  bool failed = false;             // issuing stopped half way: the stream was drained on the spot
};
constexpr int kChunkRecs = 16;     // chunks in flight per device (four deep on each slot's stream)

struct Device {
  int dev = -1;
  Slot slot[kSlots];
  Chunk rec[kChunkRecs];
  int64_t issued = 0, retired = 0; // chunk sequence numbers; chunk i lives in rec[i % kChunkRecs]
  std::thread th;
  std::mutex mu;
  std::condition_variable cv;
  std::deque<Job*> q;
  std::atomic<int> qn{0};          // q.size(), readable without the lock
  bool stop = false;
  std::atomic<int64_t> load{0};    // kept bytes queued or in flight
};

struct Runtime {
  std::vector<std::unique_ptr<Device>> devs;
  int64_t max_wait_us = 0;
  int64_t min_device_bytes = 4 << 20;
  // Measured (profiles/r02_latency.txt): the four timestamped event records cost a 64-row request
  // 20 of its 73 us and a C1 request 50 of 384 us, and four pieces beat eight once every piece
  // carries its own completion event.  Hence: every 16th chunk is timed, and a job is cut into
  // 16 chunks of at least 2 MiB and at most one slot (C1: four chunks of 2 MiB; 512 MiB: sixteen
  // of 32 MiB, so the un-overlapped first H2D and last D2H stay a small part of the whole).
  int64_t timing_every = 16;       // EMBED_GPU_TIMING: time every n-th staged chunk (1 = all, 0 = none)
  int64_t pieces = 16;             // EMBED_GPU_PIECES: chunks a lone job is cut into (floor 2 MiB each)
  ~Runtime();
};

std::mutex g_mu;
std::shared_ptr<Runtime> g_rt;                       // guarded by g_mu

std::mutex g_ticket_mu;
std::unordered_map<int64_t, Call*> g_tickets;        // guarded by g_ticket_mu
int64_t g_next_ticket = 1;

int64_t env_i64(const char* name, int64_t dflt) {    // config.GetenvInt's rule (config.go:17-24): unparsable -> default
  const char* v = getenv(name);
  if (!v || !*v) return dflt;
  char* end = nullptr;
  const long long n = strtoll(v, &end, 10);
  return (end && *end == 0 && n >= 0) ? (int64_t)n : dflt;
}

// ------------------------------------------------------------------- completion ----
void complete(Device* d, Job* j) {
  d->load.fetch_sub(j->bytes, std::memory_order_relaxed);
  Call* c = j->call;
  const int rc = j->rc;
  std::lock_guard<std::mutex> lk(c->mu);             // after this block `c` and `j` may be gone
  if (rc && !c->rc) c->rc = rc;
  if (--c->remaining == 0) {
    c->done.store(true, std::memory_order_release);
    c->cv.notify_all();
  }
}

// Waits for the oldest chunk in flight, feeds its event times to the histograms and wakes the
// callers whose last piece it carried.
void retire_oldest(Device* d) {
  Chunk& c = d->rec[d->retired % kChunkRecs];
  cudaError_t e = cudaSuccess;
  if (!c.failed) {
    e = cudaEventSynchronize(c.done);
    if (e != cudaSuccess) cudaGetLastError();
  }
  for (Job* j : c.jobs) {                              // callers first, bookkeeping after
    if (e != cudaSuccess && !j->rc) j->rc = (int)e;
    if (--j->pending == 0 && j->issued_all) complete(d, j);
  }
  c.jobs.clear();
  if (!c.failed && c.timed && e == cudaSuccess) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c.ev[0], c.ev[1]) == cudaSuccess) observe(ETN_HIST_H2D_SECONDS, ms * 1e-3);
    if (c.has_kernel && cudaEventElapsedTime(&ms, c.ev[1], c.ev[2]) == cudaSuccess)
      observe(ETN_HIST_KERNEL_SECONDS, ms * 1e-3);
    if (cudaEventElapsedTime(&ms, c.ev[2], c.ev[3]) == cudaSuccess) observe(ETN_HIST_D2H_SECONDS, ms * 1e-3);
  }
  ++d->retired;
}

// h_meta of a slot is written by the host: the chunk that last used it must have finished.
void wait_for_meta(Device* d, Slot& s) {
  while (d->retired <= s.meta_seq) retire_oldest(d);
}

// ---------------------------------------------------------------------- chunks ----
#define ETN_TRY(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { cudaGetLastError(); return (int)e_; } } while (0)

int h2d_rows(Slot& s, int64_t dst_off, int64_t dst_pitch, const char* src, int64_t src_pitch, int64_t width,
             int64_t rows) {
  if (width <= 0 || rows <= 0) return 0;
  if (src_pitch == width && dst_pitch == width)
    ETN_TRY(cudaMemcpyAsync(s.d_in + dst_off, src, (size_t)(width * rows), cudaMemcpyHostToDevice, s.stream));
  else
    ETN_TRY(cudaMemcpy2DAsync(s.d_in + dst_off, (size_t)dst_pitch, src, (size_t)src_pitch, (size_t)width,
                              (size_t)rows, cudaMemcpyHostToDevice, s.stream));
  return 0;
}

struct Seg { Job* job; int64_t r0, n; };             // rows [r0, r0 + n) of the CALL (absolute row numbers)

// Dense segments (each with its own k, D, pitch and buffers; one element type and mode).
// One k for all of them: the staged rows form one dense matrix and the dense kernels run on it.
// Mixed k: per-row metadata and the gathering page-ring kernel.
int issue_dense(Device* d, Slot& s, Chunk& ck, const std::vector<Seg>& segs) {
  const Call* c0 = segs[0].job->call;
  const int64_t es = c0->es;
  bool uniform = true;
  int64_t total = 0;
  for (const Seg& g : segs) { uniform = uniform && g.job->call->kk == c0->kk; total += g.n; }
  int64_t* m_in = reinterpret_cast<int64_t*>(s.h_meta);
  int64_t* m_out = m_in + kMaxMetaRows;
  int32_t* m_k = reinterpret_cast<int32_t*>(m_out + kMaxMetaRows);
  if (!uniform) { wait_for_meta(d, s); s.meta_seq = d->issued; }
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[0], s.stream));
  std::vector<int64_t> out_off(segs.size());
  int64_t in_b = 0, out_b = 0, row = 0;
  for (size_t i = 0; i < segs.size(); ++i) {
    const Seg& g = segs[i];
    const Call* c = g.job->call;
    const int64_t w = c->kk * es, ip = up16(w);
    if (int rc = h2d_rows(s, in_b, ip, c->hin + g.r0 * c->pitch, c->pitch, w, g.n)) return rc;
    if (!uniform) {
      out_b = up16(out_b);
      for (int64_t r = 0; r < g.n; ++r, ++row) {
        m_in[row] = (in_b + r * ip) / es;
        m_out[row] = out_b / es + r * c->kk;
        m_k[row] = (int32_t)c->kk;
      }
    }
    out_off[i] = out_b;
    in_b += g.n * ip;
    out_b += g.n * w;
  }
  const int64_t w0 = c0->kk * es, ip0 = up16(w0);
  ck.has_kernel = false;
  bool from_in = false;                               // TRUNCATE of one-k segments: copy back what came in
  if (uniform && c0->mode == ETN_TRUNCATE) {
    from_in = true;
    if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[1], s.stream));
  } else if (uniform) {
    if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[1], s.stream));
    if (int rc = etn_rt::run_dense(total, ip0 / es, c0->kk, c0->dtype, c0->mode, s.d_in, s.d_out, s.stream)) return rc;
    ck.has_kernel = true;
  } else {
    int64_t* d_in_off = reinterpret_cast<int64_t*>(s.d_meta);
    int64_t* d_out_off = d_in_off + kMaxMetaRows;
    int32_t* d_k = reinterpret_cast<int32_t*>(d_out_off + kMaxMetaRows);
    ETN_TRY(cudaMemcpyAsync(d_in_off, m_in, (size_t)total * 8, cudaMemcpyHostToDevice, s.stream));
    ETN_TRY(cudaMemcpyAsync(d_out_off, m_out, (size_t)total * 8, cudaMemcpyHostToDevice, s.stream));
    ETN_TRY(cudaMemcpyAsync(d_k, m_k, (size_t)total * 4, cudaMemcpyHostToDevice, s.stream));
    if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[1], s.stream));
    if (int rc = etn_rt::run_rows(total, INT64_MAX / 16, d_k, d_out_off, d_in_off, c0->dtype, c0->mode, s.d_in,
                                  s.d_out, s.stream))
      return rc;
    ck.has_kernel = true;
  }
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[2], s.stream));
  int64_t in_seg = 0;
  for (size_t i = 0; i < segs.size(); ++i) {
    const Seg& g = segs[i];
    const Call* c = g.job->call;
    const int64_t w = c->kk * es, ip = up16(w);
    char* dst = c->hout + g.r0 * w;
    if (!from_in)
      ETN_TRY(cudaMemcpyAsync(dst, s.d_out + out_off[i], (size_t)(g.n * w), cudaMemcpyDeviceToHost, s.stream));
    else if (ip == w)
      ETN_TRY(cudaMemcpyAsync(dst, s.d_in + in_seg, (size_t)(g.n * w), cudaMemcpyDeviceToHost, s.stream));
    else
      ETN_TRY(cudaMemcpy2DAsync(dst, (size_t)w, s.d_in + in_seg, (size_t)ip, (size_t)w, (size_t)g.n,
                                cudaMemcpyDeviceToHost, s.stream));
    in_seg += g.n * ip;
  }
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[3], s.stream));
  ETN_TRY(cudaEventRecord(ck.done, s.stream));
  return 0;
}

// How many rows of a ragged job, from row r0 on, make the next chunk: until `target` staged
// bytes, the stage or the metadata block is full.  Rows travel in blocks of one width: a run
// of rows with one k is one block however long; a run shorter than kShortRun absorbs its
// neighbours and is copied at their largest k (a few over-fetched bytes instead of one copy
// call per row).  Fills blocks as (rows, kmax) pairs.
int64_t plan_ragged(const Call* c, int64_t r0, int64_t r_end, int64_t target,
                    std::vector<std::pair<int64_t, int64_t>>& blocks) {
  const int64_t es = c->es;
  int64_t in_b = 0, rows = 0, r = r0;
  blocks.clear();
  while (r < r_end && in_b < target && rows < kMaxMetaRows) {
    int64_t kmax = c->krow(r), n = 1;
    if (in_b + up16(kmax * es) > kStageBytes) break;
    while (r + n < r_end && rows + n < kMaxMetaRows) {
      const int64_t kn = c->krow(r + n);
      int64_t km = kmax;
      if (kn != kmax) {
        if (n >= kShortRun) break;
        km = std::max(kmax, kn);
      }
      if (in_b + (n + 1) * up16(km * es) > kStageBytes) break;
      if (in_b + n * up16(km * es) >= target && n >= kShortRun) break;
      kmax = km;
      ++n;
    }
    blocks.emplace_back(n, kmax);
    in_b += n * up16(kmax * es);
    rows += n;
    r += n;
  }
  return rows;
}

int issue_ragged(Device* d, Slot& s, Chunk& ck, Job* j, int64_t r0, const std::vector<std::pair<int64_t, int64_t>>& blocks,
                 int64_t rows) {
  const Call* c = j->call;
  const int64_t es = c->es;
  int64_t* m_in = reinterpret_cast<int64_t*>(s.h_meta);
  int64_t* m_out = m_in + kMaxMetaRows;
  int32_t* m_k = reinterpret_cast<int32_t*>(m_out + kMaxMetaRows);
  wait_for_meta(d, s);
  s.meta_seq = d->issued;
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[0], s.stream));
  const int64_t out0 = c->csum[r0];
  int64_t in_b = 0, r = r0, row = 0;
  for (const auto& b : blocks) {
    const int64_t w = b.second * es, ip = up16(w);
    if (int rc = h2d_rows(s, in_b, ip, c->hin + r * c->pitch, c->pitch, w, b.first)) return rc;
    for (int64_t i = 0; i < b.first; ++i, ++row) {
      m_in[row] = (in_b + i * ip) / es;
      m_out[row] = c->csum[r + i] - out0;
      m_k[row] = (int32_t)c->krow(r + i);
    }
    in_b += b.first * ip;
    r += b.first;
  }
  int64_t* d_in_off = reinterpret_cast<int64_t*>(s.d_meta);
  int64_t* d_out_off = d_in_off + kMaxMetaRows;
  int32_t* d_k = reinterpret_cast<int32_t*>(d_out_off + kMaxMetaRows);
  ETN_TRY(cudaMemcpyAsync(d_in_off, m_in, (size_t)rows * 8, cudaMemcpyHostToDevice, s.stream));
  ETN_TRY(cudaMemcpyAsync(d_out_off, m_out, (size_t)rows * 8, cudaMemcpyHostToDevice, s.stream));
  ETN_TRY(cudaMemcpyAsync(d_k, m_k, (size_t)rows * 4, cudaMemcpyHostToDevice, s.stream));
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[1], s.stream));
  if (int rc = etn_rt::run_rows(rows, INT64_MAX / 16, d_k, d_out_off, d_in_off, c->dtype, c->mode, s.d_in, s.d_out,
                                s.stream))
    return rc;
  ck.has_kernel = true;
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[2], s.stream));
  const int64_t out_n = c->csum[r0 + rows] - out0;
  if (out_n > 0)
    ETN_TRY(cudaMemcpyAsync(c->hout + out0 * es, s.d_out, (size_t)(out_n * es), cudaMemcpyDeviceToHost, s.stream));
  if (ck.timed) ETN_TRY(cudaEventRecord(ck.ev[3], s.stream));
  ETN_TRY(cudaEventRecord(ck.done, s.stream));
  return 0;
}

// ---------------------------------------------------------------------- worker ----
bool small_dense(const Job* j) { return !j->call->kpr && j->cursor == 0 && j->bytes <= kSmallJobBytes; }

void worker(Runtime* rt, Device* d) {
  if (cudaSetDevice(d->dev) != cudaSuccess) cudaGetLastError();
  std::vector<Job*> take;
  std::vector<Seg> segs;
  std::vector<std::pair<int64_t, int64_t>> blocks;
  for (;;) {
    take.clear();
    const int64_t inflight = d->issued - d->retired;
    if (inflight == 0 && d->qn.load(std::memory_order_acquire) == 0) {
      const auto t0 = Clock::now();
      while (d->qn.load(std::memory_order_acquire) == 0 &&
             Clock::now() - t0 < std::chrono::microseconds(kWorkerSpinUs))
        cpu_relax();
    }
    {
      std::unique_lock<std::mutex> lk(d->mu);
      if (inflight == 0) d->cv.wait(lk, [&] { return d->stop || !d->q.empty(); });
      if (d->q.empty()) {
        if (inflight == 0) { if (d->stop) break; continue; }
      } else {
        Job* h = d->q.front();
        if (rt->max_wait_us > 0 && inflight == 0 && d->q.size() == 1 && small_dense(h) && !d->stop)
          d->cv.wait_for(lk, std::chrono::microseconds(rt->max_wait_us), [&] { return d->stop || d->q.size() > 1; });
        take.push_back(h);
        if (small_dense(h) && d->q.size() > 1) {
          // several small requests are waiting: they share this chunk, whole
          const Call* c0 = h->call;
          int64_t in_b = h->rows * up16(c0->kk * c0->es), rows = h->rows;
          bool uniform = true;
          for (size_t i = 1; i < d->q.size() && take.size() < kMaxJobsPerChunk; ++i) {
            Job* j = d->q[i];
            const Call* c = j->call;
            if (!small_dense(j) || c->dtype != c0->dtype || c->mode != c0->mode) break;
            const int64_t add = j->rows * up16(c->kk * c->es);
            const bool uni = uniform && c->kk == c0->kk;
            if (in_b + add + 16 * (int64_t)take.size() > kStageBytes) break;
            if (!uni && rows + j->rows > kMaxMetaRows) break;
            uniform = uni;
            in_b += add;
            rows += j->rows;
            take.push_back(j);
          }
          if (take.size() > 1) {
            for (size_t i = 0; i < take.size(); ++i) d->q.pop_front();
            d->qn.store((int)d->q.size(), std::memory_order_release);
          }
        }
      }
    }
    if (take.empty()) {                                // nothing to issue: wait for the oldest chunk, wake its callers
      retire_oldest(d);
      continue;
    }
    if (inflight == kChunkRecs) retire_oldest(d);
    Chunk& ck = d->rec[d->issued % kChunkRecs];
    ck.slot = (int)(d->issued % kSlots);
    ck.has_kernel = false;
    ck.failed = false;
    ck.timed = rt->timing_every > 0 && d->issued % rt->timing_every == 0;
    Slot& s = d->slot[ck.slot];
    int rc = 0;
    segs.clear();
    if (take.size() > 1) {
      for (Job* j : take) { segs.push_back({j, j->row0, j->rows}); j->cursor = j->rows; j->issued_all = true; }
      rc = issue_dense(d, s, ck, segs);
      count(ETN_CTR_COALESCED, (int64_t)take.size());
    } else {
      Job* j = take[0];
      const Call* c = j->call;
      const int64_t left = j->rows - j->cursor;
      // a job is cut into a few chunks so that its D2H overlaps its own H2D
      const int64_t target = std::min(kStageBytes, std::max((j->bytes + rt->pieces - 1) / rt->pieces, kMinChunkBytes));
      int64_t n = 0;
      if (!c->kpr) {
        const int64_t w = c->kk * c->es, ip = up16(w);
        n = std::min(left, std::max<int64_t>(1, std::min(kStageBytes / ip, (target + w - 1) / w)));
        segs.push_back({j, j->row0 + j->cursor, n});
        rc = issue_dense(d, s, ck, segs);
      } else {
        n = plan_ragged(c, j->row0 + j->cursor, j->row0 + j->rows, target, blocks);
        rc = n > 0 ? issue_ragged(d, s, ck, j, j->row0 + j->cursor, blocks, n) : ETN_ERR_SHAPE;
      }
      j->cursor += n;
      if (rc != 0 || j->cursor >= j->rows) {
        j->issued_all = true;
        std::lock_guard<std::mutex> lk(d->mu);
        d->q.pop_front();                              // only this thread pops, so the head is still j
        d->qn.store((int)d->q.size(), std::memory_order_release);
      }
    }
    if (rc != 0) {                                     // stopped half way: drain what did get enqueued, now
      if (cudaStreamSynchronize(s.stream) != cudaSuccess) cudaGetLastError();
      ck.failed = true;
    }
    for (Job* j : take) {
      if (rc && !j->rc) j->rc = rc;
      ++j->pending;
      ck.jobs.push_back(j);
    }
    observe(ETN_HIST_REQS_PER_LAUNCH, (double)take.size());
    ++d->issued;
    // wake callers whose chunks have finished in the meantime, without blocking
    while (d->retired < d->issued) {
      const Chunk& o = d->rec[d->retired % kChunkRecs];
      if (!o.failed && cudaEventQuery(o.done) != cudaSuccess) { cudaGetLastError(); break; }
      retire_oldest(d);
    }
  }
}

Runtime::~Runtime() {
  for (auto& d : devs) {
    { std::lock_guard<std::mutex> lk(d->mu); d->stop = true; }
    d->cv.notify_all();
  }
  for (auto& d : devs) if (d->th.joinable()) d->th.join();
  int prev = 0;
  const bool have_prev = cudaGetDevice(&prev) == cudaSuccess;
  for (auto& d : devs) {
    if (cudaSetDevice(d->dev) != cudaSuccess) { cudaGetLastError(); continue; }
    for (Chunk& c : d->rec) {
      for (cudaEvent_t e : c.ev) if (e) cudaEventDestroy(e);
      if (c.done) cudaEventDestroy(c.done);
    }
    for (Slot& s : d->slot) {
      if (s.stream) { cudaStreamSynchronize(s.stream); cudaStreamDestroy(s.stream); }
      if (s.d_in) cudaFree(s.d_in);
      if (s.d_out) cudaFree(s.d_out);
      if (s.d_meta) cudaFree(s.d_meta);
      if (s.h_meta) cudaFreeHost(s.h_meta);
    }
  }
  if (have_prev) cudaSetDevice(prev);
}

std::shared_ptr<Runtime> runtime() {
  std::lock_guard<std::mutex> lk(g_mu);
  return g_rt;
}

// ---------------------------------------------------------------------- submit ----
// Contiguous row ranges balanced by kept elements: cuts[g] = first row r with
// csum[r] * n >= total * g.
void cut_by_kept(const std::vector<int64_t>& csum, int n, std::vector<int64_t>& cuts) {
  const int64_t batch = (int64_t)csum.size() - 1, total = csum[batch];
  cuts.assign((size_t)n + 1, batch);
  cuts[0] = 0;
  for (int g = 1; g < n; ++g) {
    const __int128 want = (__int128)total * g;
    int64_t lo = cuts[g - 1], hi = batch;
    while (lo < hi) {
      const int64_t mid = lo + (hi - lo) / 2;
      if ((__int128)csum[mid] * n >= want) hi = mid; else lo = mid + 1;
    }
    cuts[g] = lo;
  }
}

// false if some k is not positive (the dense entry points refuse k <= 0 the same way:
// the handler never calls with dims <= 0, handlers.go:1891)
bool build_csum(int64_t batch, int64_t D, const int32_t* kpr, std::vector<int64_t>& csum) {
  csum.resize((size_t)batch + 1);
  int64_t acc = 0;
  bool ok = true;
  for (int64_t r = 0; r < batch; ++r) {
    csum[r] = acc;
    const int64_t k = kpr[r];
    ok = ok && k > 0;
    acc += k < 0 ? 0 : (k > D ? D : k);
  }
  csum[batch] = acc;
  return ok;
}

int submit(int64_t batch, int64_t D, int64_t k, const int32_t* kpr, int dtype, int mode, const void* host_in,
           int64_t pitch, void* host_out, int n_devices, Call** out_call) {
  *out_call = nullptr;
  if ((dtype != ETN_F32 && dtype != ETN_F64) || (mode != ETN_TRUNCATE && mode != ETN_TRUNCATE_RENORM))
    return ETN_ERR_ENUM;
  if (batch < 0 || D <= 0 || (!kpr && k <= 0)) return ETN_ERR_SHAPE;
  const int64_t es = dtype == ETN_F32 ? 4 : 8;
  if (pitch == 0) pitch = D * es;
  if (pitch < D * es || pitch % es != 0 || pitch > INT_MAX) return ETN_ERR_SHAPE;
  if (n_devices < 0) return ETN_ERR_DEVICE;
  if (batch == 0) return ETN_OK;
  if (!host_in || !host_out) return ETN_ERR_NULL;
  if (!aligned_to(host_in, (size_t)es) || !aligned_to(host_out, (size_t)es)) return ETN_ERR_ALIGN;
  std::shared_ptr<Runtime> rt = runtime();
  if (!rt) return ETN_ERR_NOT_INIT;
  const int ndev = (int)rt->devs.size();
  if (n_devices == 0) n_devices = ndev;
  if (n_devices > ndev) return ETN_ERR_DEVICE;
  const int64_t kk = kpr ? 0 : std::min(k, D);
  if (up16(std::min(kpr ? D : kk, D) * es) > kStageBytes) return ETN_ERR_SHAPE;   // one row prefix larger than a slot

  std::unique_ptr<Call> c(new Call());
  c->rt = rt;
  c->dtype = dtype; c->mode = mode; c->es = es; c->batch = batch; c->D = D; c->kk = kk; c->kpr = kpr;
  c->hin = (const char*)host_in; c->pitch = pitch; c->hout = (char*)host_out;
  c->t0 = Clock::now();
  int64_t total;
  if (kpr) {
    if (!build_csum(batch, D, kpr, c->csum)) return ETN_ERR_SHAPE;
    total = c->csum[batch] * es;
  }
  else total = batch * kk * es;

  // how many devices, and which: one more device per min_device_bytes, the least loaded first
  int nd = (int)std::max<int64_t>(1, std::min<int64_t>({(int64_t)n_devices, batch, total / std::max<int64_t>(1, rt->min_device_bytes)}));
  // Least loaded first; among equally loaded ones the lowest ordinal, NOT a rotation: a lone
  // caller then keeps hitting the same worker, which is still polling its queue, instead of
  // waking a sleeping one on every call (measured on the 8-GPU box: p99 of a C1 request 1.06 ms
  // with a rotating start, against a p50 of 0.19 ms).  Concurrent callers spread through the load.
  std::vector<int> order((size_t)ndev);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    return rt->devs[a]->load.load(std::memory_order_relaxed) < rt->devs[b]->load.load(std::memory_order_relaxed);
  });

  std::vector<int64_t> cuts;
  if (kpr) {
    cut_by_kept(c->csum, nd, cuts);
  } else {
    const int64_t per = (batch + nd - 1) / nd;       // contiguous row blocks
    cuts.resize((size_t)nd + 1);
    for (int g = 0; g <= nd; ++g) cuts[g] = std::min(batch, g * per);
  }
  c->jobs.reserve((size_t)nd);
  std::vector<int> where;
  for (int g = 0; g < nd; ++g) {
    if (cuts[g + 1] <= cuts[g]) continue;
    Job j;
    j.call = c.get();
    j.row0 = cuts[g]; j.rows = cuts[g + 1] - cuts[g];
    j.bytes = kpr ? (c->csum[cuts[g + 1]] - c->csum[cuts[g]]) * es : j.rows * kk * es;
    c->jobs.push_back(j);
    where.push_back(order[g]);
  }
  c->remaining = (int)c->jobs.size();
  count(ETN_CTR_HOST_CALLS); count(ETN_CTR_HOST_ROWS, batch);
  observe(ETN_HIST_BATCH_ROWS, (double)batch);
  Call* raw = c.release();
  *out_call = raw;
  const size_t njobs = raw->jobs.size();              // `raw` may complete the moment its last job is queued
  for (size_t i = 0; i < njobs; ++i) {
    Device* d = rt->devs[where[i]].get();
    Job* j = &raw->jobs[i];
    d->load.fetch_add(j->bytes, std::memory_order_relaxed);
    {
      std::lock_guard<std::mutex> lk(d->mu);
      d->q.push_back(j);
      d->qn.store((int)d->q.size(), std::memory_order_release);
    }
    d->cv.notify_one();
  }
  return ETN_OK;
}

int wait_call(Call* c) {
  int rc;
  if (g_spinning_callers.fetch_add(1, std::memory_order_relaxed) < kMaxSpinningCallers) {
    const auto t0 = Clock::now();
    while (!c->done.load(std::memory_order_acquire) && Clock::now() - t0 < std::chrono::microseconds(kCallerSpinUs))
      cpu_relax();
  }
  g_spinning_callers.fetch_sub(1, std::memory_order_relaxed);
  {
    std::unique_lock<std::mutex> lk(c->mu);            // also when polling saw `done`: the worker has left the lock
    c->cv.wait(lk, [&] { return c->remaining == 0; });
    rc = c->rc;
  }
  observe(ETN_HIST_CALL_SECONDS, std::chrono::duration<double>(Clock::now() - c->t0).count());
  if (rc == 0) {
    const int64_t bytes = (c->kpr ? c->csum[c->batch] : c->batch * c->kk) * c->es;
    count(ETN_CTR_H2D_BYTES, bytes); count(ETN_CTR_D2H_BYTES, bytes);
  }
  delete c;                                            // may drop the last reference to a runtime that was shut down
  return rc;
}

}  // namespace

namespace etn_rt {
int least_loaded_device() {
  std::shared_ptr<Runtime> rt = runtime();
  if (!rt || rt->devs.empty()) return -1;
  Device* best = rt->devs[0].get();
  for (auto& d : rt->devs)
    if (d->load.load(std::memory_order_relaxed) < best->load.load(std::memory_order_relaxed)) best = d.get();
  return best->dev;
}
}  // namespace etn_rt

// =============================================================== C-ABI ======
extern "C" {

int etn_init(const int* device_ids, int n_devices) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (g_rt) return ETN_OK;
  int visible = 0;
  cudaError_t e = cudaGetDeviceCount(&visible);
  if (e != cudaSuccess) { cudaGetLastError(); return (int)e; }
  std::vector<int> ids;
  if (!device_ids || n_devices <= 0) {
    for (int i = 0; i < visible; ++i) ids.push_back(i);
  } else {
    for (int i = 0; i < n_devices; ++i) {
      if (device_ids[i] < 0 || device_ids[i] >= visible) return ETN_ERR_DEVICE;
      ids.push_back(device_ids[i]);
    }
  }
  if (ids.empty()) return ETN_ERR_DEVICE;
  int prev = 0;
  cudaGetDevice(&prev);
  auto rt = std::make_shared<Runtime>();
  rt->max_wait_us = env_i64("EMBED_GPU_MAX_WAIT_US", 0);
  rt->min_device_bytes = std::max<int64_t>(1, env_i64("EMBED_GPU_MIN_DEVICE_BYTES", 4 << 20));
  rt->timing_every = env_i64("EMBED_GPU_TIMING", rt->timing_every);
  rt->pieces = std::max<int64_t>(1, env_i64("EMBED_GPU_PIECES", rt->pieces));
  int rc = 0;
  for (int id : ids) {
    std::unique_ptr<Device> d(new Device());
    d->dev = id;
    if ((e = cudaSetDevice(id)) != cudaSuccess) rc = (int)e;
    for (int i = 0; i < kSlots && rc == 0; ++i) {
      Slot& s = d->slot[i];
      if ((e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking)) != cudaSuccess) rc = (int)e;
      if (rc == 0 && (e = cudaMalloc(&s.d_in, (size_t)kStageBytes)) != cudaSuccess) rc = (int)e;
      if (rc == 0 && (e = cudaMalloc(&s.d_out, (size_t)kStageBytes + 4096)) != cudaSuccess) rc = (int)e;
      if (rc == 0 && (e = cudaMalloc(&s.d_meta, (size_t)kMetaBytes)) != cudaSuccess) rc = (int)e;
      if (rc == 0 && (e = cudaHostAlloc(&s.h_meta, (size_t)kMetaBytes, cudaHostAllocPortable)) != cudaSuccess) rc = (int)e;
    }
    for (int i = 0; i < kChunkRecs && rc == 0; ++i) {
      for (int v = 0; v < 4 && rc == 0; ++v)
        if ((e = cudaEventCreate(&d->rec[i].ev[v])) != cudaSuccess) rc = (int)e;
      if (rc == 0 && (e = cudaEventCreateWithFlags(&d->rec[i].done, cudaEventDisableTiming)) != cudaSuccess) rc = (int)e;
    }
    rt->devs.push_back(std::move(d));                 // owned by the runtime from here on, freed by ~Runtime
    if (rc) break;
  }
  cudaSetDevice(prev);
  if (rc) { cudaGetLastError(); return rc; }           // rt goes out of scope: everything made so far is released
  for (auto& d : rt->devs) d->th = std::thread(worker, rt.get(), d.get());
  g_rt = rt;
  return ETN_OK;
}

void etn_shutdown(void) {
  std::shared_ptr<Runtime> rt;
  {
    std::lock_guard<std::mutex> lock(g_mu);
    rt.swap(g_rt);
  }
  etn_rt::release_json_staging();
  etn_rt::json_gpu_shutdown();
  // Calls still in flight hold their own reference; the workers drain their queues and the
  // last reference (here or in the last waiter) joins them and frees the device memory.
}

int etn_device_count(void) {
  std::shared_ptr<Runtime> rt = runtime();
  return rt ? (int)rt->devs.size() : 0;
}

void* etn_alloc_pinned(size_t bytes) {
  void* p = nullptr;
  if (bytes == 0) return nullptr;
  if (cudaHostAlloc(&p, bytes, cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}

void etn_free_pinned(void* p) {
  if (p) cudaFreeHost(p);
}

int embed_truncate_norm_host(int64_t batch, int64_t D, int64_t k, int dtype, int mode,
                             const void* host_in, int64_t in_pitch_bytes, void* host_out,
                             int n_devices) {
  Call* c = nullptr;
  const int rc = submit(batch, D, k, nullptr, dtype, mode, host_in, in_pitch_bytes, host_out, n_devices, &c);
  if (rc != 0 || !c) return rc == 0 ? ETN_OK : finish(rc);
  return finish(wait_call(c));
}

int embed_truncate_norm_ragged_host(int64_t batch, int64_t D, const int32_t* k_per_row, int dtype, int mode,
                                    const void* host_in, int64_t in_pitch_bytes, void* host_out_packed,
                                    int n_devices) {
  if (batch > 0 && !k_per_row) return ETN_ERR_NULL;
  static const int32_t none = 1;
  Call* c = nullptr;
  const int rc = submit(batch, D, 0, batch > 0 ? k_per_row : &none, dtype, mode, host_in, in_pitch_bytes,
                        host_out_packed, n_devices, &c);
  if (rc != 0 || !c) return rc == 0 ? ETN_OK : finish(rc);
  return finish(wait_call(c));
}

int etn_submit_host(int64_t batch, int64_t D, int64_t k, const int32_t* k_per_row, int dtype, int mode,
                    const void* host_in, int64_t in_pitch_bytes, void* host_out, int n_devices,
                    int64_t* ticket) {
  if (!ticket) return ETN_ERR_NULL;
  *ticket = 0;
  Call* c = nullptr;
  const int rc = submit(batch, D, k, k_per_row, dtype, mode, host_in, in_pitch_bytes, host_out, n_devices, &c);
  if (rc != 0) return finish(rc);
  std::lock_guard<std::mutex> lk(g_ticket_mu);
  *ticket = g_next_ticket++;
  g_tickets[*ticket] = c;                              // nullptr for an empty batch: nothing to wait for
  return ETN_OK;
}

int etn_wait(int64_t ticket) {
  Call* c = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_ticket_mu);
    auto it = g_tickets.find(ticket);
    if (it == g_tickets.end()) return ETN_ERR_TICKET;
    c = it->second;
    g_tickets.erase(it);
  }
  return c ? finish(wait_call(c)) : ETN_OK;
}

int etn_shard_ragged(int64_t batch, int64_t D, const int32_t* k_per_row, int n_shards, int64_t* cuts) {
  if (batch < 0 || D <= 0 || n_shards < 1) return ETN_ERR_SHAPE;
  if (!cuts || (batch > 0 && !k_per_row)) return ETN_ERR_NULL;
  std::vector<int64_t> csum, c;
  if (!build_csum(batch, D, k_per_row, csum)) return ETN_ERR_SHAPE;
  cut_by_kept(csum, n_shards, c);
  for (int g = 0; g <= n_shards; ++g) cuts[g] = c[(size_t)g];
  return ETN_OK;
}

int etn_raw_copy_seconds(const void* host_in, void* host_out, size_t bytes, int reps, int directions,
                         double* seconds) {
  if (!seconds || ((directions & 1) && !host_in) || ((directions & 2) && !host_out)) return ETN_ERR_NULL;
  if (bytes == 0 || reps < 1 || !(directions & 3)) return ETN_ERR_SHAPE;
  void *d_in = nullptr, *d_out = nullptr;
  cudaStream_t s0 = nullptr, s1 = nullptr;
  int rc = 0;
  auto ck = [&](cudaError_t e) { if (e != cudaSuccess && rc == 0) { rc = (int)e; cudaGetLastError(); } return e == cudaSuccess; };
  if ((directions & 1)) ck(cudaMalloc(&d_in, bytes));
  if ((directions & 2) && rc == 0) ck(cudaMalloc(&d_out, bytes));
  if (rc == 0) ck(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
  if (rc == 0) ck(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
  if (rc == 0) {
    auto pass = [&] {
      if (directions & 1) ck(cudaMemcpyAsync(d_in, host_in, bytes, cudaMemcpyHostToDevice, s0));
      if (directions & 2) ck(cudaMemcpyAsync(host_out, d_out, bytes, cudaMemcpyDeviceToHost, s1));
    };
    pass();
    ck(cudaStreamSynchronize(s0)); ck(cudaStreamSynchronize(s1));
    const auto t0 = Clock::now();
    for (int r = 0; r < reps && rc == 0; ++r) pass();
    ck(cudaStreamSynchronize(s0)); ck(cudaStreamSynchronize(s1));
    *seconds = std::chrono::duration<double>(Clock::now() - t0).count();
  }
  if (s0) cudaStreamDestroy(s0);
  if (s1) cudaStreamDestroy(s1);
  if (d_in) cudaFree(d_in);
  if (d_out) cudaFree(d_out);
  return rc;
}

}  // extern "C"
