// dma_matrix.cu — raw host<->device copy matrix of one box (development tool, not product).
//
// Question it answers ("next" item 2): when embed_truncate_norm_host runs on
// N GPUs the whole box moves only ~52-66 GB/s each way.  Is that the copy path of the box
// (then it is the ceiling of any end-to-end number) or the library's pipeline?
//
// For N = 1, 2, 4, 8 GPUs at once it times pinned cudaMemcpyAsync H2D-only, D2H-only and both
// directions together (two streams per GPU), `bytes` per copy, from ONE process with a thread
// per device and from N processes, with the H2D source in ordinary pinned memory and in
// write-combined pinned memory.  It also reports where the pinned pages live (NUMA node via
// move_pages), the host's own memcpy bandwidth for 1..all cores (the DMA engines read and
// write the same DRAM), and the PCIe link each GPU trained to.
//
//   make -C tools dma_matrix && gpurun --gpus 8 -- ./tools/dma_matrix > gpurun_out/dma_matrix.txt
//
// CUDA is initialised only in forked children, never in the parent, so every group starts clean.
#include <cuda_runtime.h>
#include <pthread.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/syscall.h>
#include <sys/wait.h>
#include <unistd.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); _exit(3); } } while (0)

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

struct Shared {
  pthread_barrier_t bar;
};

static Shared* make_shared(int parties) {
  void* p = mmap(nullptr, sizeof(Shared), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
  if (p == MAP_FAILED) { perror("mmap"); exit(2); }
  Shared* s = (Shared*)p;
  pthread_barrierattr_t a;
  pthread_barrierattr_init(&a);
  pthread_barrierattr_setpshared(&a, PTHREAD_PROCESS_SHARED);
  pthread_barrier_init(&s->bar, &a, parties);
  return s;
}

static void numa_report(const char* label, int g, void* buf, size_t bytes) {
  const int n = 64;
  void* pages[n];
  int status[n];
  const size_t step = bytes / n & ~(size_t)4095;
  for (int i = 0; i < n; ++i) pages[i] = (char*)buf + i * step;
  long rc = syscall(SYS_move_pages, 0, (unsigned long)n, pages, nullptr, status, 0);
  if (rc != 0) { printf("#   gpu %d %s: move_pages query failed\n", g, label); return; }
  int hist[16] = {0}, other = 0;
  for (int i = 0; i < n; ++i) { if (status[i] >= 0 && status[i] < 16) ++hist[status[i]]; else ++other; }
  std::string s;
  for (int i = 0; i < 16; ++i) if (hist[i]) s += " node" + std::to_string(i) + ":" + std::to_string(hist[i]);
  if (other) s += " unresolved:" + std::to_string(other);
  printf("#   gpu %d %s pages (64 sampled):%s\n", g, label, s.c_str());
}

struct Opt { size_t bytes = 1ull << 30; int reps = 3; bool pin = false; };

// One agent = one GPU.  Every agent of the group walks the same test list; agents with
// g >= N sit a test out but still meet the others at the barriers.
static void agent(int g, int G, Shared* sh, const Opt& o, const char* mode) {
  if (o.pin) {
    cpu_set_t set; CPU_ZERO(&set);
    const int ncpu = (int)sysconf(_SC_NPROCESSORS_ONLN);
    CPU_SET((g * (ncpu / (G > 0 ? G : 1))) % ncpu, &set);
    sched_setaffinity(0, sizeof set, &set);
  }
  CK(cudaSetDevice(g));
  void *h_in, *h_wc, *h_out, *d_in, *d_out;
  CK(cudaHostAlloc(&h_in, o.bytes, cudaHostAllocPortable));
  CK(cudaHostAlloc(&h_wc, o.bytes, cudaHostAllocPortable | cudaHostAllocWriteCombined));
  CK(cudaHostAlloc(&h_out, o.bytes, cudaHostAllocPortable));
  memset(h_in, 1, o.bytes); memset(h_wc, 1, o.bytes); memset(h_out, 0, o.bytes);
  CK(cudaMalloc(&d_in, o.bytes)); CK(cudaMalloc(&d_out, o.bytes));
  CK(cudaMemset(d_out, 2, o.bytes));
  cudaStream_t s0, s1;
  CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
  CK(cudaDeviceSynchronize());
  pthread_barrier_wait(&sh->bar);
  if (g == 0) printf("# %s mode, %d agent(s), %zu MiB per copy, %d reps%s\n", mode, G, o.bytes >> 20, o.reps, o.pin ? ", one core per agent" : "");
  for (int q = 0; q < G; ++q) {                      // page placement, printed in turn
    if (q == g) { numa_report("pinned in ", g, h_in, o.bytes); numa_report("pinned out", g, h_out, o.bytes); fflush(stdout); }
    pthread_barrier_wait(&sh->bar);
  }
  struct Test { const char* name; bool h2d, d2h, wc; };
  const Test tests[] = {{"H2D", true, false, false}, {"D2H", false, true, false}, {"duplex", true, true, false},
                        {"H2D (write-combined source)", true, false, true}, {"duplex (write-combined source)", true, true, true}};
  for (const Test& t : tests) {
    for (int N = 1; N <= G; N *= 2) {
      const bool active = g < N;
      auto pass = [&]() {
        if (!active) return;
        if (t.h2d) CK(cudaMemcpyAsync(d_in, t.wc ? h_wc : h_in, o.bytes, cudaMemcpyHostToDevice, s0));
        if (t.d2h) CK(cudaMemcpyAsync(h_out, d_out, o.bytes, cudaMemcpyDeviceToHost, s1));
      };
      pass();
      if (active) { CK(cudaStreamSynchronize(s0)); CK(cudaStreamSynchronize(s1)); }
      pthread_barrier_wait(&sh->bar);
      const double t0 = now_s();
      for (int r = 0; r < o.reps; ++r) pass();
      if (active) { CK(cudaStreamSynchronize(s0)); CK(cudaStreamSynchronize(s1)); }
      const double mine = now_s() - t0;
      pthread_barrier_wait(&sh->bar);
      const double all = now_s() - t0;
      if (g == 0) {
        const double per_dir = (double)N * o.reps * o.bytes / all / 1e9;
        printf("%-8s %-32s N=%d  %8.1f GB/s per direction in total  %7.1f GB/s per GPU   (gpu0 alone finished at %.1f GB/s)\n",
               mode, t.name, N, per_dir, per_dir / N, (double)o.reps * o.bytes / mine / 1e9);
        fflush(stdout);
      }
    }
  }
  CK(cudaFreeHost(h_in)); CK(cudaFreeHost(h_wc)); CK(cudaFreeHost(h_out));
  CK(cudaFree(d_in)); CK(cudaFree(d_out));
}

static int device_count_in_child() {
  pid_t p = fork();
  if (p == 0) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) n = 0; _exit(n); }
  int st = 0; waitpid(p, &st, 0);
  return WIFEXITED(st) ? WEXITSTATUS(st) : 0;
}

static void run_group(int G, const Opt& o, bool procs) {
  Shared* sh = make_shared(G);
  fflush(stdout);
  if (procs) {
    std::vector<pid_t> kids;
    for (int g = 0; g < G; ++g) {
      pid_t p = fork();
      if (p == 0) { agent(g, G, sh, o, "procs"); fflush(stdout); _exit(0); }
      kids.push_back(p);
    }
    for (pid_t p : kids) { int st; waitpid(p, &st, 0); if (!WIFEXITED(st) || WEXITSTATUS(st)) printf("# child failed (status %d)\n", st); }
  } else {
    pid_t p = fork();
    if (p == 0) {
      std::vector<std::thread> th;
      for (int g = 0; g < G; ++g) th.emplace_back(agent, g, G, sh, std::cref(o), "threads");
      for (auto& t : th) t.join();
      fflush(stdout);
      _exit(0);
    }
    int st; waitpid(p, &st, 0);
    if (!WIFEXITED(st) || WEXITSTATUS(st)) printf("# thread group failed (status %d)\n", st);
  }
  munmap(sh, sizeof(Shared));
}

static void host_memcpy_bw() {
  const int ncpu = (int)sysconf(_SC_NPROCESSORS_ONLN);
  const size_t bytes = 256u << 20;
  printf("# host memcpy bandwidth (read + write counted once each way: GB/s copied), %d cpus online\n", ncpu);
  for (int T = 1; T <= ncpu; T *= 2) {
    std::vector<char*> src(T), dst(T);
    for (int i = 0; i < T; ++i) { src[i] = (char*)malloc(bytes); dst[i] = (char*)malloc(bytes); memset(src[i], 1, bytes); memset(dst[i], 2, bytes); }
    const int reps = 4;
    const double t0 = now_s();
    std::vector<std::thread> th;
    for (int i = 0; i < T; ++i) th.emplace_back([&, i] { for (int r = 0; r < reps; ++r) memcpy(dst[i], src[i], bytes); });
    for (auto& t : th) t.join();
    const double el = now_s() - t0;
    printf("host memcpy %2d thread(s): %7.1f GB/s copied\n", T, (double)T * reps * bytes / el / 1e9);
    for (int i = 0; i < T; ++i) { free(src[i]); free(dst[i]); }
  }
  fflush(stdout);
}

static void shell(const char* cmd) {
  printf("$ %s\n", cmd); fflush(stdout);
  int rc = system(cmd);
  if (rc != 0) printf("# (exit %d)\n", rc);
  fflush(stdout);
}

int main(int argc, char** argv) {
  Opt o;
  bool info = true, threads = true, procs = true;
  for (int i = 1; i < argc; ++i) {
    if (!strcmp(argv[i], "--bytes") && i + 1 < argc) o.bytes = strtoull(argv[++i], nullptr, 0);
    else if (!strcmp(argv[i], "--reps") && i + 1 < argc) o.reps = atoi(argv[++i]);
    else if (!strcmp(argv[i], "--pin")) o.pin = true;
    else if (!strcmp(argv[i], "--no-info")) info = false;
    else if (!strcmp(argv[i], "--procs-only")) threads = false;
    else if (!strcmp(argv[i], "--threads-only")) procs = false;
  }
  const int visible = device_count_in_child();
  int G = 1;
  while (G * 2 <= visible) G *= 2;
  printf("# dma_matrix: %d visible GPU(s), using %d\n", visible, G);
  if (visible < 1) return 1;
  if (info) {
    shell("lscpu | grep -E 'Model name|Socket|Core|Thread|NUMA|^CPU\\(s\\)|Hypervisor|L3'");
    shell("(numactl --hardware 2>/dev/null || (echo 'numactl not installed; /sys/devices/system/node:'; ls /sys/devices/system/node | grep node; cat /sys/devices/system/node/node*/meminfo | grep -E 'MemTotal|MemFree'))");
    shell("nvidia-smi --query-gpu=index,pci.bus_id,pcie.link.gen.current,pcie.link.gen.max,pcie.link.width.current,pcie.link.width.max --format=csv");
    shell("nvidia-smi topo -m 2>/dev/null | head -24");
    shell("grep -E 'MemTotal|MemAvailable|Hugepagesize|AnonHugePages' /proc/meminfo");
    host_memcpy_bw();
  }
  if (threads) run_group(G, o, false);
  if (procs) run_group(G, o, true);
  return 0;
}
