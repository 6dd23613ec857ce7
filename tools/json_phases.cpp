// Where the packer's host time goes for a C1-sized response in TRUNCATE mode (development tool):
// builds etn_json.cpp directly (device calls stubbed, as in tests/native/json_fuzz.cpp) and times
// the span search (parallel pre-scan + sequential skeleton) apart from the splice.
//   g++ -O2 -std=c++17 -pthread -o tools/json_phases tools/json_phases.cpp && ./tools/json_phases
#define ETN_NO_CUDA
#include "../llm-mcp_b200/csrc/etn_json.cpp"

#include <chrono>
#include <cstdio>
#include <random>

namespace etn_rt {
int json_renorm_gpu(const char*, const int64_t*, const int64_t*, size_t, int64_t, JsonGpuResult*) { return kJsonGpuUnavailable; }
void json_gpu_release(JsonGpuResult*) {}
}  // namespace etn_rt
extern "C" {
void* etn_alloc_pinned(size_t n) { return malloc(n); }
void etn_free_pinned(void* p) { free(p); }
void etn_note_fallback(int64_t) {}
int embed_truncate_norm_host(int64_t, int64_t, int64_t, int, int, const void*, int64_t, void*, int) { return 0; }
}

static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main() {
  std::mt19937_64 rng(1);
  std::normal_distribution<double> nrm(0.0, 1.0 / 64);
  std::string body = "{\"object\": \"list\", \"data\": [";
  char buf[64];
  for (int v = 0; v < 1024; ++v) {
    body += v ? ", {\"object\": \"embedding\", \"index\": " : "{\"object\": \"embedding\", \"index\": ";
    body += std::to_string(v) + ", \"embedding\": [";
    for (int i = 0; i < 4096; ++i) {
      auto r = std::to_chars(buf, buf + sizeof buf, nrm(rng));
      if (i) body += ", ";
      body.append(buf, r.ptr);
    }
    body += "]}";
  }
  body += "], \"usage\": {\"prompt_tokens\": 9}}";
  std::vector<char> out(body.size() + 2);
  printf("body %.1f MB, %u hardware threads\n", body.size() / 1e6, std::thread::hardware_concurrency());
  for (int rep = 0; rep < 4; ++rep) {
    double t0 = now();
    PreScan ps;
    prescan(body.data(), body.size(), 1024, &ps);
    double t1 = now();
    std::vector<Span> spans;
    int rc = find_embeddings(body.data(), body.size(), 1024, false, &spans);
    double t2 = now();
    int64_t stats[3];
    int64_t n = etn_json_truncate_response(body.data(), body.size(), 1024, ETN_TRUNCATE, out.data(), out.size(), 1, stats);
    double t3 = now();
    printf("prescan alone %.2f ms (%zu spans) | find_embeddings (prescan + skeleton) %.2f ms rc=%d | whole call %.2f ms -> %ld bytes  => splice ~%.2f ms\n",
           (t1 - t0) * 1e3, ps.spans.size(), (t2 - t1) * 1e3, rc, (t3 - t2) * 1e3, (long)n, (t3 - t2 - (t2 - t1)) * 1e3);
  }
  return 0;
}
