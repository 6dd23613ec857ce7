// Sanitizer harness for the JSON packer (host code that parses network input).
// Built by tests/test_json_sanitize.py with -fsanitize=address,undefined: the packer's source
// is compiled in directly, the CUDA-side entry points it calls are stubbed (TRUNCATE mode
// never reaches them), and a few thousand byte-level mutations of provider-like responses are
// pushed through.  Any out-of-bounds access or UB aborts the process.
#define ETN_NO_CUDA
#include "../../llm-mcp_b200/csrc/etn_json.cpp"

#include <cstdio>
#include <cstdlib>
#include <random>

namespace etn_rt {   // the device text path is not part of this build: "use the host path"
int json_renorm_gpu(const char*, const int64_t*, const int64_t*, size_t, int64_t, JsonGpuResult*) { return kJsonGpuUnavailable; }
void json_gpu_release(JsonGpuResult*) {}
}  // namespace etn_rt

extern "C" {
void* etn_alloc_pinned(size_t n) { return malloc(n); }
void etn_free_pinned(void* p) { free(p); }
void etn_note_fallback(int64_t) {}
int embed_truncate_norm_host(int64_t batch, int64_t, int64_t k, int dtype, int, const void* in, int64_t, void* out, int) {
  if (in != out) memcpy(out, in, (size_t)(batch * k) * (dtype == ETN_F32 ? 4 : 8));
  return 0;   // identity "renorm": enough to drive the parse -> format path
}
}

int main() {
  std::mt19937_64 rng(12345);
  std::vector<std::string> seeds = {
      R"({"object":"list","data":[{"object":"embedding","index":0,"embedding":[0.1,-2.5e-3,3,4e10,5.5,6,7,8,9]},{"index":1,"embedding":[1,2]}],"model":"m","usage":{"prompt_tokens":3}})",
      R"({"data":[{"embedding":"AAAAPwAAAD8AAAA/AAAAPwAAAD8AAAA/"},{"embedding":null},"x",{"embedding":[1,"a",[2,3],{"b":4},5]}]})",
      R"( { "data" : [ { "embedding" : [ 1.0 , 2.0 ,	3.0 ] , "k\"]}" : "v\\" } ] , "z" : [ [ ] , { } ] } )",
      // keys spelled with escapes (compared after unescaping, as Go does) and numbers at the edge of float64
      R"({"d\u0061ta":[{"emb\u0065dding":[1e308,2E+307,3,4,5e-324],"\ud83d\ude00\/\b":1}],"us\u0061ge":{"prompt\u005ftokens":7e0},"big":17976931348623157e292})",
  };
  std::vector<char> out(1 << 16);
  int64_t stats[3];
  size_t ok = 0, rejected = 0;
  for (int iter = 0; iter < 6000; ++iter) {
    std::string s = seeds[iter % seeds.size()];
    const int nmut = (int)(rng() % 4);
    for (int m = 0; m < nmut && !s.empty(); ++m) {
      const size_t pos = rng() % s.size();
      switch (rng() % 4) {
        case 0: s[pos] = (char)(rng() % 256); break;
        case 1: s.erase(pos, 1 + rng() % 3); break;
        case 2: { static const char alphabet[] = "[]{}\",:0123456789.eE-+ \\"; s.insert(pos, 1, alphabet[rng() % (sizeof alphabet - 1)]); break; }
        default: s.resize(pos); break;
      }
    }
    { char* in = (char*)malloc(s.size() ? s.size() : 1); memcpy(in, s.data(), s.size()); (void)etn_json_prompt_tokens(in, s.size()); free(in); }
    for (int mode : {ETN_TRUNCATE, ETN_TRUNCATE_RENORM, ETN_TRUNCATE | ETN_JSON_BASE64, ETN_TRUNCATE_RENORM | ETN_JSON_BASE64}) {
      for (int64_t dims : {1, 2, 4}) {
        // exact-size heap copies so that ASan sees any read past the end of the input
        char* in = (char*)malloc(s.size() ? s.size() : 1);
        memcpy(in, s.data(), s.size());
        const int64_t n = etn_json_truncate_response(in, s.size(), dims, mode, out.data(), out.size(), 0, stats);
        free(in);
        if (n >= 0) ++ok; else if (n == ETN_ERR_JSON || n == ETN_ERR_JSON_DUP) ++rejected; else { printf("unexpected code %ld\n", (long)n); return 1; }
        if (n > (int64_t)out.size()) return 2;
      }
    }
  }
  // number formatting over raw bit patterns
  char buf[64];
  for (int i = 0; i < 200000; ++i) {
    uint64_t b = rng();
    double d; memcpy(&d, &b, 8);
    if (!std::isfinite(d)) continue;
    const int64_t n = etn_json_format_array(&d, 1, buf, sizeof buf);
    if (n < 3 || n > 40) return 3;
    double back = 0;
    if (etn_json_parse_array(buf, (size_t)n, &back, 1) != 1 || memcmp(&back, &d, 8) != 0) return 4;
  }
  printf("ok=%zu rejected=%zu\n", ok, rejected);
  return 0;
}
