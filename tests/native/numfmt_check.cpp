// CPU check of llm-mcp_b200/csrc/etn_numfmt.cuh (the per-number code of the packer's GPU path)
// against the library's host formatter and parser, which stand on std::to_chars /
// std::from_chars: format_go_float64 must give the same bytes as etn_json_format_array, and
// parse_float64, whenever it answers at all, the same bits as etn_json_parse_array.
// Built with g++ and run by tests/test_numfmt.py; argv[1] = millions of random cases.
#include <cinttypes>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <vector>

#include "../../include/embed_truncate_norm.h"
#include "../../llm-mcp_b200/csrc/etn_numfmt.cuh"

static const etn_num::Pow10Row kTab[] = {
#include "../../llm-mcp_b200/csrc/etn_pow10_table.inc"
};

static long g_fail = 0;
static void fail(const char* what, double v, const char* a, const char* b) {
  if (g_fail++ < 20) fprintf(stderr, "FAIL %s: %.17g (%016" PRIx64 "): ours '%s' want '%s'\n", what, v, etn_num::double_to_bits(v), a, b);
}

static void check_format(double v) {
  if (!std::isfinite(v)) return;
  char ours[64], want[96];
  const int n = etn_num::format_go_float64(v, kTab, ours);
  ours[n] = 0;
  const int64_t w = etn_json_format_array(&v, 1, want, sizeof want);
  if (w < 2) { fail("format(lib)", v, ours, "?"); return; }
  want[w - 1] = 0;                                    // drop ']'
  if (strcmp(ours, want + 1) != 0) fail("format", v, ours, want + 1);
}

static long g_fallback = 0, g_parsed = 0;
static void check_parse(const std::string& tok) {
  double ours = 0, want = 0;
  const std::string arr = "[" + tok + "]";
  const int64_t n = etn_json_parse_array(arr.c_str(), arr.size(), &want, 1);
  const bool ok = etn_num::parse_float64(tok.c_str(), (int)tok.size(), kTab, ours);
  if (n != 1) {                                       // the host refuses it (overflow): we must not answer
    if (ok && std::isfinite(ours)) fail("parse accepted what the host refuses", ours, tok.c_str(), "refused");
    return;
  }
  if (!ok) { ++g_fallback; return; }
  ++g_parsed;
  if (etn_num::double_to_bits(ours) != etn_num::double_to_bits(want)) {
    char a[40], b[40];
    snprintf(a, sizeof a, "%.17g", ours); snprintf(b, sizeof b, "%.17g", want);
    fail(("parse " + tok).c_str(), want, a, b);
  }
}

int main(int argc, char** argv) {
  const long millions = argc > 1 ? atol(argv[1]) : 2;
  std::mt19937_64 rng(20261107);
  // ---- fixed cases: Go's own spellings, boundaries of the 'e' form, powers of ten, subnormals
  const double fixed[] = {0.0, -0.0, 1.0, -1.0, 0.1, 0.5, 1e-6, 9.999999999999999e-7, 1e-7, 1e21, 999999999999999934463.0, 1e20, 123456789.0,
                          1e22, 1e23, 5e-324, 2.2250738585072014e-308, 2.2250738585072009e-308, 1.7976931348623157e308, 0.30000000000000004,
                          100.0, 1e15, 1e16, 1e17, 9007199254740993.0, 4.35, 2.5e-5, 123456.789e3, 1.0 / 3.0, 2.0 / 3.0, 0.000001, 0.0000011};
  for (double v : fixed) { check_format(v); check_format(-v); }
  for (int e = -330; e <= 310; ++e) {
    const double v = std::pow(10.0, e);
    check_format(v); check_format(std::nextafter(v, 0.0)); check_format(std::nextafter(v, INFINITY));
    char t[32];
    snprintf(t, sizeof t, "1e%d", e); check_parse(t);
    snprintf(t, sizeof t, "-9.999999999999999e%d", e); check_parse(t);
  }
  for (int e = 0; e < 2047; ++e)                       // every binade: smallest, largest, a middle one
    for (uint64_t m : {0ull, 1ull, 0x000FFFFFFFFFFFFFull, 0x0008000000000000ull, 0x000FFFFFFFFFFFFEull})
      check_format(etn_num::bits_to_double(((uint64_t)e << 52) | m));
  const char* toks[] = {"0", "-0", "0.0", "1", "-1", "10", "1.5", "0.1", "3", "4", "1E5", "1e+5", "1e-5", "0e10", "0.000", "123456789012345678",
                        "1234567890123456789", "12345678901234567890", "0.1234567890123456789", "0.12345678901234567891", "1e999", "-1e999", "1e-999",
                        "4.9e-324", "2e-324", "2.2250738585072011e-308", "1.7976931348623157e308", "1.7976931348623159e308", "9007199254740993",
                        "9007199254740992.5", "0.30000000000000004", "100000000000000000000000", "0.00000000000000000000000000000001",
                        "1.00000000000000011102230246251565404236316680908203125", "1.00000000000000011102230246251565404236316680908203124"};
  for (const char* t : toks) check_parse(t);
  // ---- random
  const long N = millions * 1000000L;
  std::uniform_real_distribution<double> uni(-1.0, 1.0);
  std::normal_distribution<double> nrm(0.0, 0.04);
  char buf[64];
  for (long i = 0; i < N; ++i) {
    double v;
    switch (i & 3) {
      case 0: v = etn_num::bits_to_double(rng()); break;                       // any bit pattern
      case 1: v = nrm(rng); break;                                             // embedding-like components
      case 2: v = uni(rng) * std::pow(10.0, (int)(rng() % 40) - 20); break;
      default: v = (double)(int64_t)(rng() >> (rng() % 64)); break;            // integers of every size
    }
    check_format(v);
    if (!std::isfinite(v)) continue;
    // parse: the shortest spelling (what providers send), a 17-digit one, a cut one, and Go-style
    const int n = etn_num::format_go_float64(v, kTab, buf);
    check_parse(std::string(buf, (size_t)n));
    if ((i & 7) == 0) {
      snprintf(buf, sizeof buf, "%.17g", v); if (!strchr(buf, 'n') && !strchr(buf, 'i')) check_parse(buf);
      snprintf(buf, sizeof buf, "%.*e", (int)(rng() % 19), v); check_parse(buf);
      snprintf(buf, sizeof buf, "%.*f", (int)(rng() % 12), std::fmod(v, 1e9)); check_parse(buf);
    }
  }
  printf("numfmt check: %ld random values, parse answered %ld, fell back %ld, failures %ld\n", N, g_parsed, g_fallback, g_fail);
  return g_fail ? 1 : 0;
}
