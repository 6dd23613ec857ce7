// etn_json.cpp — SURVEY.md §8(f) N1: the embedding arrays of an OpenAI-format response go
// from JSON text straight into a pinned [batch x k] float64 buffer and back out, without the
// map[string]any round trip the reference pays for (json.NewDecoder(...).Decode(&result),
// core/internal/api/handlers.go:2047-2050, and json.NewEncoder(w).Encode(payload),
// core/internal/api/helpers.go:11-15 — every number boxed into an interface value on the way
// in, every number formatted on the way out, although only the first `dims` of each vector
// survive).
//
// What is parsed and what is not: only the first `dims` numbers of the arrays that are
// actually cut are ever converted, and only in TRUNCATE_RENORM mode.  In TRUNCATE mode the
// kept number tokens are copied as text, so they are bit-identical by construction.
// Everything outside the `data[*].embedding` arrays passes through byte for byte.  (The
// reference re-encodes the whole map, which sorts object keys and drops insignificant
// white space: the output here is the same JSON value, not the same bytes.)
//
// Numbers: std::from_chars / std::to_chars are correctly rounded and shortest-round-trip, the
// guarantees Go's strconv.ParseFloat(s, 64) and strconv.AppendFloat(b, f, fmt, -1, 64) give;
// layout follows encoding/json's floatEncoder: 'f' form unless |x| < 1e-6 or |x| >= 1e21,
// then 'e' form with "e-0d" shortened to "e-d".
#include "../../include/embed_truncate_norm.h"
#include "etn_internal.h"

#if defined(__SSE2__)
#include <emmintrin.h>
#endif

#include <algorithm>
#include <atomic>
#include <charconv>
#include <cmath>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <pthread.h>
#include <string.h>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

// Behind the run of decimal digits that starts at p.  The fraction of an embedding component is
// 16-17 digits, most of a token: sixteen bytes are classified at once where the ISA allows (SSE2
// is part of every x86-64), so a token costs a couple of iterations instead of one per digit.
inline const char* skip_digits(const char* p, const char* e) {
#if defined(__SSE2__)
  const __m128i zero = _mm_set1_epi8('0'), nine = _mm_set1_epi8(9);
  while (e - p >= 16) {
    const __m128i t = _mm_sub_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i*>(p)), zero);
    const unsigned not_digit = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_min_epu8(t, nine), t)) ^ 0xFFFFu;
    if (not_digit) return p + __builtin_ctz(not_digit);
    p += 16;
  }
#endif
  while (p < e && *p >= '0' && *p <= '9') ++p;
  return p;
}

struct Scanner {
  const char* b;
  const char* e;
  const char* p;
  bool ok = true;
  bool escaped = false;    // the string last skipped held a backslash

  void ws() { while (p < e && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) ++p; }
  bool fail() { ok = false; return false; }

  // p at the opening quote; on return p is one past the closing quote.
  bool skip_string() {
    if (p >= e || *p != '"') return fail();
    ++p;
    escaped = false;
    while (p < e) {
      const char c = *p++;
      if (c == '"') return true;
      if (c == '\\') { escaped = true; if (p >= e) return fail(); ++p; }
      else if ((unsigned char)c < 0x20) return fail();
    }
    return fail();
  }
  // Same, for an object key, and tells whether the key equals `key` (plain ASCII) the way Go
  // compares it: after unescaping.  A key spelled with escapes (a backslash-u-0061 for the
  // second letter of "data", say) IS that key to the reference (handlers.go:2064, :2073 index
  // the decoded map), so it is one here too.  An escape Go's decoder rejects fails the scan.
  bool string_is(const char* key, bool* same) {
    const char* s = p + 1;
    if (!skip_string()) return false;
    const char* t = p - 1;                          // closing quote
    const size_t klen = strlen(key);
    if (!escaped) {
      *same = (size_t)(t - s) == klen && memcmp(s, key, klen) == 0;
      return true;
    }
    size_t m = 0;                                   // characters of `key` matched so far
    bool eq = true;
    while (s < t) {
      unsigned cp;
      if (*s != '\\') { cp = (unsigned char)*s++; }
      else {
        if (++s >= t) return fail();
        switch (*s++) {
          case '"': cp = '"'; break;   case '\\': cp = '\\'; break; case '/': cp = '/'; break;
          case 'b': cp = '\b'; break; case 'f': cp = '\f'; break;   case 'n': cp = '\n'; break;
          case 'r': cp = '\r'; break; case 't': cp = '\t'; break;
          case 'u': {
            if (t - s < 4) return fail();
            cp = 0;
            for (int i = 0; i < 4; ++i) {
              const char h = *s++;
              cp = cp * 16 + (h >= '0' && h <= '9' ? h - '0' : h >= 'a' && h <= 'f' ? h - 'a' + 10 : h >= 'A' && h <= 'F' ? h - 'A' + 10 : 1u << 20);
            }
            if (cp >= (1u << 20)) return fail();     // not four hex digits
            break;
          }
          default: return fail();
        }
      }
      // the keys looked for are ASCII: any other code point (surrogates included) is a mismatch
      if (eq && m < klen && cp == (unsigned char)key[m]) ++m; else eq = false;
    }
    *same = eq && m == klen;
    return true;
  }
  // A number the reference cannot hold fails its whole decode (json.Decode into map[string]any
  // parses EVERY number with strconv.ParseFloat(s, 64), handlers.go:2047-2050, and an overflow
  // is an error -> 502), so it fails this scan too, wherever in the document it stands; the
  // exact check only runs for tokens whose digit count plus exponent could reach 1e308.
  bool skip_number() {
    const char* s = p;
    if (p < e && *p == '-') ++p;
    if (p >= e) return fail();
    const char* i0 = p;
    if (*p == '0') ++p;
    else if (*p >= '1' && *p <= '9') { while (p < e && *p >= '0' && *p <= '9') ++p; }
    else return fail();
    const int64_t nint = p - i0;
    if (p < e && *p == '.') {
      ++p;
      if (p >= e || *p < '0' || *p > '9') return fail();
      p = skip_digits(p + 1, e);
    }
    int64_t ex = 0;
    if (p < e && (*p == 'e' || *p == 'E')) {
      ++p;
      const bool neg = p < e && *p == '-';
      if (p < e && (*p == '+' || *p == '-')) ++p;
      if (p >= e || *p < '0' || *p > '9') return fail();
      while (p < e && *p >= '0' && *p <= '9') { if (ex < 100000) ex = ex * 10 + (*p - '0'); ++p; }
      if (neg) ex = -ex;
    }
    if (nint + ex > 308 && overflows(s, p)) return fail();
    return p > s;
  }
  static bool overflows(const char* s, const char* t) {
    double d = 0;
    auto r = std::from_chars(s, t, d);
    if (r.ec != std::errc::result_out_of_range) return false;
    std::string z(s, t);                            // out of range both ways: only overflow is an error in Go
    return std::isinf(strtod(z.c_str(), nullptr));
  }
  bool skip_literal(const char* lit) {
    const size_t n = strlen(lit);
    if ((size_t)(e - p) < n || memcmp(p, lit, n) != 0) return fail();
    p += n;
    return true;
  }
  bool skip_value(int depth = 0) {
    if (depth > 512) return fail();
    ws();
    if (p >= e) return fail();
    switch (*p) {
      case '"': return skip_string();
      case '{': {
        ++p; ws();
        if (p < e && *p == '}') { ++p; return true; }
        for (;;) {
          ws();
          if (!skip_string()) return false;
          ws();
          if (p >= e || *p != ':') return fail();
          ++p;
          if (!skip_value(depth + 1)) return false;
          ws();
          if (p < e && *p == ',') { ++p; continue; }
          if (p < e && *p == '}') { ++p; return true; }
          return fail();
        }
      }
      case '[': {
        ++p; ws();
        if (p < e && *p == ']') { ++p; return true; }
        for (;;) {
          if (!skip_value(depth + 1)) return false;
          ws();
          if (p < e && *p == ',') { ++p; continue; }
          if (p < e && *p == ']') { ++p; return true; }
          return fail();
        }
      }
      case 't': return skip_literal("true");
      case 'f': return skip_literal("false");
      case 'n': return skip_literal("null");
      default: return skip_number();
    }
  }
};

// One array-valued "embedding" of a data item.
struct Span {
  int64_t open;        // offset of '['
  int64_t close;       // offset one past ']'
  int64_t count;       // elements
  int64_t cut;         // offset one past the `dims`-th element (valid when count > dims)
  bool numeric;        // the first min(count, dims) elements are all JSON numbers
  bool b64 = false;    // a base64 string of little-endian float32 (open/close = the quotes)
};

// ---- base64 (RFC 4648, standard alphabet, '=' padding): OpenAI's encoding_format "base64"
// carries each vector as little-endian float32 bytes.
const char kB64[] = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz0123456789+/";
struct B64Table {
  signed char v[256];
  B64Table() { memset(v, -1, sizeof v); for (int i = 0; i < 64; ++i) v[(unsigned char)kB64[i]] = (signed char)i; }
};
const B64Table kB64Dec;

// Decoded byte count of a padded base64 text, or -1 if it is not one.
int64_t b64_decoded_len(const char* t, size_t n) {
  if (n == 0 || n % 4 != 0) return n == 0 ? 0 : -1;
  size_t pad = 0;
  if (t[n - 1] == '=') { ++pad; if (t[n - 2] == '=') ++pad; }
  for (size_t i = 0; i < n - pad; ++i) if (kB64Dec.v[(unsigned char)t[i]] < 0) return -1;
  return (int64_t)(n / 4 * 3 - pad);
}
// Decodes the first nbytes bytes (nbytes <= decoded length).
void b64_decode_prefix(const char* t, size_t nbytes, unsigned char* dst) {
  size_t o = 0;
  for (size_t i = 0; o < nbytes; i += 4) {
    const unsigned a = (unsigned)kB64Dec.v[(unsigned char)t[i]], b = (unsigned)kB64Dec.v[(unsigned char)t[i + 1]];
    const unsigned c = t[i + 2] == '=' ? 0u : (unsigned)kB64Dec.v[(unsigned char)t[i + 2]];
    const unsigned d = t[i + 3] == '=' ? 0u : (unsigned)kB64Dec.v[(unsigned char)t[i + 3]];
    const unsigned w = (a << 18) | (b << 12) | (c << 6) | d;
    dst[o++] = (unsigned char)(w >> 16);
    if (o < nbytes) dst[o++] = (unsigned char)(w >> 8);
    if (o < nbytes) dst[o++] = (unsigned char)w;
  }
}
size_t b64_encoded_len(size_t nbytes) { return (nbytes + 2) / 3 * 4; }
void b64_encode(const unsigned char* src, size_t nbytes, char* dst) {
  size_t i = 0, o = 0;
  for (; i + 3 <= nbytes; i += 3) {
    const unsigned w = ((unsigned)src[i] << 16) | ((unsigned)src[i + 1] << 8) | src[i + 2];
    dst[o++] = kB64[w >> 18]; dst[o++] = kB64[(w >> 12) & 63]; dst[o++] = kB64[(w >> 6) & 63]; dst[o++] = kB64[w & 63];
  }
  if (i < nbytes) {
    const unsigned w = ((unsigned)src[i] << 16) | (i + 1 < nbytes ? (unsigned)src[i + 1] << 8 : 0u);
    dst[o++] = kB64[w >> 18]; dst[o++] = kB64[(w >> 12) & 63];
    dst[o++] = i + 1 < nbytes ? kB64[(w >> 6) & 63] : '=';
    dst[o++] = '=';
  }
}

// (A table-driven automaton for plain number arrays — one dependent table load per byte instead
// of the branches of skip_number — was tried and was SLOWER: 14.4 ms against 6.5 ms for the
// pre-scan of a C1-sized body on 16 cores; the branches predict well and the load chain does not
// pipeline.  tools/json_phases.cpp times the phases.)
// p at '[' of an embedding array.  Records the span; `dims` bounds the numeric check.
bool scan_embedding(Scanner& sc, int64_t dims, Span* out) {
  out->open = sc.p - sc.b;
  out->count = 0;
  out->cut = -1;
  out->numeric = true;
  ++sc.p; sc.ws();
  if (sc.p < sc.e && *sc.p == ']') { ++sc.p; out->close = sc.p - sc.b; return true; }
  for (;;) {
    sc.ws();
    if (sc.p >= sc.e) return sc.fail();
    const char c = *sc.p;
    const bool is_num = c == '-' || (c >= '0' && c <= '9');
    if (out->count < dims && !is_num) out->numeric = false;
    if (is_num) { if (!sc.skip_number()) return false; }
    else if (!sc.skip_value(2)) return false;
    ++out->count;
    if (out->count == dims) out->cut = sc.p - sc.b;
    sc.ws();
    if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
    if (sc.p < sc.e && *sc.p == ']') { ++sc.p; out->close = sc.p - sc.b; return true; }
    return sc.fail();
  }
}

// Vectors are independent, so number parsing and formatting spread over host threads
// (handlers run one goroutine per connection; a large response still deserves the cores).
// Host threads for the packer.  Creating and joining sixteen threads costs a few hundred
// microseconds, three times per response — a tenth of a C1-sized call — so the threads are kept:
// a small pool, created on first use and never torn down (its threads sleep on a condition
// variable).  One parallel region at a time; a caller that finds the pool taken (another handler
// is in its own region) spawns threads of its own as before.  Items are handed out one at a
// time from an atomic counter, so uneven items balance themselves.
std::atomic<bool> g_pool_lost{false};                // set in the child of a fork(): the pool's threads did not come along

class Pool {
 public:
  static Pool& get() { static Pool* p = new Pool(); return *p; }      // leaked on purpose: no join at exit
  size_t size() const { return workers_ + 1; }
  template <typename F>
  bool run(size_t n, size_t threads, F& f) {
    if (g_pool_lost.load(std::memory_order_relaxed)) return false;
    std::unique_lock<std::mutex> region(region_, std::try_to_lock);
    if (!region.owns_lock()) return false;
    const size_t helpers = std::min(workers_, threads > 0 ? threads - 1 : 0);
    {
      std::lock_guard<std::mutex> lk(mu_);
      fn_ = [&f](size_t i) { f(i); };
      n_ = n;
      next_.store(0, std::memory_order_relaxed);
      wanted_ = helpers;
      active_ = helpers;
      ++gen_;
    }
    if (helpers == workers_) cv_.notify_all(); else for (size_t i = 0; i < helpers; ++i) cv_.notify_one();
    for (size_t i; (i = next_.fetch_add(1, std::memory_order_relaxed)) < n;) f(i);
    std::unique_lock<std::mutex> lk(mu_);
    done_.wait(lk, [&] { return active_ == 0; });
    fn_ = nullptr;
    return true;
  }

 private:
  Pool() {
    const size_t hw = std::max<size_t>(1, std::min<size_t>(std::thread::hardware_concurrency(), 16));
    workers_ = hw - 1;
    pthread_atfork(nullptr, nullptr, [] { g_pool_lost.store(true); });
    for (size_t t = 0; t < workers_; ++t) std::thread([this] { loop(); }).detach();
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      std::function<void(size_t)> fn;
      size_t n;
      {
        std::unique_lock<std::mutex> lk(mu_);
        cv_.wait(lk, [&] { return gen_ != seen && wanted_ > 0; });
        seen = gen_;
        --wanted_;
        fn = fn_;
        n = n_;
      }
      for (size_t i; (i = next_.fetch_add(1, std::memory_order_relaxed)) < n;) fn(i);
      std::lock_guard<std::mutex> lk(mu_);
      if (--active_ == 0) done_.notify_all();
    }
  }
  std::mutex region_, mu_;
  std::condition_variable cv_, done_;
  std::function<void(size_t)> fn_;
  std::atomic<size_t> next_{0};
  size_t n_ = 0, workers_ = 0, wanted_ = 0, active_ = 0;
  uint64_t gen_ = 0;
};

// f(i) for i in [0, n) on up to min(16, cores, n / grain) threads.
template <typename F>
void parallel_for(size_t n, size_t grain, F f) {
  size_t nt = std::min<size_t>(std::min<size_t>(std::thread::hardware_concurrency(), 16), n / std::max<size_t>(grain, 1));
  if (nt <= 1) { for (size_t i = 0; i < n; ++i) f(i); return; }
  if (Pool::get().run(n, nt, f)) return;
  std::vector<std::thread> th;                       // the pool is taken: threads of this call's own
  for (size_t t = 0; t < nt; ++t)
    th.emplace_back([=] { for (size_t i = t; i < n; i += nt) f(i); });
  for (auto& x : th) x.join();
}

// Speculative parallel pre-scan.  Nearly all bytes of a response sit inside the embedding
// arrays, and validating them number by number is the scan's whole cost, so host threads
// each take a slice of the text, look for `"embedding"` followed by `: [` and scan the
// array that starts there.  The sequential skeleton scan below stays the only judge of what
// is really a data item's embedding; when it arrives at an array it has a finished span for,
// it jumps to the closing bracket instead of walking the numbers again.  A hit inside a
// string or at the wrong depth is simply never asked for.
struct PreScan {
  std::vector<Span> spans;                         // sorted by .open
  const Span* find(int64_t open) const {
    auto it = std::lower_bound(spans.begin(), spans.end(), open, [](const Span& s, int64_t o) { return s.open < o; });
    return (it != spans.end() && it->open == open) ? &*it : nullptr;
  }
};

bool scan_embedding(Scanner& sc, int64_t dims, Span* out);

void prescan(const char* json, size_t len, int64_t dims, PreScan* ps) {
  const size_t kMin = 4u << 20;                     // not worth threads below 4 MiB
  if (len < kMin) return;
  const size_t nthreads = std::min<size_t>(std::min<size_t>(std::thread::hardware_concurrency(), 16), len / (kMin / 4));
  if (nthreads <= 1) return;
  const size_t nt = nthreads * 4;                   // four slices per thread, handed out as threads come free
  std::vector<std::vector<Span>> found(nt);
  const size_t chunk = (len + nt - 1) / nt;
  static const char key[] = "\"embedding\"";
  const size_t klen = sizeof key - 1;
  parallel_for(nt, 4, [&](size_t t) {
    const char* p = json + t * chunk;
    const char* stop = json + std::min(len, (t + 1) * chunk);
    const char* end = json + len;
    while (p < stop) {
      const char* q = (const char*)memmem(p, (size_t)(end - p) < (size_t)(stop - p) + klen ? (size_t)(end - p) : (size_t)(stop - p) + klen,
                                         key, klen);
      if (!q || q >= stop) break;
      const char* r = q + klen;
      while (r < end && (*r == ' ' || *r == '\t' || *r == '\n' || *r == '\r')) ++r;
      if (r < end && *r == ':') {
        ++r;
        while (r < end && (*r == ' ' || *r == '\t' || *r == '\n' || *r == '\r')) ++r;
        if (r < end && *r == '[') {
          Scanner sc{json, end, r};
          Span s;
          if (scan_embedding(sc, dims, &s)) { found[t].push_back(s); p = sc.p; continue; }
        }
      }
      p = q + klen;
    }
  });
  for (auto& v : found) ps->spans.insert(ps->spans.end(), v.begin(), v.end());
  std::sort(ps->spans.begin(), ps->spans.end(), [](const Span& a, const Span& b) { return a.open < b.open; });
}

// Finds the embedding arrays of result["data"][*]["embedding"] (handlers.go:2064-2073).
// Returns 0, ETN_ERR_JSON for malformed input, ETN_ERR_JSON_DUP when "data" or an item's
// "embedding" occurs twice (Go keeps the last one and drops the other on re-encode; a splice
// cannot mimic that, the caller should take the reference path).
int find_embeddings(const char* json, size_t len, int64_t dims, bool b64, std::vector<Span>* spans) {
  PreScan ps;
  prescan(json, len, dims, &ps);
  Scanner sc{json, json + len, json};
  sc.ws();
  if (sc.p >= sc.e || *sc.p != '{') {            // not an object: Go's Decode(&map) would fail earlier
    if (!sc.skip_value()) return ETN_ERR_JSON;
    return ETN_OK;
  }
  ++sc.p; sc.ws();
  int data_seen = 0;
  if (sc.p < sc.e && *sc.p == '}') return ETN_OK;
  for (;;) {
    sc.ws();
    bool is_data = false;
    if (!sc.string_is("data", &is_data)) return ETN_ERR_JSON;
    sc.ws();
    if (sc.p >= sc.e || *sc.p != ':') return ETN_ERR_JSON;
    ++sc.p; sc.ws();
    if (is_data && ++data_seen > 1) return ETN_ERR_JSON_DUP;
    if (is_data && sc.p < sc.e && *sc.p == '[') {
      ++sc.p; sc.ws();
      if (sc.p < sc.e && *sc.p == ']') { ++sc.p; }
      else for (;;) {
        sc.ws();
        if (sc.p < sc.e && *sc.p == '{') {       // a data item: look for its "embedding"
          ++sc.p; sc.ws();
          int emb_seen = 0;
          if (sc.p < sc.e && *sc.p == '}') { ++sc.p; }
          else for (;;) {
            sc.ws();
            bool is_emb = false;
            if (!sc.string_is("embedding", &is_emb)) return ETN_ERR_JSON;
            sc.ws();
            if (sc.p >= sc.e || *sc.p != ':') return ETN_ERR_JSON;
            ++sc.p; sc.ws();
            if (is_emb && ++emb_seen > 1) return ETN_ERR_JSON_DUP;
            if (is_emb && sc.p < sc.e && *sc.p == '[') {
              Span s;
              if (const Span* hit = ps.find(sc.p - sc.b)) { s = *hit; sc.p = sc.b + s.close; }
              else if (!scan_embedding(sc, dims, &s)) return ETN_ERR_JSON;
              spans->push_back(s);
            } else if (is_emb && b64 && sc.p < sc.e && *sc.p == '"') {
              Span s;
              s.open = sc.p - sc.b;
              if (!sc.skip_string()) return ETN_ERR_JSON;
              s.close = sc.p - sc.b;
              const int64_t nb = b64_decoded_len(json + s.open + 1, (size_t)(s.close - s.open - 2));
              if (nb > 0 && nb % 4 == 0) {          // float32 vector; anything else stays untouched
                s.count = nb / 4; s.cut = -1; s.numeric = true; s.b64 = true;
                spans->push_back(s);
              }
            } else if (!sc.skip_value(2)) return ETN_ERR_JSON;
            sc.ws();
            if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
            if (sc.p < sc.e && *sc.p == '}') { ++sc.p; break; }
            return ETN_ERR_JSON;
          }
        } else if (!sc.skip_value(1)) return ETN_ERR_JSON;   // non-object item: skipped (:2069-2072)
        sc.ws();
        if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
        if (sc.p < sc.e && *sc.p == ']') { ++sc.p; break; }
        return ETN_ERR_JSON;
      }
    } else if (!sc.skip_value(1)) return ETN_ERR_JSON;
    sc.ws();
    if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
    if (sc.p < sc.e && *sc.p == '}') { ++sc.p; break; }
    return ETN_ERR_JSON;
  }
  sc.ws();
  return sc.p == sc.e ? ETN_OK : ETN_ERR_JSON;
}

// encoding/json floatEncoder for float64.  Returns the number of bytes written (<= 32).
size_t format_go_float(double f, char* out) {
  char* o = out;
  if (f == 0) {                                   // "0" or "-0"
    if (std::signbit(f)) *o++ = '-';
    *o++ = '0';
    return (size_t)(o - out);
  }
  char sci[40];
  auto r = std::to_chars(sci, sci + sizeof sci, f, std::chars_format::scientific);   // d[.ddd]e[+-]XX
  const char* s = sci;
  if (*s == '-') { *o++ = '-'; ++s; }
  const char* ep = s;
  while (ep < r.ptr && *ep != 'e') ++ep;
  char digits[24];
  int nd = 0;
  for (const char* q = s; q < ep; ++q) if (*q != '.') digits[nd++] = *q;
  int exp10 = 0;
  std::from_chars(ep + 1 + (ep[1] == '+'), r.ptr, exp10);
  const double a = std::fabs(f);
  if (a < 1e-6 || a >= 1e21) {                    // 'e' form: d[.ddd]e[+-]dd, then e-0d -> e-d
    *o++ = digits[0];
    if (nd > 1) { *o++ = '.'; for (int i = 1; i < nd; ++i) *o++ = digits[i]; }
    *o++ = 'e';
    *o++ = exp10 < 0 ? '-' : '+';
    const int ax = exp10 < 0 ? -exp10 : exp10;
    if (ax < 10) { if (exp10 >= 0) *o++ = '0'; *o++ = (char)('0' + ax); }   // "e+07" stays, "e-07" -> "e-7"
    else { auto t = std::to_chars(o, o + 8, ax); o = t.ptr; }
    return (size_t)(o - out);
  }
  // 'f' form, shortest digits, no exponent
  if (exp10 < 0) {
    *o++ = '0'; *o++ = '.';
    for (int i = 0; i < -exp10 - 1; ++i) *o++ = '0';
    for (int i = 0; i < nd; ++i) *o++ = digits[i];
  } else {
    for (int i = 0; i <= exp10; ++i) *o++ = i < nd ? digits[i] : '0';
    if (nd > exp10 + 1) { *o++ = '.'; for (int i = exp10 + 1; i < nd; ++i) *o++ = digits[i]; }
  }
  return (size_t)(o - out);
}

// ETN_JSON_GPU=0 keeps the packer's number work on the host cores (read at use, like config.Getenv).
bool gpu_text_enabled() {
  const char* v = getenv("ETN_JSON_GPU");
  return !(v && v[0] == '0');
}

// Reusable pinned staging for the rewrite (one per thread would be better still; handlers
// call concurrently, so a mutex guards the shared one and overflow callers allocate their own).
struct PinPool {
  std::mutex mu;
  void* in = nullptr;
  size_t cap = 0;
} g_pin;

}  // namespace

namespace etn_rt {
void release_json_staging() {                       // etn_shutdown: nothing of the library's stays pinned
  std::lock_guard<std::mutex> lk(g_pin.mu);
  if (g_pin.in) etn_free_pinned(g_pin.in);
  g_pin.in = nullptr;
  g_pin.cap = 0;
}
}  // namespace etn_rt

extern "C" {

int64_t etn_json_format_array(const double* v, int64_t n, char* buf, size_t cap) {
  if (n < 0 || (!v && n > 0) || !buf) return ETN_ERR_NULL;
  size_t o = 0;
  if (cap < 2) return ETN_ERR_SHAPE;
  buf[o++] = '[';
  for (int64_t i = 0; i < n; ++i) {
    if (o + 34 > cap) return ETN_ERR_SHAPE;
    if (i) buf[o++] = ',';
    if (!std::isfinite(v[i])) return ETN_ERR_JSON;   // Go's encoder refuses NaN/Inf (helpers.go:14 drops the body)
    o += format_go_float(v[i], buf + o);
  }
  if (o + 1 > cap) return ETN_ERR_SHAPE;
  buf[o++] = ']';
  return (int64_t)o;
}

int64_t etn_json_parse_array(const char* text, size_t len, double* dst, int64_t max_n) {
  if (!text || (!dst && max_n > 0)) return ETN_ERR_NULL;
  Scanner sc{text, text + len, text};
  sc.ws();
  if (sc.p >= sc.e || *sc.p != '[') return ETN_ERR_JSON;
  ++sc.p; sc.ws();
  int64_t n = 0;
  if (sc.p < sc.e && *sc.p == ']') return 0;
  for (;;) {
    sc.ws();
    const char* s = sc.p;
    if (!sc.skip_number()) return ETN_ERR_JSON;
    if (n < max_n) {
      auto r = std::from_chars(s, sc.p, dst[n]);
      if (r.ec == std::errc::result_out_of_range) {
        // overflow never gets here (skip_number refused it: Go's "cannot unmarshal number ... into
        // float64"); underflow is no error in strconv.ParseFloat: the nearest value, 0 or a subnormal
        std::string z(s, sc.p);
        dst[n] = strtod(z.c_str(), nullptr);
      } else if (r.ec != std::errc() || r.ptr != sc.p) return ETN_ERR_JSON;
    }
    ++n;
    sc.ws();
    if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
    if (sc.p < sc.e && *sc.p == ']') return n;
    return ETN_ERR_JSON;
  }
}

int64_t etn_json_prompt_tokens(const char* json, size_t len) {
  if (!json) return ETN_ERR_NULL;
  Scanner sc{json, json + len, json};
  sc.ws();
  if (sc.p >= sc.e || *sc.p != '{') return 0;
  ++sc.p; sc.ws();
  if (sc.p < sc.e && *sc.p == '}') return 0;
  int64_t tokens = 0;                              // Go keeps the last duplicate; so does this loop
  for (;;) {
    sc.ws();
    bool is_usage = false;
    if (!sc.string_is("usage", &is_usage)) return ETN_ERR_JSON;
    sc.ws();
    if (sc.p >= sc.e || *sc.p != ':') return ETN_ERR_JSON;
    ++sc.p; sc.ws();
    if (is_usage) tokens = 0;
    if (is_usage && sc.p < sc.e && *sc.p == '{') {
      ++sc.p; sc.ws();
      if (sc.p < sc.e && *sc.p == '}') { ++sc.p; }
      else for (;;) {
        sc.ws();
        bool is_pt = false;
        if (!sc.string_is("prompt_tokens", &is_pt)) return ETN_ERR_JSON;
        sc.ws();
        if (sc.p >= sc.e || *sc.p != ':') return ETN_ERR_JSON;
        ++sc.p; sc.ws();
        const char* v0 = sc.p;
        const bool num = sc.p < sc.e && (*sc.p == '-' || (*sc.p >= '0' && *sc.p <= '9'));
        if (!sc.skip_value(2)) return ETN_ERR_JSON;
        if (is_pt) {
          tokens = 0;
          if (num) {
            double d = 0;
            auto r = std::from_chars(v0, sc.p, d);
            if (r.ec == std::errc() && std::fabs(d) < 9.2e18) tokens = (int64_t)d;   // int(pt): truncation toward zero
          }
        }
        sc.ws();
        if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
        if (sc.p < sc.e && *sc.p == '}') { ++sc.p; break; }
        return ETN_ERR_JSON;
      }
    } else if (!sc.skip_value(1)) return ETN_ERR_JSON;
    sc.ws();
    if (sc.p < sc.e && *sc.p == ',') { ++sc.p; continue; }
    if (sc.p < sc.e && *sc.p == '}') break;
    return ETN_ERR_JSON;
  }
  return tokens;
}

int64_t etn_json_truncate_response(const char* json, size_t len, int64_t dims, int mode, char* out,
                                   size_t out_cap, int n_devices, int64_t* stats) {
  const bool b64 = (mode & ETN_JSON_BASE64) != 0;
  mode &= ~ETN_JSON_BASE64;
  if (mode != ETN_TRUNCATE && mode != ETN_TRUNCATE_RENORM) return ETN_ERR_ENUM;
  if (!json || !out) return ETN_ERR_NULL;
  if (stats) { stats[0] = stats[1] = stats[2] = 0; }
  std::vector<Span> spans;
  if (dims > 0) {
    const int rc = find_embeddings(json, len, dims, b64, &spans);
    if (rc != ETN_OK) return rc;
  }
  // vectors that go through the GPU: longer than dims, numeric prefix, renorm requested
  std::vector<size_t> gpu64, gpu32;
  if (mode == ETN_TRUNCATE_RENORM)
    for (size_t i = 0; i < spans.size(); ++i)
      if (spans[i].count > dims && spans[i].numeric) (spans[i].b64 ? gpu32 : gpu64).push_back(i);
  double* vals = nullptr;
  float* vals32 = nullptr;
  void* own = nullptr;
  // float64 arrays: first choice is the device text path (etn_json_gpu.cu) — the kept TEXT goes
  // to the GPU, numbers are parsed, renormalised and formatted there, text comes back; if it is
  // not available (workspace busy, shapes too large for shared memory, ETN_JSON_GPU=0) the host
  // cores parse and format as before.
  etn_rt::JsonGpuResult gres;
  struct GpuRelease { etn_rt::JsonGpuResult* r; ~GpuRelease() { etn_rt::json_gpu_release(r); } } gpu_release{&gres};
  bool text_on_gpu = false;
  if (!gpu64.empty() && gpu_text_enabled()) {
    std::vector<int64_t> tb(gpu64.size()), te(gpu64.size());
    for (size_t g = 0; g < gpu64.size(); ++g) { tb[g] = spans[gpu64[g]].open + 1; te[g] = spans[gpu64[g]].cut; }
    const int rc = etn_rt::json_renorm_gpu(json, tb.data(), te.data(), gpu64.size(), dims, &gres);
    if (rc == 0) text_on_gpu = true;
    else if (rc != etn_rt::kJsonGpuUnavailable && rc != ETN_ERR_NOT_INIT) return rc < 0 ? rc : -1000 - rc;
  }
  std::vector<size_t> host64;                        // float64 vectors parsed and formatted on the host
  if (!text_on_gpu) host64 = gpu64;
  std::unique_lock<std::mutex> pin_lock(g_pin.mu, std::defer_lock);
  if (!host64.empty() || !gpu32.empty()) {
    const size_t need64 = host64.size() * (size_t)dims * sizeof(double);
    const size_t need = need64 + gpu32.size() * (size_t)dims * sizeof(float);
    if (pin_lock.try_lock()) {
      if (g_pin.cap < need) {
        etn_free_pinned(g_pin.in);
        g_pin.in = etn_alloc_pinned(need);
        g_pin.cap = g_pin.in ? need : 0;
      }
      vals = (double*)g_pin.in;
      if (!vals) pin_lock.unlock();
    }
    if (!vals) { own = etn_alloc_pinned(need); vals = (double*)own; }
    if (!vals) return -1000 - 2;                   // cudaErrorMemoryAllocation
    vals32 = (float*)((char*)vals + need64);
    std::atomic<int64_t> parse_err{0};
    parallel_for(host64.size(), 8, [&](size_t g) {
      const Span& s = spans[host64[g]];
      const int64_t n = etn_json_parse_array(json + s.open, (size_t)(s.close - s.open), vals + g * dims, dims);
      if (n < 0) parse_err.store(n);
    });
    if (parse_err.load() < 0) { if (own) etn_free_pinned(own); return parse_err.load(); }
    parallel_for(gpu32.size(), 64, [&](size_t g) {  // zero-parse path: bytes straight into pinned float32
      b64_decode_prefix(json + spans[gpu32[g]].open + 1, (size_t)dims * 4, (unsigned char*)(vals32 + g * dims));
    });
    // prefix-only buffers: D == k == dims, renormalised in place through the host path
    int rc = 0;
    if (!host64.empty())
      rc = embed_truncate_norm_host((int64_t)host64.size(), dims, dims, ETN_F64, ETN_TRUNCATE_RENORM, vals, 0, vals, n_devices);
    if (rc == 0 && !gpu32.empty())
      rc = embed_truncate_norm_host((int64_t)gpu32.size(), dims, dims, ETN_F32, ETN_TRUNCATE_RENORM, vals32, 0, vals32, n_devices);
    if (rc != 0) { if (own) etn_free_pinned(own); return rc < 0 ? rc : -1000 - rc; }
  }
  // format the renormalised float64 vectors on all cores, then splice sequentially
  const size_t fstride = (size_t)dims * 34 + 2;
  std::vector<char> ftext(host64.size() * fstride);
  std::vector<int64_t> flen(host64.size());
  parallel_for(host64.size(), 8, [&](size_t g) {
    flen[g] = etn_json_format_array(vals + g * dims, dims, ftext.data() + g * fstride, fstride);
  });
  // splice: first lay the output out as a list of pieces (what goes where), then let the host
  // threads copy them — the pieces are independent and a C1-sized answer is 20 MB of them
  struct Piece { const char* p; size_t n, o; int kind; };   // kind 0: copy, 1: base64 of n raw bytes at p, 2: base64 prefix of a base64 text at p
  std::vector<Piece> pieces;
  pieces.reserve(spans.size() * 3 + 1);
  size_t o = 0, src = 0, g64 = 0, g32 = 0;
  auto put = [&](const char* p, size_t n) -> bool {
    if (o + n > out_cap) return false;
    if (n) pieces.push_back({p, n, o, 0});
    o += n;
    return true;
  };
  bool fits = true;
  for (size_t i = 0; i < spans.size() && fits; ++i) {
    const Span& s = spans[i];
    if (s.count <= dims) continue;                 // left whole, never padded (handlers.go:2074)
    fits = put(json + src, (size_t)s.open - src);
    if (!fits) break;
    if (s.b64) {
      const size_t nbytes = (size_t)dims * 4, enc = b64_encoded_len(nbytes);
      if (o + enc + 2 > out_cap) { fits = false; break; }
      put("\"", 1);
      if (g32 < gpu32.size() && gpu32[g32] == i) {
        pieces.push_back({(const char*)(vals32 + g32 * dims), nbytes, o, 1});
        ++g32;
        if (stats) ++stats[0];
      } else {
        pieces.push_back({json + s.open + 1, nbytes, o, 2});
        if (stats) ++stats[1];
      }
      o += enc;
      put("\"", 1);
    } else if (g64 < gpu64.size() && gpu64[g64] == i) {
      const int64_t w = text_on_gpu ? (int64_t)gres.len[g64] : flen[g64];
      if (w < 0) { if (own) etn_free_pinned(own); return w; }     // NaN/Inf: Go's encoder would refuse too
      fits = put(text_on_gpu ? gres.text + g64 * gres.stride : ftext.data() + g64 * fstride, (size_t)w);
      if (!fits) break;
      ++g64;
      if (stats) ++stats[0];
    } else {                                       // the reference reslice, on the text: tokens kept verbatim
      fits = put(json + s.open, (size_t)(s.cut - s.open)) && put("]", 1);
      if (stats) ++stats[1];
    }
    src = (size_t)s.close;
  }
  if (fits) fits = put(json + src, len - src);
  if (fits) {
    auto run = [&](const Piece& q) {
      if (q.kind == 0) { memcpy(out + q.o, q.p, q.n); return; }
      if (q.kind == 1) { b64_encode((const unsigned char*)q.p, q.n, out + q.o); return; }
      std::vector<unsigned char> raw(q.n);
      b64_decode_prefix(q.p, q.n, raw.data());
      b64_encode(raw.data(), q.n, out + q.o);
    };
    // a few large groups of pieces per thread rather than one piece at a time
    const size_t per = 64;
    const size_t groups = (pieces.size() + per - 1) / per;
    parallel_for(groups, o >= (4u << 20) ? 2 : groups + 1, [&](size_t g) {
      for (size_t i = g * per; i < std::min(pieces.size(), (g + 1) * per); ++i) run(pieces[i]);
    });
  }
  if (own) etn_free_pinned(own);
  if (!fits) return ETN_ERR_SHAPE;
  if (stats) stats[2] = (int64_t)spans.size();
  if (mode == ETN_TRUNCATE_RENORM) {                 // renormalisation was asked for and these vectors only got the cut
    int64_t only_cut = 0;
    for (size_t i = 0; i < spans.size(); ++i) only_cut += spans[i].count > dims && !spans[i].numeric;
    if (only_cut) etn_note_fallback(only_cut);
  }
  return (int64_t)o;
}

}  // extern "C"
