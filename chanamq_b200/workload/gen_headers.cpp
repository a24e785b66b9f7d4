// Deterministic synthetic workload for the headers exchange (SplitMix64).
// 64 field names, 4096 values per name; a pool of "header templates" (8 distinct names, one value each) used
// Zipf(1.0); every message carries the 8 fields of one template.  Every binding has 8 argument slots plus
// x-match: `any` bindings (half of them) name 8 random (name, value) pairs; `all` bindings copy a sampled
// template and turn 0..6 of its 8 arguments into void (presence-only) arguments.  Three of the 64 names
// (4.7 % of the fields) carry 'I', 'T' and 'l' values instead of long strings, so type-tagged equality is
// exercised.  Tables are written as AMQP field-table CONTENTS (no leading length word), the form
// chq_bind_args / chq_route take.  Nothing here comes from the reference (it has no headers matcher,
// chana-mq-server/.../entity/ExchangeEntity.scala:210).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

namespace {

struct SplitMix64 {
  uint64_t s;
  explicit SplitMix64(uint64_t seed) : s(seed) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  double unit() { return (next() >> 11) * (1.0 / 9007199254740992.0); }
  uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
};

constexpr int kNames = 64, kValues = 4096, kFields = 8;

struct Field {
  uint8_t name;
  uint16_t value;
};

struct HeadersGen {
  uint64_t seed;
  std::vector<std::string> names;
  std::vector<Field> templates;  // n_templates * kFields
  std::vector<double> cdf, cdf_bind;
  uint32_t n_templates;

  HeadersGen(uint64_t sd, uint32_t nt) : seed(sd), n_templates(nt) {
    for (int i = 0; i < kNames; ++i) {
      char b[16];
      snprintf(b, sizeof b, "h%02d", i);
      names.emplace_back(std::string(b) + std::string(i % 5, 'x'));  // 3..7 bytes
    }
    templates.resize((size_t)nt * kFields);
    cdf.resize(nt);
    cdf_bind.resize(nt);
    double acc = 0, acc_b = 0;
    for (uint32_t t = 0; t < nt; ++t) {
      SplitMix64 r(seed * 0x100000001B3ull + 0x7E3779B97F4A7C15ull + t);
      bool used[kNames] = {};
      for (int f = 0; f < kFields; ++f) {
        uint32_t nm;
        do nm = r.below(kNames); while (used[nm]);
        used[nm] = true;
        templates[(size_t)t * kFields + f] = Field{(uint8_t)nm, (uint16_t)r.below(kValues)};
      }
      acc += 1.0 / (double)(t + 1);
      cdf[t] = acc;
      acc_b += 1.0 / std::sqrt((double)(t + 1));
      cdf_bind[t] = acc_b;
    }
    for (auto& v : cdf) v /= acc;
    for (auto& v : cdf_bind) v /= acc_b;
  }
  // messages use the templates Zipf(1.0); `all` bindings copy them Zipf(0.5) — calibrated so that a message
  // meets about as many `all` as `any` bindings (mean matches per message ~15 at 50 k bindings)
  uint32_t draw_template(SplitMix64& r, bool for_binding = false) const {
    const double u = r.unit();
    const std::vector<double>& c = for_binding ? cdf_bind : cdf;
    return (uint32_t)std::min<size_t>(std::lower_bound(c.begin(), c.end(), u) - c.begin(), c.size() - 1);
  }
  // one field: shortstr name, type tag, value; `as_void` writes a 'V' (presence-only) argument
  size_t put(uint8_t* dst, Field f, bool as_void) const {
    size_t n = 0;
    const std::string& nm = names[f.name];
    dst[n++] = (uint8_t)nm.size();
    memcpy(dst + n, nm.data(), nm.size());
    n += nm.size();
    if (as_void) { dst[n++] = 'V'; return n; }
    if (f.name == 61) {  // 'I': 32-bit integer, big-endian
      dst[n++] = 'I';
      const uint32_t v = f.value;
      for (int i = 3; i >= 0; --i) dst[n++] = (uint8_t)(v >> (8 * i));
    } else if (f.name == 62) {  // 'T': timestamp, 64-bit (a boolean would match half of all messages)
      dst[n++] = 'T';
      const uint64_t v = 1700000000ull + f.value;
      for (int i = 7; i >= 0; --i) dst[n++] = (uint8_t)(v >> (8 * i));
    } else if (f.name == 63) {  // 'l': 64-bit integer
      dst[n++] = 'l';
      const uint64_t v = (uint64_t)f.value * 0x100000001ull;
      for (int i = 7; i >= 0; --i) dst[n++] = (uint8_t)(v >> (8 * i));
    } else {  // 'S': long string of 4..12 lowercase letters, a function of (name, value)
      SplitMix64 r(seed ^ (((uint64_t)f.name << 32) | f.value) * 0xD6E8FEB86659FD93ull);
      const uint32_t len = 4 + r.below(9);
      dst[n++] = 'S';
      dst[n++] = 0; dst[n++] = 0; dst[n++] = 0; dst[n++] = (uint8_t)len;
      for (uint32_t i = 0; i < len; ++i) dst[n++] = (uint8_t)('a' + r.below(26));
    }
    return n;
  }
};

constexpr uint32_t kMaxField = 1 + 7 + 1 + 4 + 12;  // name + tag + longest value
constexpr uint32_t kXMatch = 1 + 7 + 1 + 4 + 3;     // "x-match" S "all"/"any"

}  // namespace

extern "C" {

void* chqgen_headers_new(uint64_t seed, uint32_t n_templates) { return new HeadersGen(seed, n_templates ? n_templates : 1); }
void chqgen_headers_free(void* g) { delete (HeadersGen*)g; }
uint32_t chqgen_headers_max_bytes(void) { return kFields * kMaxField + kXMatch; }

// n messages of stream `stream`: headers tables packed into blob, off[n+1].  Returns the bytes used.
uint64_t chqgen_headers_messages(void* gp, uint64_t stream, uint32_t n, uint8_t* blob, uint32_t* off) {
  const HeadersGen& g = *(HeadersGen*)gp;
  SplitMix64 r(g.seed * 0x9E3779B97F4A7C15ull + 0x5151 + stream * 0xD1B54A32D192ED03ull);
  uint64_t at = 0;
  for (uint32_t i = 0; i < n; ++i) {
    off[i] = (uint32_t)at;
    const Field* t = &g.templates[(size_t)g.draw_template(r) * kFields];
    for (int f = 0; f < kFields; ++f) at += g.put(blob + at, t[f], false);
  }
  off[n] = (uint32_t)at;
  return at;
}

// p bindings: argument tables (x-match first) packed into blob, off[p+1]; queue index (< n_queues) and
// durable flag per binding.  Returns the bytes used.
uint64_t chqgen_headers_bindings(void* gp, uint32_t p, uint32_t n_queues, uint8_t* blob, uint32_t* off, uint32_t* queue_idx,
                                 uint8_t* durable) {
  const HeadersGen& g = *(HeadersGen*)gp;
  SplitMix64 r(g.seed * 0xBF58476D1CE4E5B9ull + 0xB1D5);
  uint64_t at = 0;
  for (uint32_t i = 0; i < p; ++i) {
    off[i] = (uint32_t)at;
    const bool all = r.unit() < 0.5;
    static const char xm[] = "\x07x-matchS\0\0\0\x03";
    memcpy(blob + at, xm, 13);
    at += 13;
    memcpy(blob + at, all ? "all" : "any", 3);
    at += 3;
    if (all) {
      const Field* t = &g.templates[(size_t)g.draw_template(r, true) * kFields];
      const uint32_t n_void = r.below(7);  // 0..6 presence-only arguments
      bool as_void[kFields] = {};
      for (uint32_t v = 0; v < n_void;) {
        const uint32_t f = r.below(kFields);
        if (!as_void[f]) { as_void[f] = true; ++v; }
      }
      for (int f = 0; f < kFields; ++f) at += g.put(blob + at, t[f], as_void[f]);
    } else {
      for (int f = 0; f < kFields; ++f) at += g.put(blob + at, Field{(uint8_t)r.below(kNames), (uint16_t)r.below(kValues)}, false);
    }
    queue_idx[i] = n_queues >= p ? i : r.below(n_queues);
    durable[i] = r.unit() < 0.5 ? 1 : 0;
  }
  off[p] = (uint32_t)at;
  return at;
}

}  // extern "C"
