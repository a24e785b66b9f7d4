// Deterministic synthetic workloads for the routing hot path (SplitMix64, fixed seeds), shared by
// the tests, bench.py and the CPU baseline.  Shape: SURVEY.md §8(d) "T(seed, P, N, mode)" —
// a region/tenant/device/metric hierarchy: key depth uniform in 3..6, level vocabularies
// V = [32, 1024, 16384, 64, 64, 64] sampled Zipf(s) with s = [1.0, 0.5, 0.0, 0.5, 0.5, 0.5],
// words of 3-8 lowercase letters.  Binding patterns are made from independently sampled keys:
// 70 % exact, 25 % one '*', 5 % two '*'; with `hash_tail` a further 10 % get their suffix after
// word j >= 2 replaced by a single '#' (their first two words are kept literal).  Nothing here comes from the reference (it ships no
// generator; its perf specs drive an external tool, chana-mq-test/perf/test.sh:2).
#include <algorithm>
#include <cmath>
#include <cstdint>
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

constexpr int kLevels = 6;
const uint32_t kVocab[kLevels] = {32, 1024, 16384, 64, 64, 64};
const double kZipf[kLevels] = {1.0, 0.5, 0.0, 0.5, 0.5, 0.5};

struct TopicGen {
  uint64_t seed;
  std::vector<std::string> words[kLevels];
  std::vector<double> cdf[kLevels];

  explicit TopicGen(uint64_t sd) : seed(sd) {
    for (int l = 0; l < kLevels; ++l) {
      words[l].resize(kVocab[l]);
      cdf[l].resize(kVocab[l]);
      double acc = 0;
      for (uint32_t i = 0; i < kVocab[l]; ++i) {
        SplitMix64 r(seed * 0x100000001B3ull + ((uint64_t)l << 32) + i + 1);
        int len = 3 + (int)r.below(6);
        std::string w;
        for (int c = 0; c < len; ++c) w.push_back((char)('a' + r.below(26)));
        words[l][i] = w;
        acc += 1.0 / std::pow((double)(i + 1), kZipf[l]);
        cdf[l][i] = acc;
      }
      for (auto& v : cdf[l]) v /= acc;
    }
  }
  uint32_t draw(int level, SplitMix64& r) const {
    double u = r.unit();
    auto& c = cdf[level];
    return (uint32_t)std::min<size_t>(std::lower_bound(c.begin(), c.end(), u) - c.begin(), c.size() - 1);
  }
  // samples one key as word indices; returns depth
  int sample(SplitMix64& r, uint32_t* idx) const {
    int depth = 3 + (int)r.below(4);
    for (int l = 0; l < depth; ++l) idx[l] = draw(l, r);
    return depth;
  }
};

size_t emit(const TopicGen& g, const uint32_t* idx, const int8_t* kind, int depth, uint8_t* dst) {
  // kind[l]: 0 literal, 1 '*', 2 '#'
  size_t n = 0;
  for (int l = 0; l < depth; ++l) {
    if (l) dst[n++] = '.';
    if (kind && kind[l] == 1) dst[n++] = '*';
    else if (kind && kind[l] == 2) dst[n++] = '#';
    else {
      const std::string& w = g.words[l][idx[l]];
      memcpy(dst + n, w.data(), w.size());
      n += w.size();
    }
  }
  return n;
}

}  // namespace

extern "C" {

void* chqgen_topic_new(uint64_t seed) { return new TopicGen(seed); }
void chqgen_topic_free(void* g) { delete (TopicGen*)g; }
// upper bound of bytes one key or pattern can take: 6 words * 8 letters + 5 dots
uint32_t chqgen_topic_max_len() { return kLevels * 8 + (kLevels - 1); }

// N routing keys of stream `stream` (batches are independent streams).  Returns blob bytes.
uint64_t chqgen_topic_keys(void* gp, uint64_t stream, uint32_t n, uint8_t* blob, uint32_t* off) {
  auto& g = *(TopicGen*)gp;
  SplitMix64 r(g.seed ^ (0xA5A5A5A5ull + stream * 0xD6E8FEB86659FD93ull));
  uint64_t pos = 0;
  uint32_t idx[kLevels];
  for (uint32_t i = 0; i < n; ++i) {
    off[i] = (uint32_t)pos;
    int d = g.sample(r, idx);
    pos += emit(g, idx, nullptr, d, blob + pos);
  }
  off[n] = (uint32_t)pos;
  return pos;
}

// P binding patterns.  queue[i] = i % n_queues (queue names are "q%07u" of that index, so the
// canonical id order equals the index order), durable[i] Bernoulli(0.5) per QUEUE.
uint64_t chqgen_topic_patterns(void* gp, uint32_t p, uint32_t n_queues, int hash_tail, uint8_t* blob,
                               uint32_t* off, uint32_t* queue, uint8_t* durable) {
  auto& g = *(TopicGen*)gp;
  SplitMix64 r(g.seed ^ 0x5EED0BADC0FFEEull);
  uint64_t pos = 0;
  uint32_t idx[kLevels];
  int8_t kind[kLevels];
  for (uint32_t i = 0; i < p; ++i) {
    off[i] = (uint32_t)pos;
    int d = g.sample(r, idx);
    memset(kind, 0, sizeof kind);
    double u = r.unit();
    if (u >= 0.70) {
      int a = (int)r.below(d);
      kind[a] = 1;
      if (u >= 0.95) {
        int b = (int)r.below(d - 1);
        if (b >= a) ++b;
        kind[b] = 1;
      }
    }
    if (hash_tail && r.unit() < 0.10) {
      int j = 2 + (int)r.below(d - 2);  // keep words 0..j-1, then '#'
      kind[j] = 2;
      d = j + 1;
      // the first two words of a '#' pattern stay literal: "*.*.#" would match every key and
      // 1 M bindings of that mixture would emit thousands of queues per message (the degenerate
      // shape SURVEY.md §8(d) warns about)
      kind[0] = kind[1] = 0;
    }
    pos += emit(g, idx, kind, d, blob + pos);
    queue[i] = i % n_queues;
    SplitMix64 qd(g.seed ^ (0xD00Dull + queue[i]));
    durable[i] = (uint8_t)(qd.next() & 1);
  }
  off[p] = (uint32_t)pos;
  return pos;
}

}  // extern "C"
