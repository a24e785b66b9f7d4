#include "host_table.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <unordered_map>

#include "../../include/chanamq_gpu.h"

namespace chq {

void java_split(std::string_view s, std::vector<std::string_view>& out) {
  out.clear();
  size_t start = 0;
  bool dot = false;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '.') {
      out.push_back(s.substr(start, i - start));
      start = i + 1;
      dot = true;
    }
  }
  if (!dot) {
    out.push_back(s);
    return;
  }
  out.push_back(s.substr(start));
  while (!out.empty() && out.back().empty()) out.pop_back();
}

int HostTable::check_key(std::string_view key) const {
  if (key.size() > (size_t)kMaxKeyBytes) return CHQ_E_KEY_TOO_LONG;
  if (type_ == CHQ_TOPIC || type_ == CHQ_HEADERS) {
    std::vector<std::string_view> w;
    java_split(key, w);
    if (w.empty()) return CHQ_E_ZERO_WORDS;  // the reference throws here (QueueMatcher.scala:420)
  }
  return CHQ_OK;
}

bool valid_field_table(std::string_view t);  // host_headers.cpp

int HostTable::bind(std::string_view key, std::string_view queue, bool durable, std::string_view args) {
  int rc = check_key(key);
  if (rc != CHQ_OK) return rc;
  if (!args.empty() && !valid_field_table(args)) return CHQ_E_ARG;
  const bool fresh = live_.emplace(std::string(key), std::string(queue), durable, std::string(args)).second;
  if (fresh && inc_.valid) pending_.push_back(Op{true, std::string(key), std::string(queue), durable});
  return CHQ_OK;
}

int HostTable::unbind(std::string_view key, std::string_view queue, bool durable) {
  int rc = check_key(key);
  if (rc != CHQ_OK) return rc;
  // queueUnbind(queue, routingKey) ignores the arguments (ExchangeEntity.scala:370-382): every variant goes
  auto it = live_.lower_bound(std::make_tuple(std::string(key), std::string(queue), durable, std::string()));
  bool any = false;
  while (it != live_.end() && std::get<0>(*it) == key && std::get<1>(*it) == queue && std::get<2>(*it) == durable) {
    it = live_.erase(it);
    any = true;
  }
  if (any && inc_.valid) pending_.push_back(Op{false, std::string(key), std::string(queue), durable});
  return CHQ_OK;
}

int HostTable::unbind_queue(std::string_view queue) {
  for (auto it = live_.begin(); it != live_.end();) {
    if (std::get<1>(*it) == queue) {
      if (inc_.valid) pending_.push_back(Op{false, std::get<0>(*it), std::get<1>(*it), std::get<2>(*it)});
      it = live_.erase(it);
    } else {
      ++it;
    }
  }
  return CHQ_OK;
}

uint32_t bytes_hash(uint32_t seed, std::string_view w);

namespace {

uint32_t chunk32(std::string_view w, size_t at) {  // four bytes from `at`, little-endian, zero padded
  uint32_t v = 0;
  for (size_t i = 0; i < 4 && at + i < w.size(); ++i) v |= (uint32_t)(unsigned char)w[at + i] << (8 * i);
  return v;
}
uint32_t word_hash(std::string_view w) { return bytes_hash(kWordBasis, w); }
uint64_t first8(std::string_view w) {
  uint64_t f = 0;
  for (size_t i = 0; i < w.size() && i < 8; ++i) f |= (uint64_t)(unsigned char)w[i] << (8 * i);
  return f;
}
uint32_t pow2_at_least(uint64_t x) {
  uint32_t p = 1;
  while (p < x) p <<= 1;
  return p;
}

struct GroupBuild {
  uint32_t shape_key;  // index into the unsorted shape list
  std::vector<uint32_t> lits, rem;
  std::vector<uint32_t> queues;
  bool durable = false;
};

}  // namespace

// Hash of a byte string folded eight bytes at a time (the device side: hash_bytes in probe.cuh).
uint32_t bytes_hash(uint32_t seed, std::string_view w) {
  uint32_t h = seed;
  size_t at = 0;
  do {
    h = word_chunk(h, chunk32(w, at), chunk32(w, at + 4));
    at += 8;
  } while (at < w.size());
  return word_fin(h, (uint32_t)w.size());
}

void HostTable::compile(CompiledTable& out) const {
  // CHQ_COMPILE_TRACE=1 prints where the time of a commit goes (stderr)
  const bool trace = getenv("CHQ_COMPILE_TRACE") != nullptr;
  auto t_prev = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    if (!trace) return;
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "compile: %-28s %7.3f s\n", what, std::chrono::duration<double>(now - t_prev).count());
    t_prev = now;
  };
  out = CompiledTable();
  inc_ = Inc();        // rebuilt at the end (topic / direct tables); the changes queued for it are in this compile
  pending_.clear();
  out.n_bindings = live_.size();
  out.exact_rng.assign(3 * (kLenCap + 1), 0);

  // queue ids: rank of the name among the names that hold a binding, ascending by bytes
  {
    std::vector<std::string_view> q;
    q.reserve(live_.size() + declared_.size());
    for (auto& t : live_) q.push_back(std::get<1>(t));
    for (auto& d : declared_) q.push_back(d);
    std::sort(q.begin(), q.end());
    q.erase(std::unique(q.begin(), q.end()), q.end());
    out.queue_names.assign(q.begin(), q.end());
  }
  std::unordered_map<std::string_view, uint32_t> qid;
  qid.reserve(out.queue_names.size() * 2);
  for (uint32_t i = 0; i < out.queue_names.size(); ++i) qid.emplace(out.queue_names[i], i);

  lap("queue names and ids");
  if (type_ == CHQ_HEADERS && mode_ == CHQ_MODE_SPEC) {  // x-match on the arguments; the routing key plays no part
    compile_headers(out);
    return;
  }
  const bool split = (type_ == CHQ_TOPIC || type_ == CHQ_HEADERS);
  const bool spec = (mode_ == CHQ_MODE_SPEC) && split;

  if (type_ == CHQ_FANOUT) {  // FanoutMatcher.lookup: every subscriber (QueueMatcher.scala:63-65)
    out.dict.assign(1, DictSlot{});
    out.wdict.assign(1, WordSlot{});
    out.wdict_off.assign(1, 0);
    out.tags.assign(kBucketSlots, 0);
    out.slot_group.assign(kBucketSlots, kNone);
    if (!live_.empty()) {
      Group g{};
      g.next = kNone;
      bool dur = false;
      for (auto& t : live_) {
        out.qlist.push_back(qid.at(std::get<1>(t)));
        dur = dur || std::get<2>(t);
      }
      std::sort(out.qlist.begin(), out.qlist.end());
      out.qlist.erase(std::unique(out.qlist.begin(), out.qlist.end()), out.qlist.end());
      g.list_off = 0;
      g.list_len = (uint32_t)out.qlist.size();
      out.qlist.insert(out.qlist.end(), 4, kNone);  // room for the 16-byte units of a bulk copy
      g.flags = dur ? kAnyDurable : 0;
      out.groups.push_back(g);
      out.fanout_group = 0;
    }
    out.n_heads = out.groups.size();
    return;
  }

  // literal words are interned as they are met (ids 1..; the bindings are walked in key order, so a rebuild of
  // the same set gives the same table)
  std::unordered_map<std::string_view, uint32_t> words;
  std::vector<std::string_view> w;
  auto is_star = [](std::string_view s) { return s.size() == 1 && s[0] == '*'; };
  auto is_hash = [&](std::string_view s) { return spec && s.size() == 1 && s[0] == '#'; };
  auto word_id = [&](std::string_view x) {
    auto it = words.find(x);
    if (it == words.end()) it = words.emplace(x, (uint32_t)words.size() + 1).first;
    return it->second;
  };

  // reduce every pattern to (shape, literal tuple, remainder) and group equal ones
  std::map<std::tuple<uint8_t, uint8_t, uint32_t>, uint32_t> shape_ids;  // (kind, len, mask)
  std::vector<std::tuple<uint8_t, uint8_t, uint32_t>> shape_list;
  std::unordered_map<std::vector<uint32_t>, uint32_t, VecHash> group_ids;  // encoded pattern -> group
  group_ids.reserve(live_.size() * 2);
  std::vector<GroupBuild> gb;
  std::vector<uint32_t> pat, enc;
  bool any_complex = false;
  uint32_t max_len = 1, last_shape = kNone;
  std::tuple<uint8_t, uint8_t, uint32_t> last_sk{};
  for (auto& t : live_) {
    std::string_view key = std::get<0>(t);
    if (split) java_split(key, w);
    else { w.clear(); w.push_back(key); }   // DirectMatcher: whole-key equality (QueueMatcher.scala:45-47)
    pat.clear();
    for (auto& x : w) {
      if (split && is_star(x)) pat.push_back(kStar);
      else if (is_hash(x)) {
        if (pat.empty() || pat.back() != kHash) pat.push_back(kHash);  // "#.#" == "#"
      } else pat.push_back(word_id(x));
    }
    out.pattern_words += pat.size();
    size_t first_hash = std::find(pat.begin(), pat.end(), kHash) - pat.begin();
    size_t prefix = first_hash;  // == pat.size() when there is no '#'
    size_t ip = std::min<size_t>(prefix, kMaxIndexedWords);
    uint8_t kind;
    if (first_hash == pat.size() && pat.size() <= (size_t)kMaxIndexedWords) kind = kExact;
    else if (first_hash < pat.size() && prefix <= (size_t)kMaxIndexedWords && first_hash + 1 == pat.size()) kind = kTail;
    else kind = kComplex;
    uint32_t mask = 0;
    for (size_t p = 0; p < ip; ++p)
      if (pat[p] != kStar) mask |= 1u << p;
    if (kind == kComplex) any_complex = true;
    auto sk = std::make_tuple(kind, (uint8_t)ip, mask);
    if (sk != last_sk || last_shape == kNone) {   // consecutive patterns (sorted by key) mostly share their shape
      auto sit = shape_ids.find(sk);
      if (sit == shape_ids.end()) {
        sit = shape_ids.emplace(sk, (uint32_t)shape_list.size()).first;
        shape_list.push_back(sk);
      }
      last_sk = sk;
      last_shape = sit->second;
    }
    max_len = std::max<uint32_t>(max_len, (uint32_t)ip);
    enc.clear();
    enc.push_back(last_shape);
    for (size_t p = 0; p < ip; ++p)
      if (pat[p] != kStar) enc.push_back(pat[p]);
    const size_t n_lits = enc.size() - 1;
    enc.push_back(kNone);
    if (kind == kComplex) enc.insert(enc.end(), pat.begin() + ip, pat.end());
    auto git = group_ids.find(enc);
    if (git == group_ids.end()) {
      git = group_ids.emplace(enc, (uint32_t)gb.size()).first;
      GroupBuild g;
      g.shape_key = last_shape;
      g.lits.assign(enc.begin() + 1, enc.begin() + 1 + n_lits);
      g.rem.assign(enc.begin() + 2 + n_lits, enc.end());
      gb.push_back(std::move(g));
    }
    GroupBuild& dst = gb[git->second];
    dst.queues.push_back(qid.at(std::get<1>(t)));
    dst.durable = dst.durable || std::get<2>(t);
  }

  out.n_words = words.size();
  lap("patterns to groups");
  // shapes: exact ones by length, then the '#' kinds by prefix length
  std::vector<uint32_t> order(shape_list.size());
  for (uint32_t i = 0; i < order.size(); ++i) order[i] = i;
  auto rank_of = [&](uint32_t i) {
    auto& s = shape_list[i];
    return std::make_tuple(std::get<0>(s) == kExact ? 0 : 1, std::get<1>(s), std::get<0>(s), std::get<2>(s));
  };
  std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return rank_of(a) < rank_of(b); });
  std::vector<uint32_t> new_shape(shape_list.size());
  out.shapes.resize(shape_list.size());
  for (uint32_t pos = 0; pos < order.size(); ++pos) {
    auto& s = shape_list[order[pos]];
    new_shape[order[pos]] = pos;
    Shape sh{};
    sh.kind = std::get<0>(s);
    sh.len = std::get<1>(s);
    sh.lit_mask = std::get<2>(s);
    sh.seed = fin(mix(mix(0x243F6A8885A308D3ull, ((uint32_t)sh.kind << 8) | sh.len), sh.lit_mask));
    out.shapes[pos] = sh;
  }
  {
    uint32_t pos = 0;
    for (int L = 0; L <= kLenCap; ++L) {
      out.exact_rng[3 * L] = (uint16_t)pos;
      while (pos < out.shapes.size() && out.shapes[pos].kind == kExact && out.shapes[pos].len == L) ++pos;
      out.exact_rng[3 * L + 1] = (uint16_t)pos;
    }
    out.hashy_begin = pos;
    out.hashy_end = (uint32_t)out.shapes.size();
    for (int L = 0; L <= kLenCap; ++L) {  // '#' shapes whose prefix fits in a key of L words
      uint32_t e = out.hashy_begin;
      while (e < out.hashy_end && out.shapes[e].len <= L) ++e;
      out.exact_rng[3 * L + 2] = (uint16_t)e;
    }
  }
  out.wstride = any_complex ? (uint32_t)kMaxWords : std::max<uint32_t>(4, (max_len + 3) & ~3u);

  // groups, literal tuples, remainders, queue lists; "shared" = some queue sits in two groups
  std::vector<uint32_t> queue_uses(out.queue_names.size(), 0), first_group(out.queue_names.size(), kNone);
  for (uint32_t gi = 0; gi < gb.size(); ++gi) {
    GroupBuild& g = gb[gi];
    std::sort(g.queues.begin(), g.queues.end());
    g.queues.erase(std::unique(g.queues.begin(), g.queues.end()), g.queues.end());
    for (uint32_t q : g.queues)
      if (queue_uses[q]++ == 0) first_group[q] = gi;
  }
  out.groups.resize(gb.size());
  // chain heads keyed by (shape, literal tuple): groups that differ only in their remainder share one
  std::unordered_map<std::vector<uint32_t>, uint32_t, VecHash> heads;
  heads.reserve(gb.size() * 2);
  std::vector<std::pair<uint32_t, uint32_t>> head_hash;  // (hash, head group)
  for (uint32_t gi = 0; gi < gb.size(); ++gi) {
    GroupBuild& g = gb[gi];
    Group& d = out.groups[gi];
    d = Group{};
    d.shape = (uint16_t)new_shape[g.shape_key];
    // lists worth a bulk copy start on a 16-byte boundary (k_write stages them with cp.async.bulk, weave.cuh)
    if (g.queues.size() >= 8)
      while (out.qlist.size() & 3u) out.qlist.push_back(kNone);
    d.list_off = (uint32_t)out.qlist.size();
    d.list_len = (uint32_t)g.queues.size();
    out.qlist.insert(out.qlist.end(), g.queues.begin(), g.queues.end());
    d.lit_off = (uint32_t)out.lits.size();
    d.nlit = (uint8_t)g.lits.size();
    out.lits.insert(out.lits.end(), g.lits.begin(), g.lits.end());
    d.rem_off = (uint32_t)out.rems.size();
    d.rem_len = (uint32_t)g.rem.size();
    {
      uint32_t li = 0;
      for (uint32_t mk = out.shapes[d.shape].lit_mask; mk; mk &= mk - 1, ++li) {
        const uint32_t pos = (uint32_t)__builtin_ctz(mk);
        if (pos < 8) d.zpos[pos] = zob(pos, g.lits[li]);
      }
    }
    out.rems.insert(out.rems.end(), g.rem.begin(), g.rem.end());
    d.flags = (g.durable ? kAnyDurable : 0) | (g.rem.empty() ? 0 : kHasRem);
    for (uint32_t q : g.queues)
      if (queue_uses[q] > 1) { d.flags |= kShared; break; }
    d.next = kNone;
    enc.clear();
    enc.push_back(d.shape);
    enc.insert(enc.end(), g.lits.begin(), g.lits.end());
    auto hit = heads.find(enc);
    if (hit == heads.end()) {
      heads.emplace(enc, gi);
      uint32_t h = (uint32_t)out.shapes[d.shape].seed;
      uint32_t li = 0;
      for (uint32_t mk = out.shapes[d.shape].lit_mask; mk; mk &= mk - 1, ++li)
        h ^= zob((uint32_t)__builtin_ctz(mk), g.lits[li]);
      head_hash.emplace_back(h, gi);
    } else {
      d.next = out.groups[hit->second].next;  // push behind the head
      out.groups[hit->second].next = gi;
    }
  }
  out.n_heads = head_hash.size();
  out.qlist.insert(out.qlist.end(), 4, kNone);  // a bulk copy moves whole 16-byte units: the last list may be read past its end

  lap("groups, lists, chain heads");
  // hash index over the chain heads: eight 16-bit tags per bucket, about 3/8 full, linear probing
  // by bucket; a bucket with an empty slot ends a probe sequence (no deletions: commit rebuilds)
  uint32_t nb = pow2_at_least(std::max<uint64_t>(1, (head_hash.size() + 2) / 3));
  out.tags.assign((size_t)nb * kBucketSlots, 0);
  out.slot_group.assign((size_t)nb * kBucketSlots, kNone);
  for (auto& hh : head_hash) {
    uint32_t b = hh.first & (nb - 1);
    for (;;) {
      uint16_t* t = &out.tags[(size_t)b * kBucketSlots];
      int i = 0;
      while (i < kBucketSlots && t[i] != 0) ++i;
      if (i < kBucketSlots) {
        t[i] = (uint16_t)tag16(hh.first);
        out.slot_group[(size_t)b * kBucketSlots + i] = hh.second;
        break;
      }
      b = (b + 1) & (nb - 1);
    }
  }

  lap("tag index");
  // word dictionary (open addressing on the word hash, at most half full)
  uint32_t nd = pow2_at_least(std::max<uint64_t>(2, words.size() * 2));
  out.dict.assign(1, DictSlot{});
  out.wdict.assign(nd, WordSlot{});
  out.wdict_off.assign(nd, 0);
  for (auto& kv : words) {
    const uint32_t h = word_hash(kv.first);
    WordSlot s{};
    s.first8 = first8(kv.first);
    s.id = kv.second;
    s.tag_len = (h & 0xFFFFFF00u) | (uint32_t)kv.first.size();
    uint32_t i = h & (nd - 1);
    while (out.wdict[i].id != 0) i = (i + 1) & (nd - 1);
    out.wdict[i] = s;
    if (kv.first.size() > 8) {
      out.wdict_off[i] = (uint32_t)out.dict_bytes.size();
      out.dict_bytes.insert(out.dict_bytes.end(), kv.first.begin(), kv.first.end());
    }
  }
  if (out.dict_bytes.empty()) out.dict_bytes.push_back(0);

  // what single binds need later to be added in place (host_inc.cpp)
  inc_.words.reserve(words.size() * 2);
  for (auto& kv : words) inc_.words.emplace(std::string(kv.first), kv.second);
  inc_.shape_ids = std::move(shape_ids);
  inc_.new_shape = std::move(new_shape);
  inc_.group_ids = std::move(group_ids);
  inc_.heads = std::move(heads);
  inc_.qid.reserve(out.queue_names.size() * 2);
  for (uint32_t i = 0; i < out.queue_names.size(); ++i) inc_.qid.emplace(out.queue_names[i], i);
  inc_.queue_uses = std::move(queue_uses);
  inc_.first_group = std::move(first_group);
  inc_.list_cap.resize(out.groups.size());
  for (uint32_t gi = 0; gi < out.groups.size(); ++gi) inc_.list_cap[gi] = out.groups[gi].list_len;
  inc_.max_len = max_len;
  inc_.valid = true;
  lap("state for incremental commits");
}

}  // namespace chq
