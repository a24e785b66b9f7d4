// ORACLE — TEST INFRASTRUCTURE ONLY (see trie_compat.hpp).  C ABI so tests and bench.py's
// cpu_baseline / --impl reference legs can drive the restatement through ctypes.
//
// orc_route() restates the per-message body of ExchangeEntity.publish (EX:293-302): lookup,
// collapse subscribers to the SET of queue names (EX:300), OR of their durable flags (EX:302).
// Queue ids are the rank of the queue name among the names that currently hold a binding,
// ascending by bytes — the same canonical numbering the product assigns at commit.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <thread>

#include "match_spec.hpp"
#include "trie_compat.hpp"

using namespace orc;

namespace {

struct Triple {
  std::string key, queue;
  bool durable;
  bool operator<(const Triple& o) const {
    if (key != o.key) return key < o.key;
    if (queue != o.queue) return queue < o.queue;
    return durable < o.durable;
  }
};

struct Handle {
  int type = 2, mode = 0;
  std::unique_ptr<QueueMatcher> m;
  std::set<Triple> live;  // the matcher holds a SET of (pattern, subscriber) pairs (QM:443-446)
  bool ranks_dirty = true;
  std::vector<std::string> names;
  std::unordered_map<std::string, uint32_t> rank;
  // last route() result
  std::vector<uint32_t> offsets, ids;
  std::vector<uint8_t> durable, err;
  double last_seconds = 0;

  void refresh() {
    if (!ranks_dirty) return;
    std::set<std::string> q;
    for (auto& t : live) q.insert(t.queue);
    names.assign(q.begin(), q.end());
    rank.clear();
    for (uint32_t i = 0; i < names.size(); ++i) rank.emplace(names[i], i);
    ranks_dirty = false;
  }
};

void route_range(Handle* h, const uint8_t* blob, const uint32_t* off, uint32_t lo, uint32_t hi,
                 std::vector<std::vector<uint32_t>>& per_msg) {
  for (uint32_t i = lo; i < hi; ++i) {
    std::string key((const char*)blob + off[i], off[i + 1] - off[i]);
    SubSet subs;
    int rc = h->m->lookup(key, subs);
    if (rc == -1) { h->err[i] = 1; continue; }
    std::vector<uint32_t>& v = per_msg[i];
    bool dur = false;
    for (auto& kv : subs) {
      v.push_back(h->rank.at(kv.first.queue));
      dur = dur || kv.first.durable;
    }
    std::sort(v.begin(), v.end());
    v.erase(std::unique(v.begin(), v.end()), v.end());
    h->durable[i] = dur;
  }
}

}  // namespace

extern "C" {

// type: 0 direct, 1 fanout, 2 topic, 3 headers (the reference routes headers as topic, EX:153/215)
// mode: 0 COMPAT (reference-exact), 1 SPEC (AMQP text; parity unpinned)
void* orc_new(int type, int mode) {
  auto* h = new Handle();
  h->type = type;
  h->mode = mode;
  if (type == 0) h->m.reset(new DirectMatcher());
  else if (type == 1) h->m.reset(new FanoutMatcher());
  else if (mode == 1) h->m.reset(new SpecTrie());
  else h->m.reset(new TrieMatcher());
  return h;
}
void orc_free(void* p) { delete (Handle*)p; }

// Bulk loading: build the same trie by in-place updates instead of the reference's copy-on-write
// (QM:183-232 copy a whole branch map per insert; single-threaded the result is identical).
void orc_set_inplace(void* p, int on) {
  auto* h = (Handle*)p;
  if (auto* t = dynamic_cast<TrieMatcher*>(h->m.get())) t->inplace_insert = on != 0;
}

int orc_bind(void* p, const uint8_t* key, uint32_t klen, const uint8_t* queue, uint32_t qlen, int durable) {
  auto* h = (Handle*)p;
  std::string k((const char*)key, klen), q((const char*)queue, qlen);
  int rc = h->m->subscribe(k, Subscriber{q, durable != 0});
  if (rc == 0) {
    h->live.insert(Triple{k, q, durable != 0});
    h->ranks_dirty = true;
  }
  return rc;
}
int orc_unbind(void* p, const uint8_t* key, uint32_t klen, const uint8_t* queue, uint32_t qlen, int durable) {
  auto* h = (Handle*)p;
  std::string k((const char*)key, klen), q((const char*)queue, qlen);
  int rc = h->m->unsubscribe(k, Subscriber{q, durable != 0});
  if (rc == 0) {
    h->live.erase(Triple{k, q, durable != 0});
    h->ranks_dirty = true;
  }
  return rc;
}

// bulk bind (blob + offsets form), same semantics as orc_bind per element; returns first error
int orc_bind_many(void* p, const uint8_t* kb, const uint32_t* ko, const uint8_t* qb, const uint32_t* qo,
                  const uint8_t* durable, uint32_t n) {
  int first = 0;
  for (uint32_t i = 0; i < n; ++i) {
    int rc = orc_bind(p, kb + ko[i], ko[i + 1] - ko[i], qb + qo[i], qo[i + 1] - qo[i], durable ? durable[i] : 0);
    if (rc && !first) first = rc;
  }
  return first;
}

uint32_t orc_n_queues(void* p) {
  auto* h = (Handle*)p;
  h->refresh();
  return (uint32_t)h->names.size();
}
int orc_queue_name(void* p, uint32_t id, const uint8_t** name, uint32_t* len) {
  auto* h = (Handle*)p;
  h->refresh();
  if (id >= h->names.size()) return -1;
  *name = (const uint8_t*)h->names[id].data();
  *len = (uint32_t)h->names[id].size();
  return 0;
}

// Routes n keys (blob + n+1 offsets) on `threads` host threads (keys sharded; the matcher is only
// read, except for the reference's own tomb clean-up, so threads > 1 is for bind-only tables).
// Returns the total number of queue ids, or -1.  Results stay in the handle.
int64_t orc_route(void* p, const uint8_t* blob, const uint32_t* off, uint32_t n, int threads) {
  auto* h = (Handle*)p;
  h->refresh();
  h->durable.assign(n, 0);
  h->err.assign(n, 0);
  std::vector<std::vector<uint32_t>> per_msg(n);
  auto t0 = std::chrono::steady_clock::now();
  if (threads <= 1 || n < 1024) {
    route_range(h, blob, off, 0, n, per_msg);
  } else {
    std::vector<std::thread> pool;
    uint32_t chunk = (n + threads - 1) / threads;
    for (int t = 0; t < threads; ++t) {
      uint32_t lo = std::min<uint64_t>((uint64_t)t * chunk, n), hi = std::min<uint64_t>((uint64_t)lo + chunk, n);
      if (lo < hi) pool.emplace_back(route_range, h, blob, off, lo, hi, std::ref(per_msg));
    }
    for (auto& th : pool) th.join();
  }
  h->last_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  h->offsets.assign(n + 1, 0);
  uint64_t total = 0;
  for (uint32_t i = 0; i < n; ++i) {
    h->offsets[i] = (uint32_t)total;
    total += per_msg[i].size();
  }
  if (total > 0xFFFFFFFFull) return -1;
  h->offsets[n] = (uint32_t)total;
  h->ids.resize(total);
  for (uint32_t i = 0; i < n; ++i)
    if (!per_msg[i].empty())
      memcpy(h->ids.data() + h->offsets[i], per_msg[i].data(), per_msg[i].size() * 4);
  return (int64_t)total;
}
const uint32_t* orc_offsets(void* p) { return ((Handle*)p)->offsets.data(); }
const uint32_t* orc_ids(void* p) { return ((Handle*)p)->ids.data(); }
const uint8_t* orc_durable(void* p) { return ((Handle*)p)->durable.data(); }
const uint8_t* orc_err(void* p) { return ((Handle*)p)->err.data(); }
// seconds spent in the lookups of the last orc_route() (the part EX:293-302 covers)
double orc_last_seconds(void* p) { return ((Handle*)p)->last_seconds; }

// Pair predicates (second and third opinions): 1 match, 0 no match, -1 zero-word input.
int orc_pair_match(int mode, const uint8_t* pat, uint32_t plen, const uint8_t* key, uint32_t klen) {
  Words pw = java_split_dot(std::string((const char*)pat, plen));
  Words kw = java_split_dot(std::string((const char*)key, klen));
  if (pw.empty() || kw.empty()) return -1;
  return mode == 1 ? spec_match(pw, kw) : compat_match(pw, kw);
}
// Java split as a NUL-free, '\n'-joined list written to out (returns word count or -1 if cap too small).
int orc_split(const uint8_t* s, uint32_t len, uint8_t* out, uint32_t cap, uint32_t* out_len) {
  Words w = java_split_dot(std::string((const char*)s, len));
  std::string j;
  for (size_t i = 0; i < w.size(); ++i) {
    if (i) j.push_back('\n');
    j += w[i];
  }
  if (j.size() > cap) return -1;
  memcpy(out, j.data(), j.size());
  *out_len = (uint32_t)j.size();
  return (int)w.size();
}

}  // extern "C"
