// ORACLE — TEST INFRASTRUCTURE ONLY (see trie_compat.hpp).
//
// Headers-exchange routing, AMQP 0-9-1 x-match all/any.  PARITY UNPINNED: the reference has no headers
// matcher ("TODO header matcher ?", chana-mq-server/.../entity/ExchangeEntity.scala:210; a headers exchange
// gets a TrieMatcher on the routing key, ExchangeEntity.scala:153,215), so nothing under /root/reference can
// be compared with.  Written from the rule restated in SURVEY.md Appendix C, as a brute force over
// (message, binding) pairs, independently of the CUDA index (atoms, anchors) it checks:
//   arguments table: x-match = "all" (default) | "any"; other names starting "x-" ignored;
//   an argument matches when the message has a header of that name and the argument is void or the
//   header has the same type tag and the same value bytes; all: every argument matches (none: always);
//   any: at least one matches (none: never).  Duplicate names in a table: the first one counts
//   (chana-mq-base/.../model/ValueReader.scala:71-72).  The routing key plays no part.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <set>
#include <string>
#include <vector>

namespace {

struct Field { std::string name; uint8_t tag; std::string value; };

bool parse(const uint8_t* t, size_t n, std::vector<Field>& out) {
  size_t p = 0;
  while (p < n) {
    size_t nl = t[p++];
    if (p + nl + 1 > n) return false;
    std::string name((const char*)t + p, nl);
    p += nl;
    uint8_t tag = t[p++];
    size_t vl;
    switch (tag) {
      case 'S': case 'x': case 'F': case 'A':
        if (p + 4 > n) return false;
        vl = ((size_t)t[p] << 24) | ((size_t)t[p + 1] << 16) | ((size_t)t[p + 2] << 8) | t[p + 3];
        p += 4;
        break;
      case 'I': case 'f': vl = 4; break;
      case 'D': vl = 5; break;
      case 'T': case 'd': case 'l': vl = 8; break;
      case 'b': case 't': vl = 1; break;
      case 's': vl = 2; break;
      case 'V': vl = 0; break;
      default: return false;
    }
    if (p + vl > n) return false;
    bool dup = false;
    for (auto& f : out) dup = dup || f.name == name;
    if (!dup) out.push_back(Field{name, tag, std::string((const char*)t + p, vl)});
    p += vl;
  }
  return true;
}

struct Binding { std::vector<Field> args; bool any; std::string queue; bool durable; };

struct H {
  std::vector<Binding> binds;
  std::vector<std::string> names;
  std::vector<uint32_t> offsets, ids;
  std::vector<uint8_t> durable, err;
};

}  // namespace

extern "C" {

void* orch_new() { return new H(); }
void orch_free(void* p) { delete (H*)p; }

int orch_bind(void* p, const uint8_t* queue, uint32_t qlen, int durable, const uint8_t* args, uint32_t alen) {
  H* h = (H*)p;
  std::vector<Field> f;
  if (!parse(args, alen, f)) return -5;
  Binding b;
  b.any = false;
  b.queue.assign((const char*)queue, qlen);
  b.durable = durable != 0;
  for (auto& x : f) {
    if (x.name == "x-match") { b.any = x.tag == 'S' && x.value == "any"; continue; }
    if (x.name.rfind("x-", 0) == 0) continue;
    b.args.push_back(x);
  }
  h->binds.push_back(b);
  return 0;
}

uint32_t orch_n_queues(void* p) {
  H* h = (H*)p;
  std::set<std::string> q;
  for (auto& b : h->binds) q.insert(b.queue);
  h->names.assign(q.begin(), q.end());
  return (uint32_t)h->names.size();
}
int orch_queue_name(void* p, uint32_t id, const uint8_t** name, uint32_t* len) {
  H* h = (H*)p;
  if (id >= h->names.size()) return -1;
  *name = (const uint8_t*)h->names[id].data();
  *len = (uint32_t)h->names[id].size();
  return 0;
}

int64_t orch_route(void* p, const uint8_t* blob, const uint32_t* off, uint32_t n) {
  H* h = (H*)p;
  orch_n_queues(p);
  std::vector<uint32_t> bq(h->binds.size());
  for (size_t i = 0; i < h->binds.size(); ++i)
    bq[i] = (uint32_t)(std::lower_bound(h->names.begin(), h->names.end(), h->binds[i].queue) - h->names.begin());
  h->offsets.assign(n + 1, 0);
  h->ids.clear();
  h->durable.assign(n, 0);
  h->err.assign(n, 0);
  for (uint32_t m = 0; m < n; ++m) {
    h->offsets[m] = (uint32_t)h->ids.size();
    std::vector<Field> msg;
    if (!parse(blob + off[m], off[m + 1] - off[m], msg)) { h->err[m] = 3; continue; }
    std::vector<uint32_t> q;
    bool dur = false;
    for (size_t i = 0; i < h->binds.size(); ++i) {
      const Binding& b = h->binds[i];
      size_t hit = 0;
      for (auto& a : b.args) {
        for (auto& f : msg)
          if (f.name == a.name) { if (a.tag == 'V' || (a.tag == f.tag && a.value == f.value)) ++hit; break; }
      }
      const bool match = b.any ? hit > 0 : hit == b.args.size();
      if (match) { q.push_back(bq[i]); dur = dur || b.durable; }
    }
    std::sort(q.begin(), q.end());
    q.erase(std::unique(q.begin(), q.end()), q.end());
    h->ids.insert(h->ids.end(), q.begin(), q.end());
    h->durable[m] = dur;
  }
  h->offsets[n] = (uint32_t)h->ids.size();
  return (int64_t)h->ids.size();
}
const uint32_t* orch_offsets(void* p) { return ((H*)p)->offsets.data(); }
const uint32_t* orch_ids(void* p) { return ((H*)p)->ids.data(); }
const uint8_t* orch_durable(void* p) { return ((H*)p)->durable.data(); }
const uint8_t* orch_err(void* p) { return ((H*)p)->err.data(); }

}  // extern "C"
