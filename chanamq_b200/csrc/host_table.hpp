// Host side of one exchange's binding table: the live binding set (what ExchangeEntity keeps in
// `binds`/`subscriptions` and feeds to QueueMatcher.subscribe/unsubscribe,
// ExchangeEntity.scala:333-409) and its compilation into the flat arrays of chq_common.h.
// Binding churn is far rarer than publishing, so a commit recompiles the whole table and the
// device copy is swapped at a batch boundary (the exchange actor already serialises bind against
// publish, ExchangeEntity.scala:187-285).
#pragma once
#include <cstdint>
#include <map>
#include <set>
#include <unordered_map>
#include <string>
#include <string_view>
#include <tuple>
#include <vector>

#include "chq_common.h"

namespace chq {

struct CompiledTable {
  std::vector<DictSlot> dict;          // header field names (headers tables)
  std::vector<uint8_t> dict_bytes;
  std::vector<WordSlot> wdict;         // pattern words (topic / direct tables)
  std::vector<uint32_t> wdict_off;
  std::vector<Shape> shapes;
  std::vector<uint16_t> exact_rng;
  uint32_t hashy_begin = 0, hashy_end = 0;
  std::vector<uint16_t> tags;          // [buckets * 8]
  std::vector<uint32_t> slot_group;    // [buckets * 8]
  std::vector<Group> groups;
  std::vector<uint32_t> lits, rems, qlist;
  uint32_t wstride = 4;
  uint32_t fanout_group = kNone;
  // headers exchange, SPEC mode
  bool headers_mode = false;
  std::vector<DictSlot> vdict;
  std::vector<uint8_t> vdict_bytes;
  std::vector<uint32_t> presence_atom, all_atoms;
  std::vector<AtomRec> atoms;
  std::vector<AllGroup> all_groups;
  uint32_t always_off = 0, always_len = 0, always_flags = 0;
  std::vector<std::string> queue_names;  // id -> name, ascending by bytes
  // statistics for chq_stats / the roofline model
  uint64_t n_bindings = 0, n_heads = 0, pattern_words = 0, n_words = 0;
};

// Java String.split("\\.") (QueueMatcher.scala:69,386,395,404): cut at every '.', drop trailing
// empty words, keep leading/interior ones; no '.' at all gives the input back as one word.
void java_split(std::string_view s, std::vector<std::string_view>& out);
// hash of a word / field name / value: word_chunk over 8-byte chunks from `seed`, then word_fin with the length
uint32_t bytes_hash(uint32_t seed, std::string_view w);

// One piece of a committed table that an incremental commit changed on the host: elements [off, off + n) of one of
// its arrays, to be copied over the device copy.
enum PatchArray { kPaWdict, kPaWdictOff, kPaDictBytes, kPaTags, kPaSlotGroup, kPaGroups, kPaLits, kPaRems, kPaQlist, kPaCount };
struct Patch { int array; size_t off, n; };

struct VecHash {  // FNV-style hash of an encoded pattern
  size_t operator()(const std::vector<uint32_t>& v) const {
    uint64_t h = 0xCBF29CE484222325ull;
    for (uint32_t x : v) h = (h ^ x) * 0x100000001B3ull;
    return (size_t)(h ^ (h >> 29));
  }
};

class HostTable {
 public:
  HostTable(int exchange_type, int mode) : type_(exchange_type), mode_(mode) {}
  // 0, or a negative chq status (zero-word pattern, key too long)
  // args = contents of the binding's arguments field table (headers exchanges; empty otherwise)
  int bind(std::string_view key, std::string_view queue, bool durable, std::string_view args = {});
  int unbind(std::string_view key, std::string_view queue, bool durable);
  int unbind_queue(std::string_view queue);
  // a queue that takes part in the id numbering although nothing is bound to it HERE (binding-slice
  // sharding: every GPU numbers the queues of the whole exchange, so ids agree across slices)
  void declare_queue(std::string_view queue) { declared_.emplace(queue); inc_.valid = false; }
  // Incremental commit (host_inc.cpp): applies the binds / unbinds made since the last commit to an already compiled
  // table in place and lists what changed.  false = this set of changes needs a full compile() (a pattern shape the
  // table has not seen, a hash index or dictionary that got too full, a '#' that widens the word rows, ...).
  bool apply_pending(CompiledTable& ct, std::vector<Patch>& patches) const;
  size_t pending() const { return pending_.size(); }
  bool incremental_ready() const { return inc_.valid; }
  void compile(CompiledTable& out) const;
  size_t size() const { return live_.size(); }
  int type() const { return type_; }
  int mode() const { return mode_; }

 private:
  int check_key(std::string_view key) const;
  int type_, mode_;
  // the matcher holds a SET of (pattern, queue, durable) triples (QueueMatcher.scala:443-446)
  void compile_headers(CompiledTable& out) const;  // host_headers.cpp
  std::set<std::tuple<std::string, std::string, bool, std::string>> live_;
  std::set<std::string> declared_;
  // what compile() knew when it built the table, kept so that single binds can be added to it (host_inc.cpp)
  struct Inc {
    bool valid = false;
    std::unordered_map<std::string, uint32_t> words;                        // pattern word -> id
    std::map<std::tuple<uint8_t, uint8_t, uint32_t>, uint32_t> shape_ids;   // (kind, len, mask) -> number in meeting order
    std::vector<uint32_t> new_shape;                                        // that number -> position in CompiledTable::shapes
    std::unordered_map<std::vector<uint32_t>, uint32_t, VecHash> group_ids; // encoded pattern -> group
    std::unordered_map<std::vector<uint32_t>, uint32_t, VecHash> heads;     // (shape, literal tuple) -> chain head
    std::unordered_map<std::string, uint32_t> qid;                          // queue name -> id
    std::vector<uint32_t> queue_uses, first_group;                          // per queue: groups that list it, one of them
    std::vector<uint32_t> list_cap;                                         // per group: room of its list in qlist
    uint32_t max_len = 1;
  };
  mutable Inc inc_;
  struct Op { bool bind; std::string key, queue; bool durable; };
  mutable std::vector<Op> pending_;
  bool apply_bind(CompiledTable& ct, std::vector<Patch>& patches, const Op& op) const;
  bool apply_unbind(CompiledTable& ct, std::vector<Patch>& patches, const Op& op) const;
};

}  // namespace chq
