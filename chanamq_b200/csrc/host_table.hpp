// Host side of one exchange's binding table: the live binding set (what ExchangeEntity keeps in
// `binds`/`subscriptions` and feeds to QueueMatcher.subscribe/unsubscribe,
// ExchangeEntity.scala:333-409) and its compilation into the flat arrays of chq_common.h.
// Binding churn is far rarer than publishing, so a commit recompiles the whole table and the
// device copy is swapped at a batch boundary (the exchange actor already serialises bind against
// publish, ExchangeEntity.scala:187-285).
#pragma once
#include <cstdint>
#include <set>
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
  void declare_queue(std::string_view queue) { declared_.emplace(queue); }
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
};

}  // namespace chq
