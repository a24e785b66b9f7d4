// ORACLE — TEST INFRASTRUCTURE ONLY (see trie_compat.hpp).
//
// SPEC-mode topic matching: AMQP 0-9-1 topic-exchange semantics, "*" = exactly one word,
// "#" = zero or more words.  PARITY UNPINNED: the reference implements neither (its WILDCARD is
// "*" only, QueueMatcher.scala:70, and "#" is an ordinary word there), so there is nothing under
// /root/reference to compare with.  This file is written from the AMQP text
// independently of the CUDA kernels; spec_match() is a plain DP and SpecTrie an index that must
// agree with it (tests/test_oracle_spec.py), and on "#"-free tables both must agree with the
// COMPAT trie.  Tokenisation stays Java split so the two modes share inputs.
#pragma once
#include <cstdint>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "trie_compat.hpp"

namespace orc {

// Plain O(|p|*|k|) dynamic programme over words.
inline bool spec_match(const Words& pw, const Words& kw) {
  const size_t np = pw.size(), nk = kw.size();
  // reach[j] = pattern prefix processed so far can consume exactly j key words
  std::vector<char> reach(nk + 1, 0), next(nk + 1, 0);
  reach[0] = 1;
  for (size_t i = 0; i < np; ++i) {
    std::fill(next.begin(), next.end(), 0);
    if (pw[i] == "#") {
      bool any = false;
      for (size_t j = 0; j <= nk; ++j) {
        any = any || reach[j];
        next[j] = any;
      }
    } else {
      const bool star = pw[i] == "*";
      for (size_t j = 0; j < nk; ++j)
        if (reach[j] && (star || pw[i] == kw[j])) next[j + 1] = 1;
    }
    reach.swap(next);
  }
  return reach[nk];
}

// Word trie with "*" and "#" edges; lookup explores (node, position) states once each.
struct SpecTrie : QueueMatcher {
  struct Node {
    std::unordered_map<std::string, int> next;
    SubSet subs;
  };
  std::vector<Node> nodes;
  SpecTrie() { nodes.emplace_back(); }

  int subscribe(const std::string& topic, const Subscriber& sub) override {
    Words w = java_split_dot(topic);
    if (w.empty()) return -1;
    int cur = 0;
    for (auto& word : w) {
      auto it = nodes[cur].next.find(word);
      if (it == nodes[cur].next.end()) {
        int id = (int)nodes.size();
        nodes[cur].next.emplace(word, id);
        nodes.emplace_back();
        cur = id;
      } else {
        cur = it->second;
      }
    }
    nodes[cur].subs.emplace(sub, 0);
    return 0;
  }
  int unsubscribe(const std::string& topic, const Subscriber& sub) override {
    Words w = java_split_dot(topic);
    if (w.empty()) return -1;
    int cur = 0;
    for (auto& word : w) {
      auto it = nodes[cur].next.find(word);
      if (it == nodes[cur].next.end()) return 0;
      cur = it->second;
    }
    nodes[cur].subs.erase(sub);
    return 0;
  }
  void walk(int node, const Words& kw, size_t pos, std::unordered_set<uint64_t>& seen, SubSet& out) const {
    if (!seen.insert(((uint64_t)node << 16) | pos).second) return;
    const Node& n = nodes[node];
    if (pos == kw.size())
      for (auto& kv : n.subs) out.emplace(kv.first, 0);
    auto hash = n.next.find("#");
    if (hash != n.next.end())
      for (size_t p = pos; p <= kw.size(); ++p) walk(hash->second, kw, p, seen, out);
    if (pos < kw.size()) {
      auto star = n.next.find("*");
      if (star != n.next.end()) walk(star->second, kw, pos + 1, seen, out);
      auto ex = n.next.find(kw[pos]);
      // a key word spelled "*" or "#" is still just a word; it must not ride the wildcard edge twice
      if (ex != n.next.end() && kw[pos] != "*" && kw[pos] != "#") walk(ex->second, kw, pos + 1, seen, out);
    }
  }
  int lookup(const std::string& topic, SubSet& out) override {
    Words kw = java_split_dot(topic);
    if (kw.empty()) return -1;
    std::unordered_set<uint64_t> seen;
    walk(0, kw, 0, seen, out);
    return 0;
  }
};

}  // namespace orc
