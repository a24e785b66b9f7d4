// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path: only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
//
// C++17 restatement of the reference's CPU routing path (no JVM exists in this image, so the
// Scala itself cannot run here).  Every function cites the reference lines it follows;
// abbreviations: QM = chana-mq-server/src/main/scala/chana/mq/amqp/server/engine/QueueMatcher.scala,
//                EX = chana-mq-server/src/main/scala/chana/mq/amqp/server/entity/ExchangeEntity.scala.
//
// Parity status: PINNED for COMPAT mode by the reference's only known-answer test,
// TrieMatcher.main (QM:75-139), reproduced in tests/test_oracle_kat.py.
#pragma once
#include <cstdint>
#include <map>
#include <memory>
#include <set>
#include <string>
#include <unordered_map>
#include <vector>

namespace orc {

// ---------------------------------------------------------------------------------------------
// Java `String.split("\\.")` (QM:69, QM:386, QM:395, QM:404): cut at every '.', then drop ALL
// trailing empty strings; an input without any '.' comes back as the single element [input]
// (so "" -> [""]), while "." and ".." come back empty.
// ---------------------------------------------------------------------------------------------
inline std::vector<std::string> java_split_dot(const std::string& s) {
  std::vector<std::string> out;
  size_t start = 0;
  bool found = false;
  for (size_t i = 0; i < s.size(); ++i) {
    if (s[i] == '.') {
      out.emplace_back(s, start, i - start);
      start = i + 1;
      found = true;
    }
  }
  if (!found) { out.clear(); out.push_back(s); return out; }
  out.emplace_back(s, start, s.size() - start);
  while (!out.empty() && out.back().empty()) out.pop_back();
  return out;
}

// ExchangeEntity.TopicSubscriber(queue, isDurable) — case-class equality on both fields (EX:97).
struct Subscriber {
  std::string queue;
  bool durable = false;
  bool operator==(const Subscriber& o) const { return durable == o.durable && queue == o.queue; }
  bool operator<(const Subscriber& o) const {
    if (queue != o.queue) return queue < o.queue;
    return durable < o.durable;
  }
};
struct SubscriberHash {
  size_t operator()(const Subscriber& s) const {
    return std::hash<std::string>()(s.queue) * 31u + (s.durable ? 1231u : 1237u);
  }
};
using SubSet = std::unordered_map<Subscriber, int, SubscriberHash>;  // Map[Subscriber, AnyRef]

struct QueueMatcher {  // QM:11-27
  virtual ~QueueMatcher() {}
  // returns 0, or a negative code: -1 zero-word topic (ArrayIndexOutOfBounds, QM:420/469/528),
  // -2 the trie is stuck on a tombstone (unbounded retry in the reference).
  virtual int subscribe(const std::string& topic, const Subscriber& sub) = 0;
  virtual int unsubscribe(const std::string& topic, const Subscriber& sub) = 0;
  virtual int lookup(const std::string& topic, SubSet& out) = 0;
};

// DirectMatcher (QM:29-48): whole-key string equality.
struct DirectMatcher : QueueMatcher {
  std::unordered_map<std::string, std::set<Subscriber>> topic_to_subs;
  int subscribe(const std::string& topic, const Subscriber& sub) override {
    topic_to_subs[topic].insert(sub);  // HashSet[Subscription]: (topic, sub) pairs, QM:32-37
    return 0;
  }
  int unsubscribe(const std::string& topic, const Subscriber& sub) override {
    auto it = topic_to_subs.find(topic);  // QM:39-43 (the emptied entry stays)
    if (it != topic_to_subs.end()) it->second.erase(sub);
    return 0;
  }
  int lookup(const std::string& topic, SubSet& out) override {
    auto it = topic_to_subs.find(topic);  // QM:45-47
    if (it != topic_to_subs.end())
      for (auto& s : it->second) out.emplace(s, 0);
    return 0;
  }
};

// FanoutMatcher (QM:50-66): every subscriber, whatever the key.
struct FanoutMatcher : QueueMatcher {
  std::set<std::pair<std::string, Subscriber>> subs;
  int subscribe(const std::string& topic, const Subscriber& sub) override {
    subs.insert({topic, sub});
    return 0;
  }
  int unsubscribe(const std::string& topic, const Subscriber& sub) override {
    subs.erase({topic, sub});
    return 0;
  }
  int lookup(const std::string&, SubSet& out) override {
    for (auto& p : subs) out.emplace(p.second, 0);
    return 0;
  }
};

// ---------------------------------------------------------------------------------------------
// TrieMatcher (QM:68-601): a "CS-trie": I-node -> main node -> (C-node | tomb).  A C-node maps a
// word to a branch; a branch carries the subscribers that END at this word and an optional
// I-node for longer topics.  Lookup follows the exact word and the literal word "*" at each level.
// Single-threaded here, so every compareAndSet in the reference succeeds.
// ---------------------------------------------------------------------------------------------
struct INode;
struct Branch {
  SubSet subs;
  std::shared_ptr<INode> inode;
};
using BranchMap = std::unordered_map<std::string, std::shared_ptr<Branch>>;
struct CNode { BranchMap branches; };
struct MainNode {
  std::shared_ptr<CNode> cnode;  // null when tombed
  bool tomb = false;             // tNode != null
};
struct INode { std::shared_ptr<MainNode> main; };

using Words = std::vector<std::string>;

struct TrieMatcher : QueueMatcher {
  static constexpr const char* WILDCARD = "*";  // QM:70
  std::shared_ptr<INode> root;
  bool inplace_insert = false;  // see subscribe()

  TrieMatcher() {
    root = std::make_shared<INode>();
    root->main = mk_main(std::make_shared<CNode>());
  }

  static std::shared_ptr<MainNode> mk_main(std::shared_ptr<CNode> c) {
    auto m = std::make_shared<MainNode>();
    m->cnode = std::move(c);
    return m;
  }
  static std::shared_ptr<MainNode> mk_tomb() {
    auto m = std::make_shared<MainNode>();
    m->tomb = true;
    return m;
  }
  static Words tail(const Words& w) { return Words(w.begin() + 1, w.end()); }  // words.slice(1, n)

  // newCNode (QM:270-278): a fresh chain of C-nodes for the remaining words.
  static std::shared_ptr<CNode> new_cnode(const Words& words, const Subscriber& sub) {
    auto cn = std::make_shared<CNode>();
    auto br = std::make_shared<Branch>();
    if (words.size() == 1) {
      br->subs.emplace(sub, 0);
    } else {
      br->inode = std::make_shared<INode>();
      br->inode->main = mk_main(new_cnode(tail(words), sub));
    }
    cn->branches.emplace(words[0], br);
    return cn;
  }
  // CNode.inserted (QM:183-196): copy of the node plus a new branch for words(0).
  static std::shared_ptr<CNode> cn_inserted(const CNode& cn, const Words& words, const Subscriber& sub) {
    auto out = std::make_shared<CNode>(cn);
    auto br = std::make_shared<Branch>();
    if (words.size() == 1) {
      br->subs.emplace(sub, 0);
    } else {
      br->inode = std::make_shared<INode>();
      br->inode->main = mk_main(new_cnode(tail(words), sub));
    }
    out->branches[words[0]] = br;
    return out;
  }
  // CNode.updated (QM:201-219): copy with `sub` added to the branch of `word`.
  static std::shared_ptr<CNode> cn_updated(const CNode& cn, const std::string& word, const Subscriber& sub) {
    auto out = std::make_shared<CNode>(cn);
    auto nb = std::make_shared<Branch>();
    nb->subs.emplace(sub, 0);
    auto it = out->branches.find(word);
    if (it != out->branches.end()) {
      for (auto& kv : it->second->subs) nb->subs.emplace(kv.first, 0);
      nb->inode = it->second->inode;
    }
    out->branches[word] = nb;
    return out;
  }
  // CNode.updatedBranch (QM:224-232) with Branch.updated (QM:150-157).
  static std::shared_ptr<CNode> cn_updated_branch(const CNode& cn, const std::string& word,
                                                  std::shared_ptr<INode> in, const Branch& br) {
    auto out = std::make_shared<CNode>(cn);
    auto nb = std::make_shared<Branch>();
    nb->subs = br.subs;
    nb->inode = std::move(in);
    out->branches[word] = nb;
    return out;
  }
  // CNode.removed (QM:237-255) with Branch.removed (QM:162-165).
  static std::shared_ptr<CNode> cn_removed(const CNode& cn, const std::string& word, const Subscriber& sub) {
    auto out = std::make_shared<CNode>(cn);
    auto it = out->branches.find(word);
    if (it != out->branches.end()) {
      auto nb = std::make_shared<Branch>();
      for (auto& kv : it->second->subs)
        if (!(kv.first == sub)) nb->subs.emplace(kv.first, 0);
      nb->inode = it->second->inode;
      if (nb->subs.empty() && !nb->inode) out->branches.erase(it);
      else it->second = nb;
    }
    return out;
  }
  // prunable (QM:367-376), toCompressed (QM:352-360), clean (QM:283-288)
  static bool prunable(const Branch& br) {
    if (!br.subs.empty()) return false;
    if (!br.inode) return true;
    return br.inode->main->tomb;
  }
  static std::shared_ptr<MainNode> to_compressed(const CNode& cn) {
    auto c = std::make_shared<CNode>();
    for (auto& kv : cn.branches)
      if (!prunable(*kv.second)) c->branches.emplace(kv.first, kv.second);
    return mk_main(c);
  }
  static void clean(INode* i) {
    auto m = i->main;
    if (m->cnode) i->main = to_compressed(*m->cnode);
  }
  // toContracted (QM:592-598)
  std::shared_ptr<MainNode> to_contracted(std::shared_ptr<CNode> cn, const INode* parent) const {
    if (root.get() != parent && cn->branches.empty()) return mk_tomb();
    return mk_main(std::move(cn));
  }
  // contract (QM:322-346)
  bool contract(INode* parents_parent, INode* parent, const MainNode& pmain) {
    auto ncn = to_compressed(*pmain.cnode);
    if (ncn->cnode->branches.empty() && parents_parent != nullptr) {
      auto ppmain = parents_parent->main;
      for (auto& kv : ppmain->cnode->branches) {
        if (kv.second->inode.get() == parent) {
          auto upd = cn_updated_branch(*ppmain->cnode, kv.first, nullptr, *kv.second);
          if (kv.second->subs.empty()) upd->branches.erase(kv.first);
          parents_parent->main = to_compressed(*upd);
          return true;
        }
      }
    } else {
      parent->main = to_contracted(ncn->cnode, parent);
      return true;
    }
    return true;
  }
  // cleanParent (QM:296-316)
  void clean_parent(INode* i, INode* parent, INode* parents_parent, const std::string& word) {
    auto main = i->main;
    auto pmain = parent->main;
    if (pmain->cnode) {
      auto it = pmain->cnode->branches.find(word);
      if (it != pmain->cnode->branches.end()) {
        if (it->second->inode.get() != i) return;
        if (main->tomb) contract(parents_parent, parent, *pmain);  // CAS cannot fail here
      }
    }
  }

  // iinsert (QM:415-460).  1 = done, 0 = retry, -2 = invalid state.
  int iinsert(INode* i, INode* parent, const Words& words, const Subscriber& sub) {
    auto main = i->main;
    if (main->cnode) {
      auto& cn = *main->cnode;
      auto it = cn.branches.find(words[0]);
      if (it == cn.branches.end()) {
        if (inplace_insert) {  // same resulting node as cn_inserted, without the map copy
          auto tmp = new_cnode(words, sub);
          cn.branches.emplace(words[0], tmp->branches.begin()->second);
        } else {
          i->main = mk_main(cn_inserted(cn, words, sub));
        }
        return 1;
      }
      auto br = it->second;
      if (words.size() > 1) {
        if (br->inode) return iinsert(br->inode.get(), i, tail(words), sub);
        auto nin = std::make_shared<INode>();
        nin->main = mk_main(new_cnode(tail(words), sub));
        if (inplace_insert) br->inode = nin;
        else i->main = mk_main(cn_updated_branch(cn, words[0], nin, *br));
        return 1;
      }
      if (br->subs.count(sub)) return 1;  // already subscribed, QM:443-446
      if (inplace_insert) br->subs.emplace(sub, 0);
      else i->main = mk_main(cn_updated(cn, words[0], sub));
      return 1;
    } else if (main->tomb) {
      clean(parent);
      return 0;
    }
    return -2;
  }

  // iremove (QM:462-516)
  int iremove(INode* i, INode* parent, INode* parents_parent, const Words& words, size_t idx,
              const Subscriber& sub) {
    auto main = i->main;
    if (main->cnode) {
      auto& cn = *main->cnode;
      auto it = cn.branches.find(words[idx]);
      if (it == cn.branches.end()) return 1;
      auto br = it->second;
      if (idx + 1 < words.size()) {
        if (br->inode) return iremove(br->inode.get(), i, parent, words, idx + 1, sub);
        return 1;
      }
      if (!br->subs.count(sub)) return 1;
      auto ncn = cn_removed(cn, words[idx], sub);
      i->main = to_contracted(ncn, i);
      if (parent != nullptr && i->main->tomb) clean_parent(i, parent, parents_parent, words[idx - 1]);
      return 1;
    } else if (main->tomb) {
      clean(parent);
      return 0;
    }
    return -2;
  }

  // ilookup (QM:523-562) and bLookup (QM:569-585).  Returns false when the reference would
  // ask for a retry (it then answers Set(), QM:409-411).
  bool ilookup(INode* i, INode* parent, const Words& words, SubSet& subs) {
    auto main = i->main;
    if (main->cnode) {
      auto& br = main->cnode->branches;
      auto exact = br.find(words[0]);       // getBranches (QM:261-263): the word itself ...
      auto wild = br.find(WILDCARD);        // ... and the single-word wildcard
      SubSet local;
      if (exact != br.end() && !blookup(i, *exact->second, words, local)) return false;
      if (wild != br.end() && !blookup(i, *wild->second, words, local)) return false;
      for (auto& kv : local) subs.emplace(kv.first, 0);
      return true;
    } else if (main->tomb) {
      clean(parent);
      return false;
    }
    return false;
  }
  bool blookup(INode* i, const Branch& b, const Words& words, SubSet& subs) {
    if (words.size() > 1) {
      if (!b.inode) return true;
      return ilookup(b.inode.get(), i, tail(words), subs);
    }
    for (auto& kv : b.subs) subs.emplace(kv.first, 0);
    return true;
  }

  int subscribe(const std::string& topic, const Subscriber& sub) override {  // QM:385-392
    Words words = java_split_dot(topic);
    if (words.empty()) return -1;
    for (int attempt = 0; attempt < 64; ++attempt) {
      int r = iinsert(root.get(), nullptr, words, sub);
      if (r == 1) return 0;
      if (r < 0) return r;
    }
    return -2;  // the reference recurses without bound here
  }
  int unsubscribe(const std::string& topic, const Subscriber& sub) override {  // QM:394-401
    Words words = java_split_dot(topic);
    if (words.empty()) return -1;
    for (int attempt = 0; attempt < 64; ++attempt) {
      int r = iremove(root.get(), nullptr, nullptr, words, 0, sub);
      if (r == 1) return 0;
      if (r < 0) return r;
    }
    return -2;
  }
  int lookup(const std::string& topic, SubSet& out) override {  // QM:403-413
    Words words = java_split_dot(topic);
    if (words.empty()) return -1;
    SubSet res;
    if (ilookup(root.get(), nullptr, words, res))
      for (auto& kv : res) out.emplace(kv.first, 0);
    return 0;
  }
};

// Brute-force restatement of what the trie computes: same word count, and every
// pattern word equals the key word or is "*".  Second opinion for the trie.
inline bool compat_match(const Words& pw, const Words& kw) {
  if (pw.size() != kw.size()) return false;
  for (size_t i = 0; i < pw.size(); ++i)
    if (pw[i] != kw[i] && pw[i] != "*") return false;
  return true;
}

}  // namespace orc
