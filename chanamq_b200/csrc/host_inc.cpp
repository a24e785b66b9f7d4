// Incremental commit: single binds / unbinds applied to an already compiled table IN PLACE, so that a Queue.Bind on
// a large exchange costs microseconds of host work and a few small copies to the device instead of a recompile of
// the whole binding set.  The reference's bind is O(depth) and visible at once (TrieMatcher.subscribe -> iinsert,
// QueueMatcher.scala:385-392, 415-460, called per Queue.Bind at ExchangeEntity.scala:333-342; unsubscribe -> iremove,
// QueueMatcher.scala:394-401, 462-516).
//
// What can be done in place on a topic / direct table:
//   * a queue name the table has not seen: it takes the next free id (ids are then no longer name ranks — ascending id
//     stays the order of a result, but equals name order again only after the next full commit);
//   * a pattern word the table has not seen: one slot of the open-addressing dictionary;
//   * a pattern the table has not seen, of a shape it has: a new group record, chained behind the group with the same
//     literal tuple or entered into the tag index (first bucket with a free slot on the probe path, as at build time);
//   * one more queue on a known pattern: inserted into the group's sorted list if the list has room, else the list
//     moves to the end of qlist with twice the room;
//   * an unbind: the id leaves the list, the durable flag is recomputed from the bindings that remain.
// What needs the full compile (apply_pending returns false): a new SHAPE (shapes are ordered by length in the table and
// numbered inside every group record), a pattern that would widen the word-id rows, a tag index or dictionary more
// than half full, fan-out and headers tables, keys that are not in canonical form (trailing dots, "#.#").
#include <algorithm>
#include <cstring>

#include "../../include/chanamq_gpu.h"
#include "host_table.hpp"

namespace chq {

namespace {

uint32_t word_hash_of(std::string_view w) { return bytes_hash(kWordBasis, w); }
uint64_t first8_of(std::string_view w) {
  uint64_t f = 0;
  for (size_t i = 0; i < w.size() && i < 8; ++i) f |= (uint64_t)(unsigned char)w[i] << (8 * i);
  return f;
}
void touch(std::vector<Patch>& patches, int array, size_t off, size_t n) { patches.push_back(Patch{array, off, n}); }

// one key string <-> one pattern: no trailing dot (Java split drops trailing empty words), no "#.#" (collapses to "#")
bool canonical_key(std::string_view key, bool split, bool spec) {
  if (!split) return true;
  if (!key.empty() && key.back() == '.') return false;
  if (spec && key.find("#.#") != std::string_view::npos) return false;
  return true;
}

}  // namespace

bool HostTable::apply_pending(CompiledTable& ct, std::vector<Patch>& patches) const {
  if (!inc_.valid) return false;
  if (type_ == CHQ_FANOUT || (type_ == CHQ_HEADERS && mode_ == CHQ_MODE_SPEC)) return false;
  if (pending_.size() > 4096) return false;   // a bulk change: the full compile is the better tool
  for (const Op& op : pending_) {
    const bool ok = op.bind ? apply_bind(ct, patches, op) : apply_unbind(ct, patches, op);
    if (!ok) { inc_.valid = false; return false; }   // ct may be half patched: the caller compiles it afresh
  }
  pending_.clear();
  ct.n_bindings = live_.size();
  return true;
}

bool HostTable::apply_bind(CompiledTable& ct, std::vector<Patch>& patches, const Op& op) const {
  Inc& st = inc_;
  const bool split = (type_ == CHQ_TOPIC || type_ == CHQ_HEADERS);
  const bool spec = (mode_ == CHQ_MODE_SPEC) && split;
  if (!canonical_key(op.key, split, spec)) return false;
  // ---- the pattern, word by word (as HostTable::compile reads it) ----
  std::vector<std::string_view> w;
  if (split) java_split(op.key, w);
  else w.push_back(op.key);
  std::vector<uint32_t> pat;
  for (auto& x : w) {
    if (split && x.size() == 1 && x[0] == '*') { pat.push_back(kStar); continue; }
    if (spec && x.size() == 1 && x[0] == '#') {
      if (pat.empty() || pat.back() != kHash) pat.push_back(kHash);
      continue;
    }
    auto it = st.words.find(std::string(x));
    if (it == st.words.end()) {  // a word no pattern had: one dictionary slot
      const uint32_t nd = (uint32_t)ct.wdict.size();
      if ((st.words.size() + 1) * 2 > nd) return false;
      const uint32_t id = (uint32_t)st.words.size() + 1, h = word_hash_of(x);
      WordSlot s{};
      s.first8 = first8_of(x);
      s.id = id;
      s.tag_len = (h & 0xFFFFFF00u) | (uint32_t)x.size();
      uint32_t i = h & (nd - 1);
      while (ct.wdict[i].id != 0) i = (i + 1) & (nd - 1);
      ct.wdict[i] = s;
      touch(patches, kPaWdict, i, 1);
      if (x.size() > 8) {
        ct.wdict_off[i] = (uint32_t)ct.dict_bytes.size();
        touch(patches, kPaWdictOff, i, 1);
        touch(patches, kPaDictBytes, ct.dict_bytes.size(), x.size());
        ct.dict_bytes.insert(ct.dict_bytes.end(), x.begin(), x.end());
      }
      it = st.words.emplace(std::string(x), id).first;
      ct.n_words = st.words.size();
    }
    pat.push_back(it->second);
  }
  const size_t first_hash = std::find(pat.begin(), pat.end(), kHash) - pat.begin();
  const size_t ip = std::min<size_t>(first_hash, kMaxIndexedWords);
  uint8_t kind;
  if (first_hash == pat.size() && pat.size() <= (size_t)kMaxIndexedWords) kind = kExact;
  else if (first_hash < pat.size() && first_hash <= (size_t)kMaxIndexedWords && first_hash + 1 == pat.size()) kind = kTail;
  else kind = kComplex;
  if (ct.wstride != (uint32_t)kMaxWords && (kind == kComplex || ip > ct.wstride)) return false;   // the word rows would widen
  uint32_t mask = 0;
  for (size_t p = 0; p < ip; ++p)
    if (pat[p] != kStar) mask |= 1u << p;
  auto sit = st.shape_ids.find(std::make_tuple(kind, (uint8_t)ip, mask));
  if (sit == st.shape_ids.end()) return false;   // a shape the table has not seen
  const uint32_t shape_key = sit->second, shape = st.new_shape[shape_key];
  // ---- the queue ----
  uint32_t q;
  auto qit = st.qid.find(op.queue);
  if (qit == st.qid.end()) {
    q = (uint32_t)ct.queue_names.size();
    ct.queue_names.push_back(op.queue);
    st.qid.emplace(op.queue, q);
    st.queue_uses.push_back(0);
    st.first_group.push_back(kNone);
  } else {
    q = qit->second;
  }
  // ---- the group of this pattern ----
  std::vector<uint32_t> enc;
  enc.push_back(shape_key);
  for (size_t p = 0; p < ip; ++p)
    if (pat[p] != kStar) enc.push_back(pat[p]);
  const size_t n_lits = enc.size() - 1;
  enc.push_back(kNone);
  if (kind == kComplex) enc.insert(enc.end(), pat.begin() + ip, pat.end());
  ct.pattern_words += pat.size();
  uint32_t gi;
  auto git = st.group_ids.find(enc);
  if (git == st.group_ids.end()) {
    gi = (uint32_t)ct.groups.size();
    Group d{};
    d.shape = (uint16_t)shape;
    d.next = kNone;
    d.nlit = (uint8_t)n_lits;
    d.lit_off = (uint32_t)ct.lits.size();
    d.rem_off = (uint32_t)ct.rems.size();
    d.rem_len = (uint32_t)(enc.size() - 2 - n_lits);
    d.flags = d.rem_len ? kHasRem : 0;
    if (n_lits) { touch(patches, kPaLits, ct.lits.size(), n_lits); ct.lits.insert(ct.lits.end(), enc.begin() + 1, enc.begin() + 1 + n_lits); }
    if (d.rem_len) { touch(patches, kPaRems, ct.rems.size(), d.rem_len); ct.rems.insert(ct.rems.end(), enc.begin() + 2 + n_lits, enc.end()); }
    uint32_t h = (uint32_t)ct.shapes[shape].seed, li = 0;
    for (uint32_t mk = ct.shapes[shape].lit_mask; mk; mk &= mk - 1, ++li) {
      const uint32_t pos = (uint32_t)__builtin_ctz(mk);
      const uint32_t z = zob(pos, enc[1 + li]);
      if (pos < 8) d.zpos[pos] = z;
      h ^= z;
    }
    // an empty list with a little room, on a 16-byte boundary
    while (ct.qlist.size() & 3u) ct.qlist.push_back(kNone);
    d.list_off = (uint32_t)ct.qlist.size();
    d.list_len = 0;
    touch(patches, kPaQlist, ct.qlist.size(), 8);
    ct.qlist.insert(ct.qlist.end(), 8, kNone);
    st.list_cap.push_back(4);
    // chain behind the group with the same (shape, literal tuple), or a new entry of the tag index
    std::vector<uint32_t> hk;
    hk.push_back(shape);
    hk.insert(hk.end(), enc.begin() + 1, enc.begin() + 1 + n_lits);
    auto hit = st.heads.find(hk);
    if (hit != st.heads.end()) {
      d.next = ct.groups[hit->second].next;
      ct.groups[hit->second].next = gi;
      touch(patches, kPaGroups, hit->second, 1);
    } else {
      const uint32_t nb = (uint32_t)(ct.tags.size() / kBucketSlots);
      if ((ct.n_heads + 1) * 2 > (uint64_t)nb * kBucketSlots) return false;   // more than half full: rebuild it larger
      uint32_t b = h & (nb - 1);
      for (;;) {
        uint16_t* t = &ct.tags[(size_t)b * kBucketSlots];
        int i = 0;
        while (i < kBucketSlots && t[i] != 0) ++i;
        if (i < kBucketSlots) {
          t[i] = (uint16_t)tag16(h);
          ct.slot_group[(size_t)b * kBucketSlots + i] = gi;
          touch(patches, kPaTags, (size_t)b * kBucketSlots + i, 1);
          touch(patches, kPaSlotGroup, (size_t)b * kBucketSlots + i, 1);
          break;
        }
        b = (b + 1) & (nb - 1);
      }
      st.heads.emplace(std::move(hk), gi);
      ct.n_heads++;
    }
    ct.groups.push_back(d);
    st.group_ids.emplace(std::move(enc), gi);
  } else {
    gi = git->second;
  }
  // ---- the queue id into the group's sorted list ----
  Group& g = ct.groups[gi];
  uint32_t* list = ct.qlist.data() + g.list_off;
  uint32_t* at = std::lower_bound(list, list + g.list_len, q);
  if (at == list + g.list_len || *at != q) {
    const size_t ins = at - list;
    if (g.list_len < st.list_cap[gi]) {
      std::memmove(at + 1, at, (g.list_len - ins) * sizeof(uint32_t));
      *at = q;
      g.list_len++;
      touch(patches, kPaQlist, g.list_off + ins, g.list_len - ins);
    } else {  // no room: the list moves to the end of qlist with twice the room
      const uint32_t cap = std::max<uint32_t>(8, 2 * (g.list_len + 1)) & ~3u;
      std::vector<uint32_t> moved(list, list + g.list_len);
      moved.insert(moved.begin() + ins, q);
      while (ct.qlist.size() & 3u) ct.qlist.push_back(kNone);
      const uint32_t new_off = (uint32_t)ct.qlist.size();
      touch(patches, kPaQlist, new_off, cap + 4);
      ct.qlist.insert(ct.qlist.end(), moved.begin(), moved.end());
      ct.qlist.insert(ct.qlist.end(), cap + 4 - moved.size(), kNone);   // (+4: a bulk copy moves whole 16-byte units)
      Group& g2 = ct.groups[gi];
      g2.list_off = new_off;
      g2.list_len = (uint32_t)moved.size();
      st.list_cap[gi] = cap;
    }
    // a queue that now sits in two groups: their unions have to de-duplicate
    const uint32_t uses = ++st.queue_uses[q];
    if (uses == 1) st.first_group[q] = gi;
    else {
      ct.groups[gi].flags |= kShared;
      const uint32_t fg = st.first_group[q];
      if (uses == 2 && fg != kNone && fg != gi && !(ct.groups[fg].flags & kShared)) {
        ct.groups[fg].flags |= kShared;
        touch(patches, kPaGroups, fg, 1);
      }
    }
  }
  if (op.durable) ct.groups[gi].flags |= kAnyDurable;
  touch(patches, kPaGroups, gi, 1);
  return true;
}

bool HostTable::apply_unbind(CompiledTable& ct, std::vector<Patch>& patches, const Op& op) const {
  Inc& st = inc_;
  const bool split = (type_ == CHQ_TOPIC || type_ == CHQ_HEADERS);
  const bool spec = (mode_ == CHQ_MODE_SPEC) && split;
  if (!canonical_key(op.key, split, spec)) return false;
  // the group of this pattern (it exists: the binding was compiled or added in place)
  std::vector<std::string_view> w;
  if (split) java_split(op.key, w);
  else w.push_back(op.key);
  std::vector<uint32_t> pat;
  for (auto& x : w) {
    if (split && x.size() == 1 && x[0] == '*') { pat.push_back(kStar); continue; }
    if (spec && x.size() == 1 && x[0] == '#') {
      if (pat.empty() || pat.back() != kHash) pat.push_back(kHash);
      continue;
    }
    auto it = st.words.find(std::string(x));
    if (it == st.words.end()) return false;
    pat.push_back(it->second);
  }
  const size_t first_hash = std::find(pat.begin(), pat.end(), kHash) - pat.begin();
  const size_t ip = std::min<size_t>(first_hash, kMaxIndexedWords);
  uint8_t kind;
  if (first_hash == pat.size() && pat.size() <= (size_t)kMaxIndexedWords) kind = kExact;
  else if (first_hash < pat.size() && first_hash <= (size_t)kMaxIndexedWords && first_hash + 1 == pat.size()) kind = kTail;
  else kind = kComplex;
  uint32_t mask = 0;
  for (size_t p = 0; p < ip; ++p)
    if (pat[p] != kStar) mask |= 1u << p;
  auto sit = st.shape_ids.find(std::make_tuple(kind, (uint8_t)ip, mask));
  if (sit == st.shape_ids.end()) return false;
  std::vector<uint32_t> enc;
  enc.push_back(sit->second);
  for (size_t p = 0; p < ip; ++p)
    if (pat[p] != kStar) enc.push_back(pat[p]);
  enc.push_back(kNone);
  if (kind == kComplex) enc.insert(enc.end(), pat.begin() + ip, pat.end());
  auto git = st.group_ids.find(enc);
  auto qit = st.qid.find(op.queue);
  if (git == st.group_ids.end() || qit == st.qid.end()) return false;
  const uint32_t gi = git->second, q = qit->second;
  Group& g = ct.groups[gi];
  // what the bindings that remain under this key still say (the set keeps (key, queue, durable, args) tuples in key order)
  bool still_bound = false, any_durable = false;
  for (auto it = live_.lower_bound(std::make_tuple(op.key, std::string(), false, std::string()));
       it != live_.end() && std::get<0>(*it) == op.key; ++it) {
    if (std::get<1>(*it) == op.queue) still_bound = true;
    if (std::get<2>(*it)) any_durable = true;
  }
  if (!still_bound) {
    uint32_t* list = ct.qlist.data() + g.list_off;
    uint32_t* at = std::lower_bound(list, list + g.list_len, q);
    if (at != list + g.list_len && *at == q) {
      const size_t pos = at - list;
      std::memmove(at, at + 1, (g.list_len - pos - 1) * sizeof(uint32_t));
      g.list_len--;
      list[g.list_len] = kNone;
      touch(patches, kPaQlist, g.list_off + pos, g.list_len - pos + 1);
      if (st.queue_uses[q] > 0) st.queue_uses[q]--;   // (the groups it still sits in keep their de-duplication flag)
    }
  }
  g.flags = (uint8_t)((g.flags & ~kAnyDurable) | (any_durable ? kAnyDurable : 0));
  touch(patches, kPaGroups, gi, 1);
  return true;
}

}  // namespace chq
