// Host-side compilation of a headers exchange (SPEC mode): binding arguments -> atoms -> index.
//
// Semantics (AMQP 0-9-1 §3.1.3.4 as restated in SURVEY.md Appendix C; the reference has no headers
// matcher at all — "TODO header matcher ?" ExchangeEntity.scala:210): the binding's arguments table
// carries x-match = "all" (default) or "any"; every other argument whose name does not start with "x-"
// must (all) / may (any) be matched by a message header of the same name that either has the same
// typed value (same tag, same bytes) or — for a void argument — merely exists.  `all` with no
// arguments matches every message, `any` with none matches nothing.  Duplicate names in a table: the
// first one counts (ValueReader.scala:71-72).
#include <algorithm>
#include <cstring>
#include <map>
#include <unordered_map>

#include "../../include/chanamq_gpu.h"
#include "host_table.hpp"

namespace chq {

namespace {

struct Field {
  std::string_view name;
  uint8_t tag;
  std::string_view value;  // the value's bytes after the tag (for S/x/F/A: after the length word)
};

// Walks one field table body; false if malformed.  Nested tables/arrays are kept as opaque bytes.
bool parse_fields(std::string_view t, std::vector<Field>* out) {
  size_t p = 0;
  auto need = [&](size_t n) { return p + n <= t.size(); };
  while (p < t.size()) {
    if (!need(1)) return false;
    const size_t nl = (uint8_t)t[p++];
    if (!need(nl)) return false;
    std::string_view name = t.substr(p, nl);
    p += nl;
    if (!need(1)) return false;
    const uint8_t tag = (uint8_t)t[p++];
    size_t vl = 0;
    switch (tag) {
      case 'S': case 'x': case 'F': case 'A': {
        if (!need(4)) return false;
        vl = ((size_t)(uint8_t)t[p] << 24) | ((size_t)(uint8_t)t[p + 1] << 16) | ((size_t)(uint8_t)t[p + 2] << 8) | (uint8_t)t[p + 3];
        p += 4;
        break;
      }
      case 'I': case 'f': vl = 4; break;
      case 'D': vl = 5; break;
      case 'T': case 'd': case 'l': vl = 8; break;
      case 'b': case 't': vl = 1; break;
      case 's': vl = 2; break;
      case 'V': vl = 0; break;
      default: return false;
    }
    if (!need(vl)) return false;
    if (out) {
      bool dup = false;
      for (auto& f : *out) dup = dup || f.name == name;
      if (!dup) out->push_back(Field{name, tag, t.substr(p, vl)});
    }
    p += vl;
  }
  return true;
}

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

}  // namespace

bool valid_field_table(std::string_view t) { return parse_fields(t, nullptr); }

// hash of a value atom; the device computes the same from (name id, tag, value bytes)
uint64_t value_atom_hash(uint32_t name_id, uint8_t tag, std::string_view value) {
  return bytes_hash(kWordBasis ^ (((name_id << 8) | tag) * 0x9E3779B1u), value);
}

void HostTable::compile_headers(CompiledTable& out) const {
  out.headers_mode = true;
  out.wstride = 4;
  out.tags.assign(kBucketSlots, 0);
  out.slot_group.assign(kBucketSlots, kNone);
  out.groups.assign(1, Group{});
  std::unordered_map<std::string_view, uint32_t> qid;
  for (uint32_t i = 0; i < out.queue_names.size(); ++i) qid.emplace(out.queue_names[i], i);

  // pass 1: field names and atoms
  std::map<std::string_view, uint32_t> names;                                   // name -> id
  std::map<std::tuple<std::string_view, uint8_t, std::string_view>, uint32_t> value_atoms;
  std::map<std::string_view, uint32_t> presence_atoms;
  struct B { std::vector<Field> f; bool any; uint32_t queue; bool durable; };
  std::vector<B> bs;
  for (auto& t : live_) {
    B b;
    b.any = false;
    b.queue = qid.at(std::get<1>(t));
    b.durable = std::get<2>(t);
    std::vector<Field> all;
    parse_fields(std::get<3>(t), &all);  // validated at bind time
    for (auto& f : all) {
      if (f.name == "x-match") { b.any = f.tag == 'S' && f.value == "any"; continue; }
      if (f.name.size() >= 2 && f.name[0] == 'x' && f.name[1] == '-') continue;
      b.f.push_back(f);
      names.emplace(f.name, 0);
      if (f.tag == 'V') presence_atoms.emplace(f.name, 0);
      else value_atoms.emplace(std::make_tuple(f.name, f.tag, f.value), 0);
    }
    bs.push_back(std::move(b));
  }
  uint32_t id = 1;
  for (auto& kv : names) kv.second = id++;
  out.n_words = names.size();
  uint32_t atom = 1;
  for (auto& kv : presence_atoms) kv.second = atom++;
  for (auto& kv : value_atoms) kv.second = atom++;
  const uint32_t n_atoms = atom - 1;

  // pass 2: per binding its atom set; `any` -> per-atom queue sets, `all` -> groups by atom set
  std::vector<std::vector<uint32_t>> any_q(n_atoms + 1);
  std::vector<uint8_t> any_dur(n_atoms + 1, 0);
  std::map<std::vector<uint32_t>, std::pair<std::vector<uint32_t>, bool>> all_sets;
  std::vector<uint32_t> always;
  bool always_dur = false;
  for (auto& b : bs) {
    std::vector<uint32_t> as;
    for (auto& f : b.f)
      as.push_back(f.tag == 'V' ? presence_atoms.at(f.name) : value_atoms.at(std::make_tuple(f.name, f.tag, f.value)));
    std::sort(as.begin(), as.end());
    as.erase(std::unique(as.begin(), as.end()), as.end());
    if (b.any) {
      for (uint32_t a : as) { any_q[a].push_back(b.queue); any_dur[a] |= b.durable; }
    } else if (as.empty()) {
      always.push_back(b.queue);
      always_dur = always_dur || b.durable;
    } else {
      auto& g = all_sets[as];
      g.first.push_back(b.queue);
      g.second = g.second || b.durable;
    }
  }
  auto put_list = [&](std::vector<uint32_t>& q, uint32_t* off, uint32_t* len) {
    std::sort(q.begin(), q.end());
    q.erase(std::unique(q.begin(), q.end()), q.end());
    *off = (uint32_t)out.qlist.size();
    *len = (uint32_t)q.size();
    out.qlist.insert(out.qlist.end(), q.begin(), q.end());
  };
  out.atoms.assign(n_atoms + 1, AtomRec{});
  for (uint32_t a = 1; a <= n_atoms; ++a) {
    put_list(any_q[a], &out.atoms[a].any_off, &out.atoms[a].any_len);
    out.atoms[a].any_flags = (any_dur[a] ? kAnyDurable : 0) | kShared;
  }
  put_list(always, &out.always_off, &out.always_len);
  out.always_flags = (always_dur ? kAnyDurable : 0) | kShared;
  // anchor every `all` group at its rarest atom (fewest groups mention it): few candidates per message
  std::vector<uint32_t> freq(n_atoms + 1, 0);
  for (auto& kv : all_sets)
    for (uint32_t a : kv.first) freq[a]++;
  std::vector<std::vector<const std::vector<uint32_t>*>> anchored(n_atoms + 1);
  for (auto& kv : all_sets) {
    uint32_t best = kv.first[0];
    for (uint32_t a : kv.first)
      if (freq[a] < freq[best]) best = a;
    anchored[best].push_back(&kv.first);
  }
  for (uint32_t a = 1; a <= n_atoms; ++a) {
    out.atoms[a].all_begin = (uint32_t)out.all_groups.size();
    for (auto* set : anchored[a]) {
      auto& g = all_sets[*set];
      AllGroup ag{};
      ag.atoms_off = (uint32_t)out.all_atoms.size();
      ag.n_atoms = (uint32_t)set->size();
      out.all_atoms.insert(out.all_atoms.end(), set->begin(), set->end());
      uint64_t sig = 0;
      for (uint32_t at : *set) sig |= 1ull << (at & 63u);
      ag.sig_lo = (uint32_t)sig;
      ag.sig_hi = (uint32_t)(sig >> 32);
      put_list(g.first, &ag.list_off, &ag.list_len);
      ag.flags = (g.second ? kAnyDurable : 0) | kShared;
      out.all_groups.push_back(ag);
    }
    out.atoms[a].all_end = (uint32_t)out.all_groups.size();
  }
  if (out.all_groups.empty()) out.all_groups.push_back(AllGroup{});
  if (out.all_atoms.empty()) out.all_atoms.push_back(0);
  if (out.qlist.empty()) out.qlist.push_back(0);
  out.n_heads = n_atoms;

  // field-name dictionary (the word dictionary's layout and hash: 16-byte slots, one load per probe)
  auto dict_put = [](std::vector<DictSlot>& tab, const DictSlot& s) {
    uint32_t i = (uint32_t)s.hash & ((uint32_t)tab.size() - 1);
    while (tab[i].id != 0) i = (i + 1) & ((uint32_t)tab.size() - 1);
    tab[i] = s;
  };
  out.dict.assign(1, DictSlot{});
  {
    const uint32_t nd = pow2_at_least(std::max<uint64_t>(2, names.size() * 2));
    out.wdict.assign(nd, WordSlot{});
    out.wdict_off.assign(nd, 0);
    for (auto& kv : names) {
      const uint32_t h = bytes_hash(kWordBasis, kv.first);
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
  }
  if (out.dict_bytes.empty()) out.dict_bytes.push_back(0);
  out.presence_atom.assign(names.size() + 1, 0);
  for (auto& kv : presence_atoms) out.presence_atom[names.at(kv.first)] = kv.second;
  out.vdict.assign(pow2_at_least(std::max<uint64_t>(2, value_atoms.size() * 2)), DictSlot{});
  for (auto& kv : value_atoms) {
    const uint32_t name_id = names.at(std::get<0>(kv.first));
    const uint8_t tag = std::get<1>(kv.first);
    std::string_view v = std::get<2>(kv.first);
    DictSlot s{};
    s.hash = value_atom_hash(name_id, tag, v);
    s.first8 = first8(v);
    s.len = (uint32_t)v.size();
    s.id = kv.second;
    s.byte_off = (uint32_t)out.vdict_bytes.size();
    s.pad = (name_id << 8) | tag;
    if (v.size() > 8) out.vdict_bytes.insert(out.vdict_bytes.end(), v.begin(), v.end());
    dict_put(out.vdict, s);
  }
  if (out.vdict_bytes.empty()) out.vdict_bytes.push_back(0);
}

}  // namespace chq
