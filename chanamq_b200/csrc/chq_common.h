// Shared between the host-side table compiler and the sm_100a kernels: the layout of one
// exchange's compiled binding table in HBM, and the hash functions both sides must agree on.
//
// What the table replaces: the reference's per-exchange TrieMatcher
// (chana-mq-server/.../engine/QueueMatcher.scala:380-601) — a pointer-chasing trie walked once
// per message.  Here the binding set is compiled into a TUPLE-SPACE index: every pattern is
// reduced to a shape (kind, word count, which positions are literal) plus the tuple of its
// literal word ids; one hash probe per (key, compatible shape) finds the group of bindings
// that share that pattern, and the group carries its sorted queue-id list.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define CHQ_HD __host__ __device__ __forceinline__
#else
#define CHQ_HD inline
#endif

namespace chq {

constexpr uint32_t kUnknownWord = 0;          // key word that no binding mentions
constexpr uint32_t kStar = 0xFFFFFFFEu;       // '*' inside a remainder
constexpr uint32_t kHash = 0xFFFFFFFDu;       // '#' inside a remainder (SPEC mode only)
constexpr uint32_t kNone = 0xFFFFFFFFu;
constexpr int kMaxIndexedWords = 16;          // a shape indexes at most the first 16 words
constexpr int kMaxKeyBytes = 255;             // AMQP shortstr (ValueReader.scala:35-39)
constexpr int kMaxWords = 256;                // 255 bytes hold at most 255 words (empty interior words count: "a....b")
constexpr int kLenCap = kMaxWords + 1;

enum ShapeKind : uint8_t {
  kExact = 0,    // no '#': key must have exactly `len` words
  kTail = 1,     // literal/'*' prefix of `len` words then only '#': key needs >= len words
  kComplex = 2,  // anything else: prefix indexed, remainder verified by the word-level NFA
};

struct Shape {            // 16 B
  uint64_t seed;          // per-shape hash seed
  uint32_t lit_mask;      // bit p set = word p of the prefix is a literal (else '*')
  uint8_t kind;
  uint8_t len;            // number of indexed prefix words (kExact: the whole pattern)
  uint16_t pad;
};

enum GroupFlags : uint8_t {
  kAnyDurable = 1,   // some binding's queue is durable (ExchangeEntity.scala:302)
  kShared = 2,       // some queue here is also bound through another group (union must de-duplicate)
  kHasRem = 4,       // a remainder has to be run through the word-level NFA
};

struct Group {            // 64 B; the first 32 B sector answers most probe hits on its own
  uint32_t list_off;      // into qlist: sorted unique queue ids bound with exactly this pattern
  uint32_t list_len;
  uint32_t next;          // next group with the same (shape, literal tuple) but another remainder
  uint16_t shape;
  uint8_t nlit;
  uint8_t flags;          // GroupFlags
  // zpos[p] = zob(p, literal word id at prefix position p) for p < 8, 0 where the shape has a '*'.
  // zob(p, .) is a bijection on 32 bits, so comparing it with the key's zob(p, word) IS comparing
  // the words — the key keeps those values in registers, which makes verification load-free.
  uint32_t zpos[8];
  uint32_t rem_off;       // into rems: pattern words after the indexed prefix (kHasRem)
  uint32_t rem_len;
  uint32_t lit_off;       // into lits: all literal word ids in position order (literals at positions >= 8)
  uint32_t pad;
};

// Hash index over the chain heads: buckets of eight slots.  tags[] holds one 16-bit tag per slot
// (0 = empty), so a probe is ONE 16-byte load; slot_group[] names the head group of a slot and is
// only read on a tag hit.  Built at most ~3/8 full; a bucket with an empty slot ends a probe.
constexpr int kBucketSlots = 8;

// Routing-key words (topic / direct tables): 16-byte slots so that a probe is one lane-load.  A word of
// at most 8 bytes is identified by (first8, len) alone; longer words also compare the hash bits and
// then the bytes kept in dict_bytes at wdict_off[slot].
struct WordSlot {
  uint64_t first8;        // first 8 bytes, little-endian packed, zero padded
  uint32_t id;            // 0 = empty slot
  uint32_t tag_len;       // word_hash & 0xFFFFFF00 | length in bytes (<= 255)
};

struct DictSlot {         // 32 B: one interned header field name / value atom (headers tables)
  uint64_t hash;
  uint64_t first8;        // first 8 bytes, little-endian packed (whole word when len <= 8)
  uint32_t len;
  uint32_t id;            // 0 = empty slot
  uint32_t byte_off;      // into dict_bytes, for words longer than 8 bytes
  uint32_t pad;
};

// ---- headers exchange (SPEC mode; AMQP 0-9-1 x-match all/any — the reference has none, ExchangeEntity.scala:210) ----
// An ATOM is one thing a binding argument can ask of a message: "field <name> is present" (a void
// argument) or "field <name> has exactly this typed value".  A message's headers table is turned into
// the set of atoms it satisfies; an `any` binding matches if one of its atoms is in the set, an `all`
// binding if every one of its atoms is.
struct AtomRec {          // 32 B
  uint32_t any_off, any_len, any_flags;   // queues of the `any` bindings that list this atom (sorted, unique)
  uint32_t all_begin, all_end;            // `all` groups anchored at this atom (their rarest atom)
  uint32_t pad[3];
};
struct AllGroup {         // 32 B: the `all` bindings with one and the same atom set
  uint32_t atoms_off, n_atoms;            // into all_atoms (ascending)
  uint32_t list_off, list_len, flags;
  uint32_t sig_lo, sig_hi;                // bit (atom & 63) set for every atom of the set: a message whose own
  uint32_t pad;                           // signature lacks one of these bits cannot satisfy the group
};
constexpr int kMaxHeaderFields = 64;      // known-name fields considered per message
constexpr int kHeaderListsPerMessage = 16; // descriptor arena = this many matched lists per message on average

// Device view of a committed table (all pointers are device pointers).
struct TableView {
  const DictSlot* dict;        // headers tables: field names
  const uint8_t* dict_bytes;
  uint32_t dict_mask;
  const WordSlot* wdict;       // topic / direct tables: pattern words
  const uint32_t* wdict_off;   // [slots] offset into dict_bytes of a word longer than 8 bytes
  uint32_t wdict_mask;
  const Shape* shapes;
  // [3*(kLenCap+1)]: for a key of L words, rng[3L], rng[3L+1] = the kExact shapes with len == L,
  // and [hashy_begin, rng[3L+2]) = the kTail/kComplex shapes (ascending len) with len <= L
  const uint16_t* exact_rng;
  uint32_t hashy_begin, hashy_end;
  const uint16_t* tags;        // [buckets * 8] 16-bit tags; a bucket is one aligned 16-byte load
  const uint32_t* slot_group;  // [buckets * 8]
  uint32_t bucket_mask;
  const Group* groups;
  const uint32_t* lits;
  const uint32_t* rems;
  const uint32_t* qlist;
  uint32_t wstride;            // word ids kept per key (multiple of 4)
  uint32_t n_groups;
  // direct / fanout exchanges reuse the same index: direct = one shape over the whole-key id,
  // fanout = a single always-matching group
  uint32_t exchange_type;
  uint32_t fanout_group;
  // headers exchange in SPEC mode (null otherwise); field names are interned in dict/dict_bytes
  const DictSlot* vdict;        // (name id, tag, value bytes) -> atom id; DictSlot.pad = name id << 8 | tag
  const uint8_t* vdict_bytes;
  uint32_t vdict_mask;
  const uint32_t* presence_atom;  // [names + 1]: atom "field present", 0 = nobody asks
  const AtomRec* atoms;           // [atoms + 1]
  const AllGroup* all_groups;
  const uint32_t* all_atoms;
  uint32_t always_off, always_len, always_flags;  // `all` bindings without arguments: match everything
};

// --- hashing -------------------------------------------------------------------------------
CHQ_HD uint64_t mix(uint64_t h, uint32_t w) {
  h = (h ^ w) * 0x9E3779B97F4A7C15ull;
  return h ^ (h >> 29);
}
CHQ_HD uint64_t fin(uint64_t h) {
  h ^= h >> 32;
  h *= 0xD6E8FEB86659FD93ull;
  h ^= h >> 32;
  return h;
}
CHQ_HD uint32_t tag_of(uint64_t h) { return (uint32_t)(h >> 32) | 1u; }
// Zobrist-style tuple hash (32 bit): hash(shape, literal tuple) = shape seed ^ XOR over the literal
// positions p of zob(p, word id at p).  A key computes zob() once per word, and every shape it is
// probed under then costs a few XORs.  The low bits pick the bucket, tag16() the slot tag; a tag
// hit is always confirmed against the stored tuple, so the hash only has to be fast.
CHQ_HD uint32_t zob(uint32_t p, uint32_t w) {
  uint32_t x = w ^ ((p + 1) * 0x9E3779B9u);
  x ^= x >> 16; x *= 0x85EBCA6Bu;
  x ^= x >> 13; x *= 0xC2B2AE35u;
  return x ^ (x >> 16);
}
CHQ_HD uint32_t tag16(uint32_t h) { return ((h * 0x9E3779B1u) >> 16) | 1u; }
// Hash of a routing-key word: folded over its bytes eight at a time (little-endian, zero padded; an
// empty word is one zero chunk), then the length.  The tokeniser reads keys as aligned 32-bit words, so
// a chunk is two registers.
constexpr uint32_t kWordBasis = 0x811C9DC5u;
CHQ_HD uint32_t word_chunk(uint32_t h, uint32_t lo, uint32_t hi) {
  h = (h ^ lo) * 0x9E3779B1u;
  h = (h ^ (h >> 15) ^ hi) * 0x85EBCA77u;
  return h ^ (h >> 13);
}
CHQ_HD uint32_t word_fin(uint32_t h, uint32_t len) {
  h = (h ^ len) * 0xC2B2AE3Du;
  return h ^ (h >> 16);
}
// FNV-1a, one byte at a time (header field names and values)
constexpr uint64_t kFnvBasis = 0xCBF29CE484222325ull;
constexpr uint64_t kFnvPrime = 0x100000001B3ull;
CHQ_HD uint64_t fnv_step(uint64_t h, uint8_t b) { return (h ^ b) * kFnvPrime; }

}  // namespace chq
