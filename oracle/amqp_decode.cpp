// ORACLE — TEST INFRASTRUCTURE ONLY (see trie_compat.hpp).
//
// C++ restatement of the reference's wire decode for the publish path.  base/ =
// chana-mq-base/src/main/scala/chana/mq/amqp/, FS = chana-mq-server/.../engine/FrameStage.scala:
//   FrameParser        base/engine/FrameParser.scala:68-194   (incremental: 7-byte header, payload, 0xCE)
//   CommandAssembler   base/engine/CommandAssembler.scala:44-131 (METHOD -> HEADER -> BODY*)
//   AMQClass.readFrom  base/method/Method.scala:14-32, Basic.readFrom case 40  base/method/Basic.scala:36-45
//   ArgumentsReader    base/method/ArgumentsReader.scala:27-97 (packed bits)
//   AMQContentHeader / BasicProperties.readFrom  base/model/AMQContentHeader.scala:12-23, BasicProperties.scala:42-96
//   ContentHeaderPropertyReader  base/model/ContentHeaderPropertyReader.scala:25-68
//   ValueReader (shortstr, table, field values)  base/model/ValueReader.scala:29-113
//   derived publish fields  FS:483-508;  heartbeat skip and error stop  FS:296-331
//
// Parity status: the reference has NO decode test; the only vector is the heartbeat frame
// 08 00 00 00 00 00 00 CE (base/model/Frame.scala:58-64).  Pinned by that vector plus hand-assembled
// frames for every field type (tests/test_oracle_decode.py); beyond that this is a restatement.
//
// Output record = what the CUDA decode emits (include/chanamq_gpu.h chq_publish_rec), so the two
// can be compared byte for byte.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace {

struct Rec {  // mirrors chq_publish_rec (64 bytes)
  uint64_t method_off, body_off, body_size;
  int64_t ttl;
  uint32_t exch_off, exch_len, key_off, key_len, hdr_tbl_off, hdr_tbl_len, seg;
  uint16_t channel;
  uint8_t kind, flags;
};
static_assert(sizeof(Rec) == 64, "record layout");
enum { F_MANDATORY = 1, F_IMMEDIATE = 2, F_PERSIST = 4, F_HAS_TTL = 8, F_HAS_HEADERS = 16, F_HAS_PROPS = 32 };
enum { ERR_FRAME = 501 /* ErrorCodes.FRAME_ERROR */, ERR_THROWN = 1 /* the reference throws an exception */ };

struct Thrown : std::runtime_error { using std::runtime_error::runtime_error; };

// java.io.DataInputStream over a byte range, big endian; reading past the end = EOFException
struct In {
  const uint8_t* base;  // batch start (for absolute offsets)
  size_t pos, end;
  uint8_t u8() { if (pos + 1 > end) throw Thrown("EOF"); return base[pos++]; }
  uint16_t u16() { uint16_t a = u8(); return (uint16_t)((a << 8) | u8()); }
  uint32_t u32() { uint32_t a = u16(); return (a << 16) | u16(); }
  uint64_t u64() { uint64_t a = u32(); return (a << 32) | u32(); }
  void skip(uint64_t n) { if (pos + n > end) throw Thrown("EOF"); pos += n; }
};

void read_field_value(In& in, int depth);
// ValueReader.readTable (VR:62-75): u32 length, then name/value pairs until the length is used up.
// The contents are parsed (so malformed tables throw, as in the reference) but not kept here.
void read_table(In& in, int depth, uint32_t* off, uint32_t* len) {
  uint32_t n = in.u32();
  if (off) { *off = (uint32_t)in.pos; *len = n; }
  // TruncatedInputStream.available() = min(limit - count, underlying.available()) (VR:153): a declared
  // length that overruns the payload simply ends the loop at the payload's end
  const size_t lim = std::min<uint64_t>((uint64_t)in.pos + n, in.end);
  if (len) *len = (uint32_t)(lim - in.pos);
  In t{in.base, in.pos, lim};
  while (t.pos < t.end) {
    uint8_t l = t.u8();
    t.skip(l);
    read_field_value(t, depth + 1);
  }
  in.pos = lim;
}
void read_field_value(In& in, int depth) {  // VR:90-113
  if (depth > 64) throw Thrown("nesting");
  switch (in.u8()) {
    case 'S': case 'x': { uint32_t n = in.u32(); if (n >= 0x7FFFFFFFu) throw Thrown("too long"); in.skip(n); break; }
    case 'I': in.skip(4); break;
    case 'D': in.skip(5); break;
    case 'T': in.skip(8); break;
    case 'F': read_table(in, depth, nullptr, nullptr); break;
    case 'A': {
      uint32_t n = in.u32();
      const size_t lim = std::min<uint64_t>((uint64_t)in.pos + n, in.end);
      In t{in.base, in.pos, lim};
      while (t.pos < t.end) read_field_value(t, depth + 1);
      in.pos = lim;
      break;
    }
    case 'b': case 't': in.skip(1); break;
    case 'd': case 'l': in.skip(8); break;
    case 'f': in.skip(4); break;
    case 's': in.skip(2); break;
    case 'V': break;
    default: throw Thrown("Unrecognised type in table");
  }
}

// Long.parseLong (FS:497 `x.toLong`): optional sign, at least one digit, decimal, must fit a long.
bool parse_long(const uint8_t* s, size_t n, int64_t* out) {
  if (n == 0) return false;
  size_t i = 0;
  bool neg = false;
  if (s[0] == '-' || s[0] == '+') { neg = s[0] == '-'; i = 1; if (n == 1) return false; }
  uint64_t limit = neg ? (uint64_t)1 << 63 : ((uint64_t)1 << 63) - 1, v = 0;
  for (; i < n; ++i) {
    if (s[i] < '0' || s[i] > '9') return false;  // ASCII digits only (non-ASCII digits: documented divergence)
    uint64_t d = s[i] - '0';
    if (v > (limit - d) / 10) return false;
    v = v * 10 + d;
  }
  *out = neg ? (int64_t)(0 - v) : (int64_t)v;
  return true;
}

bool has_content(uint16_t cls, uint16_t mth) {  // Method.hasContent: Basic.Publish/Return/Deliver/GetOk
  return cls == 60 && (mth == 40 || mth == 50 || mth == 60 || mth == 71);
}
bool known_class(uint16_t c) { return c == 10 || c == 20 || c == 30 || c == 40 || c == 50 || c == 60 || c == 90 || c == 85; }

struct Frame { uint8_t type; uint16_t channel; uint32_t size; uint64_t off; };  // off = position of the type byte

// CommandAssembler + the publish-field derivation of FS:483-508, one instance per connection (FS:251)
struct Assembler {
  enum { METHOD, HEADER, BODY } st = METHOD;
  Rec rec{};
  int64_t remaining = 0;
  // returns 1 when a command completed (rec valid), 0 to go on, ERR_FRAME on an assembler error
  int on_frame(const uint8_t* b, const Frame& f, uint32_t seg) {
    In in{b, (size_t)f.off + 7, (size_t)f.off + 7 + f.size};
    if (st == METHOD) {
      if (f.type != 1) return ERR_FRAME;  // "expected type Frame.METHOD" (CA:90-91)
      rec = Rec{};
      rec.method_off = f.off;
      rec.seg = seg;
      rec.channel = f.channel;
      uint16_t cls = in.u16();
      if (!known_class(cls)) throw Thrown("UnknownClassOrMethodId");
      uint16_t mth = in.u16();
      if (cls == 60 && mth == 40) {  // Basic.Publish (BA:45): reserved short, exchange, routingKey, bits
        rec.kind = 1;
        in.u16();
        rec.exch_len = in.u8(); rec.exch_off = (uint32_t)in.pos; in.skip(rec.exch_len);
        rec.key_len = in.u8(); rec.key_off = (uint32_t)in.pos; in.skip(rec.key_len);
        uint8_t bits = in.u8();  // AR:69-78: first readBit pulls one octet, bit 0 then bit 1
        if (bits & 1) rec.flags |= F_MANDATORY;
        if (bits & 2) rec.flags |= F_IMMEDIATE;
      } else {  // any other method: handed back undecoded (control plane, out of scope here)
        rec.kind = 0;
        rec.exch_off = (uint32_t)(f.off + 7);
        rec.exch_len = f.size;
        rec.key_off = cls;
        rec.key_len = mth;
      }
      if (has_content(cls, mth)) { st = HEADER; return 0; }
      return 1;
    }
    if (st == HEADER) {
      if (f.type != 2) return ERR_FRAME;  // CA:106-107
      uint16_t cls = in.u16();
      if (cls != 60) throw Thrown("UnknownClassOrMethodId");  // CH:19-21
      in.u16();                                               // weight, ignored (BP:43)
      int64_t body_size = (int64_t)in.u64();
      uint16_t flags = in.u16();                              // PR:44-62: one flag word for 14 properties
      if (flags & 1) throw Thrown("Unexpected continuation flag word");  // PR:64-68
      rec.flags |= F_HAS_PROPS;
      auto shortstr = [&](uint32_t* off, uint32_t* len) { uint8_t l = in.u8(); if (off) { *off = (uint32_t)in.pos; *len = l; } in.skip(l); };
      if (flags & 0x8000) shortstr(nullptr, nullptr);         // contentType
      if (flags & 0x4000) shortstr(nullptr, nullptr);         // contentEncoding
      if (flags & 0x2000) { read_table(in, 0, &rec.hdr_tbl_off, &rec.hdr_tbl_len); rec.flags |= F_HAS_HEADERS; }
      if (flags & 0x1000) { if (in.u8() == 2) rec.flags |= F_PERSIST; }   // deliveryMode (FS:483-489)
      if (flags & 0x0800) in.u8();                            // priority
      if (flags & 0x0400) shortstr(nullptr, nullptr);         // correlationId
      if (flags & 0x0200) shortstr(nullptr, nullptr);         // replyTo
      if (flags & 0x0100) {                                   // expiration (FS:491-504)
        uint32_t o = 0, l = 0;
        shortstr(&o, &l);
        int64_t v;
        if (parse_long(b + o, l, &v)) { rec.ttl = v; rec.flags |= F_HAS_TTL; }
      }
      if (flags & 0x0080) shortstr(nullptr, nullptr);         // messageId
      if (flags & 0x0040) in.u64();                           // timestamp
      if (flags & 0x0020) shortstr(nullptr, nullptr);         // type
      if (flags & 0x0010) shortstr(nullptr, nullptr);         // userId
      if (flags & 0x0008) shortstr(nullptr, nullptr);         // appId
      if (flags & 0x0004) shortstr(nullptr, nullptr);         // clusterId
      rec.body_size = 0;
      remaining = body_size;
      if (remaining > 0) { rec.body_size = (uint64_t)body_size; st = BODY; return 0; }
      st = METHOD;
      return 1;  // bodySize <= 0 completes at the header with an empty body (CA:100-105)
    }
    // BODY
    if (f.type != 3) return ERR_FRAME;  // CA:127-128
    if (f.size > 0) {
      if (rec.body_off == 0) rec.body_off = f.off + 7;
      remaining -= f.size;
      if (remaining > 0) return 0;
      if (remaining == 0) { st = METHOD; return 1; }
      return ERR_FRAME;  // overshoot (CA:121-122)
    }
    return 0;  // zero-length BODY frames are ignored (CA:124-126)
  }
};

// FrameParser, fed chunk by chunk; keeps the unread tail like `input = concat(input, newInput)` (FP:78-81)
struct Parser {
  std::vector<uint8_t> buf;  // unconsumed bytes
  uint64_t consumed = 0;     // absolute offset of buf[0] within the segment
  // appends a chunk, returns frames completed by it; err = ERR_FRAME on a bad end marker (FP:149-154)
  int feed(const uint8_t* p, size_t n, std::vector<Frame>& out, uint64_t seg_base) {
    buf.insert(buf.end(), p, p + n);
    size_t pos = 0;
    int err = 0;
    while (buf.size() - pos >= 7) {
      uint8_t type = buf[pos];
      uint16_t ch = (uint16_t)((buf[pos + 1] << 8) | buf[pos + 2]);
      uint32_t size = ((uint32_t)buf[pos + 3] << 24) | (buf[pos + 4] << 16) | (buf[pos + 5] << 8) | buf[pos + 6];
      if (size & 0x80000000u) { err = ERR_THROWN; break; }  // readInt is signed (FP:177-186): NegativeArraySizeException at FP:135
      if (buf.size() - pos < 7 + (size_t)size + 1) break;                  // ExpectData / ExpectEnd: wait for more
      if (buf[pos + 7 + size] != 0xCE) { err = ERR_FRAME; break; }
      out.push_back(Frame{type, ch, size, seg_base + consumed + pos});
      pos += 8 + (size_t)size;
    }
    buf.erase(buf.begin(), buf.begin() + pos);
    consumed += pos;
    return err;
  }
};

}  // namespace

extern "C" {

// Decodes n_seg connection segments (seg_offsets has n_seg+1 entries) the way FrameStage would, feeding
// each segment to its own FrameParser in chunks of `chunk` bytes (0 = the whole segment at once).
// recs (capacity recs_cap) receives the completed commands in segment order; seg_consumed = bytes of
// the segment up to the end of its last completed command (a stateless caller re-feeds the rest);
// seg_err = 0, 501 (frame/assembly error, FS:320-331) or 1 (the reference throws); seg_err_off = the
// offending frame.  Returns 0, or -1 if recs_cap is too small.
int orc_decode(const uint8_t* bytes, const uint64_t* seg_offsets, uint32_t n_seg, uint64_t chunk, void* recs_out,
               uint64_t recs_cap, uint64_t* n_recs, uint64_t* seg_consumed, int32_t* seg_err, uint64_t* seg_err_off) {
  Rec* recs = (Rec*)recs_out;
  uint64_t n = 0;
  for (uint32_t s = 0; s < n_seg; ++s) {
    const uint64_t b0 = seg_offsets[s], b1 = seg_offsets[s + 1];
    Parser parser;
    Assembler as;
    seg_consumed[s] = 0;
    seg_err[s] = 0;
    seg_err_off[s] = 0;
    uint64_t fed = b0;
    bool stop = false;
    while (fed < b1 && !stop) {
      uint64_t take = chunk ? std::min<uint64_t>(chunk, b1 - fed) : b1 - fed;
      std::vector<Frame> frames;
      int perr = 0;
      uint64_t last_off = 0;
      bool in_frame = false;  // an exception while a frame is being assembled points at that frame
      try {
        perr = parser.feed(bytes + fed, take, frames, b0);
        for (auto& f : frames) {
          last_off = f.off;
          in_frame = true;
          if (f.type == 8) {  // heartbeat: skipped before the assembler sees it (FS:301-303)
            if (as.st == Assembler::METHOD) seg_consumed[s] = f.off + 8 + f.size - b0;
            continue;
          }
          int r = as.on_frame(bytes, f, s);
          if (r == 1) {
            if (n >= recs_cap) return -1;
            recs[n++] = as.rec;
            seg_consumed[s] = f.off + 8 + f.size - b0;
            as = Assembler();  // "reset for next command" (FS:319)
          } else if (r != 0) {
            seg_err[s] = r;
            seg_err_off[s] = f.off;
            stop = true;
            break;
          }
        }
        if (!stop && perr) {
          seg_err[s] = perr;
          seg_err_off[s] = b0 + parser.consumed;
          stop = true;
        }
      } catch (Thrown&) {
        seg_err[s] = ERR_THROWN;
        seg_err_off[s] = in_frame ? last_off : b0 + parser.consumed;
        stop = true;
      }
      fed += take;
    }
  }
  *n_recs = n;
  return 0;
}

}  // extern "C"
