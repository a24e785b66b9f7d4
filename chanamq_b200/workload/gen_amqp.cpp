// Synthetic AMQP 0-9-1 byte streams for the decode half of the hot path.
// The bytes follow what the reference's ENCODERS emit (there is no decode fixture in the reference):
//   frame      type u8 | channel u16 | size u32 | payload | 0xCE      base/model/Frame.scala:217-226
//   method     class u16 | method u16 | arguments                     base/method/Method.scala:58-67
//   publish    reserved u16 | exchange shortstr | routingKey shortstr | bits   base/method/Basic.scala:155-170
//   header     class u16 | weight u16 | bodySize u64 | flags u16 | properties  base/model/AMQContentHeader.scala:40-57,
//              base/model/ContentHeaderPropertyWriter.scala:31-53, BasicProperties.scala:121-153
//   table      u32 len | { name shortstr | tag | value }*                 base/model/ValueWriter.scala:46-97
// (base/ = chana-mq-base/src/main/scala/chana/mq/amqp/).  Every connection segment starts on a frame
// boundary and holds whole messages: METHOD(Basic.Publish) + HEADER + one BODY frame; 10 % of the
// messages also carry `expiration` and an 8-field `headers` table; 0.1 % heartbeats in between.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

extern "C" uint64_t chqgen_topic_keys(void* gp, uint64_t stream, uint32_t n, uint8_t* blob, uint32_t* off);

namespace {

struct Rng {
  uint64_t s;
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  uint32_t below(uint32_t n) { return (uint32_t)(((next() >> 32) * (uint64_t)n) >> 32); }
};

struct W {
  uint8_t* p;
  void u8(uint8_t v) { *p++ = v; }
  void u16(uint16_t v) { *p++ = (uint8_t)(v >> 8); *p++ = (uint8_t)v; }
  void u32(uint32_t v) { u16((uint16_t)(v >> 16)); u16((uint16_t)v); }
  void u64(uint64_t v) { u32((uint32_t)(v >> 32)); u32((uint32_t)v); }
  void bytes(const void* s, size_t n) { memcpy(p, s, n); p += n; }
  void shortstr(const char* s, size_t n) { u8((uint8_t)n); bytes(s, n); }
};

// begins a frame, returns where its size field lives
uint8_t* frame_begin(W& w, uint8_t type, uint16_t channel) {
  w.u8(type);
  w.u16(channel);
  uint8_t* size_at = w.p;
  w.u32(0);
  return size_at;
}
void frame_end(W& w, uint8_t* size_at) {
  uint32_t size = (uint32_t)(w.p - (size_at + 4));
  W s{size_at};
  s.u32(size);
  w.u8(0xCE);
}

}  // namespace

extern "C" {

// Upper bound of one message on the wire (for sizing the output buffer).
uint64_t chqgen_amqp_max_message(uint32_t body_bytes) { return 8 + 4 + 2 + 12 + 256 + 1 + 8 + 14 + 4 + 8 * 40 + 1 + 24 + 8 + body_bytes + 8; }

// Fills `out` (capacity cap) with n_seg connection segments of about target_bytes / n_seg bytes each;
// seg_offsets gets n_seg + 1 entries.  Routing keys come from the topic generator `topic_gen`
// (stream `stream`).  Returns the bytes written; *n_messages = publishes emitted.
uint64_t chqgen_amqp_stream(void* topic_gen, uint64_t stream, uint64_t target_bytes, uint32_t n_seg, uint32_t body_bytes,
                            uint8_t* out, uint64_t cap, uint64_t* seg_offsets, uint64_t* n_messages) {
  Rng r{0xA3B1C2D3E4F50617ull ^ (stream * 0x9E3779B97F4A7C15ull)};
  const uint64_t per_seg = target_bytes / (n_seg ? n_seg : 1);
  const uint64_t max_msg = chqgen_amqp_max_message(body_bytes);
  static const char kExchange[] = "bench.topic";
  // routing keys in blocks from the shared topic generator
  const uint32_t kBlock = 4096;
  std::string keys;
  keys.resize((size_t)kBlock * 64);
  uint32_t off[kBlock + 1];
  uint32_t have = 0, used = 0;
  uint64_t block_no = 0, msgs = 0;
  W w{out};
  for (uint32_t s = 0; s < n_seg; ++s) {
    seg_offsets[s] = (uint64_t)(w.p - out);
    const uint8_t* seg_start = w.p;
    while ((uint64_t)(w.p - seg_start) + max_msg <= per_seg && (uint64_t)(w.p - out) + max_msg <= cap) {
      if (used == have) {
        chqgen_topic_keys(topic_gen, stream * 1000003ull + block_no++, kBlock, (uint8_t*)&keys[0], off);
        have = kBlock;
        used = 0;
      }
      const char* key = &keys[off[used]];
      const uint32_t klen = off[used + 1] - off[used];
      ++used;
      const uint16_t channel = (uint16_t)(1 + r.below(4));
      if (r.below(1000) == 0) {  // heartbeat: type 8, channel 0, no payload (Frame.scala:58-64)
        uint8_t* sz = frame_begin(w, 8, 0);
        frame_end(w, sz);
      }
      // METHOD: Basic.Publish
      uint8_t* sz = frame_begin(w, 1, channel);
      w.u16(60); w.u16(40); w.u16(0);
      w.shortstr(kExchange, sizeof(kExchange) - 1);
      w.shortstr(key, klen);
      w.u8(0);
      frame_end(w, sz);
      // HEADER
      const bool rich = r.below(10) == 0;
      const uint8_t delivery = (uint8_t)(1 + r.below(2));
      sz = frame_begin(w, 2, channel);
      w.u16(60); w.u16(0); w.u64(body_bytes);
      w.u16(rich ? (0x2000 | 0x1000 | 0x0100) : 0x1000);
      if (rich) {  // headers table: 8 fields, mostly long strings
        uint8_t* tlen = w.p;
        w.u32(0);
        for (int f = 0; f < 8; ++f) {
          char name[8];
          int nl = snprintf(name, sizeof name, "h%02u", r.below(64));
          w.shortstr(name, (size_t)nl);
          const uint32_t kind = r.below(100);
          if (kind < 95) {
            const uint32_t vl = 4 + r.below(9);
            w.u8('S'); w.u32(vl);
            for (uint32_t i = 0; i < vl; ++i) w.u8((uint8_t)('a' + r.below(26)));
          } else if (kind < 97) { w.u8('I'); w.u32((uint32_t)r.next()); }
          else if (kind < 99) { w.u8('t'); w.u8((uint8_t)r.below(2)); }
          else { w.u8('l'); w.u64(r.next()); }
        }
        W t{tlen};
        t.u32((uint32_t)(w.p - (tlen + 4)));
      }
      w.u8(delivery);
      if (rich) {
        char exp[24];
        int el = snprintf(exp, sizeof exp, "%u", 1000 + r.below(600000));
        w.shortstr(exp, (size_t)el);
      }
      frame_end(w, sz);
      // BODY: one frame (frame-max 131072 in reference.conf:151 >> 4 KB)
      if (body_bytes) {
        sz = frame_begin(w, 3, channel);
        memset(w.p, (int)(msgs & 0x7F), body_bytes);
        w.p += body_bytes;
        frame_end(w, sz);
      }
      ++msgs;
    }
  }
  seg_offsets[n_seg] = (uint64_t)(w.p - out);
  if (n_messages) *n_messages = msgs;
  return (uint64_t)(w.p - out);
}

}  // extern "C"
