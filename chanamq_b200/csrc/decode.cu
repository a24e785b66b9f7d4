// AMQP frame scan + Basic.Publish / content-header decode on the device (K4/K5 of SURVEY.md §2b).
//
// Replaces, for many connection segments at once (base/ = chana-mq-base/src/main/scala/chana/mq/amqp/):
//   FrameParser.onReceive/process/parse      base/engine/FrameParser.scala:78-158
//   CommandAssembler.onReceive               base/engine/CommandAssembler.scala:56-130
//   AMQClass.readFrom -> Basic.Publish       base/method/Method.scala:14-32, Basic.scala:36-45
//   AMQContentHeader / BasicProperties       base/model/AMQContentHeader.scala:12-23, BasicProperties.scala:42-96
//   ValueReader tables                       base/model/ValueReader.scala:62-113
//   derived publish fields                   chana-mq-server/.../engine/FrameStage.scala:483-508
//
// A frame chain is a pointer chase (next = pos + 8 + size), so the scan is serial inside a segment and
// parallel across segments: k_frame_scan gives every connection segment one thread that walks its
// frames, runs the METHOD -> HEADER -> BODY* state machine with full validation, and appends the
// offset of each completed command to a scratch list.  k_publish_decode then parses all commands in
// parallel (one thread each) into fixed 64-byte records.  Bodies are never read: of a 4 KB message
// only ~100 bytes of frame headers, method arguments and properties are touched.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <string>

#include "kernels.h"

#include "../../include/chanamq_gpu.h"

namespace {

enum : int32_t { kErrFrame = 501 /* ErrorCodes.FRAME_ERROR */, kErrThrown = 1 /* the reference throws */ };
enum : uint8_t { F_MANDATORY = 1, F_IMMEDIATE = 2, F_PERSIST = 4, F_HAS_TTL = 8, F_HAS_HEADERS = 16, F_HAS_PROPS = 32 };

// big-endian reader over [pos, end); a read past the end sets eof (EOFException in the reference)
struct Rd {
  const uint8_t* b;
  uint64_t pos, end;
  bool eof;
  __device__ uint32_t u8() {
    if (pos >= end) { eof = true; return 0; }
    return __ldg(b + pos++);
  }
  __device__ uint32_t u16() { const uint32_t a = u8(); return (a << 8) | u8(); }
  __device__ uint32_t u32() { const uint32_t a = u16(); return (a << 16) | u16(); }
  __device__ uint64_t u64() { const uint64_t a = u32(); return (a << 32) | u32(); }
  __device__ void skip(uint64_t n) {
    if (pos + n > end) { eof = true; pos = end; } else pos += n;
  }
};

// Validates the fields of a table body occupying [rd.pos, lim) (ValueReader.readTable/readFieldValue/
// readArray, VR:62-113), nested containers included.  Returns false where the reference throws.
__device__ bool walk_table(Rd& rd, uint64_t lim) {
  constexpr int kDepth = 16;
  uint64_t ends[kDepth];
  uint8_t is_array[kDepth];
  int sp = 0;
  const uint64_t outer_end = rd.end;
  ends[0] = lim; is_array[0] = 0; sp = 1;
  bool ok = true;
  while (sp > 0 && ok) {
    rd.end = ends[sp - 1];
    if (rd.pos >= rd.end) { rd.pos = rd.end; --sp; continue; }
    if (!is_array[sp - 1]) { const uint32_t l = rd.u8(); rd.skip(l); }
    const uint32_t tag = rd.u8();
    switch (tag) {
      case 'S': case 'x': { const uint32_t n = rd.u32(); if (n >= 0x7FFFFFFFu) ok = false; else rd.skip(n); break; }
      case 'I': case 'f': rd.skip(4); break;
      case 'D': rd.skip(5); break;
      case 'T': case 'd': case 'l': rd.skip(8); break;
      case 'b': case 't': rd.skip(1); break;
      case 's': rd.skip(2); break;
      case 'V': break;
      case 'F': case 'A': {
        const uint32_t n = rd.u32();
        if (sp >= kDepth) { ok = false; break; }  // deeper nesting than any broker traffic: refused
        const uint64_t inner = min(rd.pos + (uint64_t)n, rd.end);
        ends[sp] = inner; is_array[sp] = tag == 'A'; ++sp;
        break;
      }
      default: ok = false;  // "Unrecognised type in table"
    }
    if (rd.eof) ok = false;
  }
  rd.end = outer_end;
  rd.pos = lim;
  return ok;
}

// Long.parseLong: optional sign, decimal ASCII digits, must fit (FrameStage.scala:497 `x.toLong`)
__device__ bool parse_long(const uint8_t* s, uint32_t n, int64_t* out) {
  if (n == 0) return false;
  uint32_t i = 0;
  bool neg = false;
  const uint8_t c0 = __ldg(s);
  if (c0 == '-' || c0 == '+') { neg = c0 == '-'; i = 1; if (n == 1) return false; }
  const uint64_t limit = neg ? (1ull << 63) : (1ull << 63) - 1;
  uint64_t v = 0;
  for (; i < n; ++i) {
    const uint8_t c = __ldg(s + i);
    if (c < '0' || c > '9') return false;
    const uint64_t d = c - '0';
    if (v > (limit - d) / 10) return false;
    v = v * 10 + d;
  }
  *out = neg ? (int64_t)(0 - v) : (int64_t)v;
  return true;
}

enum FrameStatus { kFrameOk = 0, kFrameIncomplete = 1, kFrameNegative = 2, kFrameBadEnd = 3 };
struct FrameHdr { uint32_t type, channel, size; };
__device__ FrameStatus read_frame(const uint8_t* b, uint64_t pos, uint64_t end, FrameHdr& f) {
  if (pos + 7 > end) return kFrameIncomplete;
  f.type = __ldg(b + pos);
  f.channel = (__ldg(b + pos + 1) << 8) | __ldg(b + pos + 2);
  f.size = (__ldg(b + pos + 3) << 24) | (__ldg(b + pos + 4) << 16) | (__ldg(b + pos + 5) << 8) | __ldg(b + pos + 6);
  if (f.size & 0x80000000u) return kFrameNegative;  // signed readInt (FP:177-186) -> NegativeArraySizeException (FP:135)
  if (pos + 8 + (uint64_t)f.size > end) return kFrameIncomplete;
  if (__ldg(b + pos + 7 + f.size) != 0xCE) return kFrameBadEnd;  // FP:149-154
  return kFrameOk;
}

__device__ bool known_class(uint32_t c) { return c == 10 || c == 20 || c == 30 || c == 40 || c == 50 || c == 60 || c == 90 || c == 85; }
__device__ bool has_content(uint32_t c, uint32_t m) { return c == 60 && (m == 40 || m == 50 || m == 60 || m == 71); }

enum CmdStatus { kCmdDone = 0, kCmdIncomplete = 1, kCmdError = 2 };
// Assembles ONE command whose METHOD frame starts at `pos` (already known to be a complete, well-ended,
// non-heartbeat frame).  On kCmdDone *next = position after the command's last frame and rec (if not
// null) is filled; on kCmdError *err / *err_off say why and where.
__device__ CmdStatus parse_command(const uint8_t* b, uint64_t pos, uint64_t seg_end, uint32_t seg, chq_publish_rec* out,
                                   uint64_t* next, int32_t* err, uint64_t* err_off) {
  chq_publish_rec rec;
  rec.method_off = pos; rec.body_off = 0; rec.body_size = 0; rec.ttl = 0;
  rec.exch_off = rec.exch_len = rec.key_off = rec.key_len = rec.hdr_tbl_off = rec.hdr_tbl_len = 0;
  rec.seg = seg; rec.channel = 0; rec.kind = 0; rec.flags = 0;
  FrameHdr f;
  read_frame(b, pos, seg_end, f);
  *err_off = pos;
  if (f.type != 1) { *err = kErrFrame; return kCmdError; }  // "expected type Frame.METHOD" (CA:90-91)
  rec.channel = (uint16_t)f.channel;
  Rd rd{b, pos + 7, pos + 7 + f.size, false};
  const uint32_t cls = rd.u16();
  if (rd.eof || !known_class(cls)) { *err = kErrThrown; return kCmdError; }  // EOF / UnknownClassOrMethodId
  const uint32_t mth = rd.u16();
  if (rd.eof) { *err = kErrThrown; return kCmdError; }
  if (cls == 60 && mth == 40) {  // Basic.Publish(reserved, exchange, routingKey, mandatory, immediate) (BA:45)
    rec.kind = 1;
    rd.u16();
    rec.exch_len = rd.u8(); rec.exch_off = (uint32_t)rd.pos; rd.skip(rec.exch_len);
    rec.key_len = rd.u8(); rec.key_off = (uint32_t)rd.pos; rd.skip(rec.key_len);
    const uint32_t bits = rd.u8();  // ArgumentsReader.readBit: one octet, bit 0 then bit 1 (AR:69-78)
    if (rd.eof) { *err = kErrThrown; return kCmdError; }
    if (bits & 1) rec.flags |= F_MANDATORY;
    if (bits & 2) rec.flags |= F_IMMEDIATE;
  } else {  // every other method is handed back undecoded
    rec.exch_off = (uint32_t)(pos + 7);
    rec.exch_len = f.size;
    rec.key_off = cls;
    rec.key_len = mth;
  }
  pos += 8 + (uint64_t)f.size;
  if (has_content(cls, mth)) {
    // ---- HEADER (heartbeats in between are skipped before the assembler, FS:301-303) ----
    for (;;) {
      const FrameStatus fs = read_frame(b, pos, seg_end, f);
      *err_off = pos;
      if (fs == kFrameIncomplete) return kCmdIncomplete;
      if (fs == kFrameNegative) { *err = kErrThrown; return kCmdError; }
      if (fs == kFrameBadEnd) { *err = kErrFrame; return kCmdError; }
      if (f.type != 8) break;
      pos += 8 + (uint64_t)f.size;
    }
    if (f.type != 2) { *err = kErrFrame; return kCmdError; }  // CA:106-107
    Rd hd{b, pos + 7, pos + 7 + f.size, false};
    if (hd.u16() != 60 || hd.eof) { *err = kErrThrown; return kCmdError; }  // CH:19-21
    hd.u16();                                                                // weight (BP:43)
    const int64_t body_size = (int64_t)hd.u64();
    const uint32_t fl = hd.u16();
    if (hd.eof || (fl & 1)) { *err = kErrThrown; return kCmdError; }        // EOF / continuation flag (PR:64-68)
    rec.flags |= F_HAS_PROPS;
    bool ok = true;
    auto shortstr = [&](uint32_t* off, uint32_t* len) {
      const uint32_t l = hd.u8();
      if (off) { *off = (uint32_t)hd.pos; *len = l; }
      hd.skip(l);
    };
    if (fl & 0x8000) shortstr(nullptr, nullptr);  // contentType
    if (fl & 0x4000) shortstr(nullptr, nullptr);  // contentEncoding
    if (fl & 0x2000) {                            // headers
      const uint32_t n = hd.u32();
      const uint64_t lim = min(hd.pos + (uint64_t)n, hd.end);
      rec.hdr_tbl_off = (uint32_t)hd.pos;
      rec.hdr_tbl_len = (uint32_t)(lim - hd.pos);
      rec.flags |= F_HAS_HEADERS;
      if (!hd.eof) ok = walk_table(hd, lim);
    }
    if (ok && (fl & 0x1000)) { if (hd.u8() == 2) rec.flags |= F_PERSIST; }  // deliveryMode (FS:483-489)
    if (ok && (fl & 0x0800)) hd.u8();                                        // priority
    if (ok && (fl & 0x0400)) shortstr(nullptr, nullptr);                     // correlationId
    if (ok && (fl & 0x0200)) shortstr(nullptr, nullptr);                     // replyTo
    if (ok && (fl & 0x0100)) {                                               // expiration (FS:491-504)
      uint32_t o = 0, l = 0;
      shortstr(&o, &l);
      int64_t v;
      if (!hd.eof && parse_long(b + o, l, &v)) { rec.ttl = v; rec.flags |= F_HAS_TTL; }
    }
    if (ok && (fl & 0x0080)) shortstr(nullptr, nullptr);                     // messageId
    if (ok && (fl & 0x0040)) hd.u64();                                       // timestamp
    if (ok && (fl & 0x0020)) shortstr(nullptr, nullptr);                     // type
    if (ok && (fl & 0x0010)) shortstr(nullptr, nullptr);                     // userId
    if (ok && (fl & 0x0008)) shortstr(nullptr, nullptr);                     // appId
    if (ok && (fl & 0x0004)) shortstr(nullptr, nullptr);                     // clusterId
    if (!ok || hd.eof) { *err = kErrThrown; return kCmdError; }
    pos += 8 + (uint64_t)f.size;
    // ---- BODY* until bodySize bytes have arrived; bodySize <= 0 completes at the header (CA:100-105) ----
    int64_t remaining = body_size;
    if (remaining > 0) rec.body_size = (uint64_t)body_size;
    while (remaining > 0) {
      const FrameStatus fs = read_frame(b, pos, seg_end, f);
      *err_off = pos;
      if (fs == kFrameIncomplete) return kCmdIncomplete;
      if (fs == kFrameNegative) { *err = kErrThrown; return kCmdError; }
      if (fs == kFrameBadEnd) { *err = kErrFrame; return kCmdError; }
      if (f.type == 8) { pos += 8 + (uint64_t)f.size; continue; }
      if (f.type != 3) { *err = kErrFrame; return kCmdError; }  // CA:127-128
      if (f.size > 0) {
        if (rec.body_off == 0) rec.body_off = pos + 7;
        remaining -= f.size;
        if (remaining < 0) { *err = kErrFrame; return kCmdError; }  // overshoot (CA:121-122)
      }
      pos += 8 + (uint64_t)f.size;
    }
  }
  *next = pos;
  if (out) *out = rec;
  return kCmdDone;
}

// scratch slot where segment s keeps its command offsets: a command takes at least 12 wire bytes
__device__ __host__ inline uint64_t seg_list_base(uint64_t seg_begin, uint32_t s) { return seg_begin / 12 + s; }

// K4: one thread per connection segment.
__global__ void k_frame_scan(const uint8_t* __restrict__ b, const uint64_t* __restrict__ seg_off, uint32_t n_seg,
                             uint32_t* __restrict__ cmd_off, uint64_t* __restrict__ seg_count,
                             uint64_t* __restrict__ seg_consumed, int32_t* __restrict__ seg_err,
                             uint64_t* __restrict__ seg_err_off) {
  const uint32_t s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_seg) return;
  const uint64_t begin = seg_off[s], end = seg_off[s + 1];
  uint32_t* list = cmd_off + seg_list_base(begin, s);
  uint64_t pos = begin, consumed = begin, count = 0, eoff = 0;
  int32_t err = 0;
  for (;;) {
    FrameHdr f;
    const FrameStatus fs = read_frame(b, pos, end, f);
    if (fs == kFrameIncomplete) break;
    if (fs == kFrameNegative) { err = kErrThrown; eoff = pos; break; }
    if (fs == kFrameBadEnd) { err = kErrFrame; eoff = pos; break; }
    if (f.type == 8) {  // heartbeat while idle (FS:301-303)
      pos += 8 + (uint64_t)f.size;
      consumed = pos;
      continue;
    }
    uint64_t next = pos;
    const CmdStatus cs = parse_command(b, pos, end, s, nullptr, &next, &err, &eoff);
    if (cs == kCmdIncomplete) break;
    if (cs == kCmdError) break;
    list[count++] = (uint32_t)pos;
    pos = next;
    consumed = pos;
  }
  seg_count[s] = count;
  seg_consumed[s] = consumed - begin;
  seg_err[s] = err;
  seg_err_off[s] = err ? eoff : 0;
}

// ------------------------------------------------------------------------------------------------
// Chunked scan: a long segment walked by many threads.
//
// The scan loop above visits a chain of IDLE positions (starts of METHOD frames or of heartbeats met between
// commands), each found from the one before.  Here a segment is cut into chunks of `chunk` bytes; the thread
// of chunk j > 0 GUESSES the first idle position at or after its chunk start (the first byte position from
// which a whole command parses and is followed by another plausible command), walks the chain from there to
// the end of its chunk, and notes where it came out.  k_chunk_validate then checks the chain exactly: chunk 0
// starts at the segment start; chunk j's walk stands only if it started where chunk j-1 really came out,
// otherwise that chunk is walked again from the true position (bodies full of frame look-alikes, frames longer
// than a chunk: slower, never wrong).  The result is the sequential scan's, bit for bit.
// ------------------------------------------------------------------------------------------------
constexpr uint64_t kNoPos = ~0ull;
struct ChunkRes {      // what one walk found
  uint64_t entry;      // idle position it started from (kNoPos: no guess found)
  uint64_t exit;       // first idle position at or after the chunk end, or where it stopped
  uint64_t eoff;       // error offset
  uint32_t count;      // commands started inside the chunk
  int32_t err;
  uint32_t stopped;    // 1: the chain ends here (incomplete tail or error)
  uint32_t pad;
};

// where chunk v keeps its command offsets: a command takes at least 12 wire bytes
__device__ __host__ inline uint64_t chunk_list_base(uint64_t chunk_start, uint32_t v) { return chunk_start / 12 + v; }

// Walks the idle positions from `pos` while they are below `limit`.
__device__ void walk_chunk(const uint8_t* __restrict__ b, uint64_t pos, uint64_t limit, uint64_t seg_end, uint32_t seg,
                           uint32_t* __restrict__ list, ChunkRes& r) {
  r.entry = pos; r.count = 0; r.err = 0; r.eoff = 0; r.stopped = 0; r.pad = 0;
  while (pos < limit) {
    FrameHdr f;
    const FrameStatus fs = read_frame(b, pos, seg_end, f);
    if (fs == kFrameIncomplete) { r.stopped = 1; break; }
    if (fs == kFrameNegative) { r.err = kErrThrown; r.eoff = pos; r.stopped = 1; break; }
    if (fs == kFrameBadEnd) { r.err = kErrFrame; r.eoff = pos; r.stopped = 1; break; }
    if (f.type == 8) { pos += 8 + (uint64_t)f.size; continue; }
    uint64_t next = pos;
    const CmdStatus cs = parse_command(b, pos, seg_end, seg, nullptr, &next, &r.err, &r.eoff);
    if (cs != kCmdDone) { if (cs == kCmdIncomplete) { r.err = 0; r.eoff = 0; } r.stopped = 1; break; }
    list[r.count++] = (uint32_t)pos;
    pos = next;
  }
  r.exit = pos;
}

// Could an idle position be at p?  0 no; else the position the chain would go to next (p itself if the command
// at p is still incomplete at the segment end).
__device__ uint64_t plausible_idle(const uint8_t* __restrict__ b, uint64_t p, uint64_t seg_end, uint32_t seg) {
  FrameHdr f;
  if (read_frame(b, p, seg_end, f) != kFrameOk) return 0;
  if (f.type == 8) return (f.size == 0 && f.channel == 0) ? p + 8 : 0;
  if (f.type != 1) return 0;
  uint64_t next = p, eoff;
  int32_t err;
  const CmdStatus cs = parse_command(b, p, seg_end, seg, nullptr, &next, &err, &eoff);
  if (cs == kCmdError) return 0;
  return cs == kCmdDone ? next : p;
}

// First position in [from, to) that looks like an idle position and is followed by another one; only a 0x01
// (METHOD) or a 0x08 (heartbeat) byte can start one.  By a whole warp: 512 bytes per step, 16 per lane; every lane tests the candidates among its own bytes in
// ascending order, and the lowest lane that found one has the first plausible position of the stripe.
__device__ uint64_t guess_idle_warp(const uint8_t* __restrict__ b, uint64_t from, uint64_t to, uint64_t seg_end, uint32_t seg,
                                    int lane) {
  auto test = [&](uint64_t p) -> bool {
    const uint64_t q = plausible_idle(b, p, seg_end, seg);
    if (q == 0) return false;
    if (q == p || q + 7 > seg_end) return true;
    return plausible_idle(b, q, seg_end, seg) != 0;
  };
  const uint64_t mis = (reinterpret_cast<uintptr_t>(b) + from) & 15u;
  for (uint64_t base = from - min(mis, from); base < to; base += 512) {
    const uint64_t lo = base + 16 * (uint64_t)lane;
    uint64_t found = kNoPos;
    if (lo < to) {
      uint32_t w[4] = {0, 0, 0, 0};
      if (lo >= from && lo + 16 <= to && ((reinterpret_cast<uintptr_t>(b) + lo) & 15u) == 0) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(b + lo));
        w[0] = v.x; w[1] = v.y; w[2] = v.z; w[3] = v.w;
      } else {
        for (int i = 0; i < 16; ++i)
          if (lo + i >= from && lo + i < to) w[i >> 2] |= (uint32_t)__ldg(b + lo + i) << (8 * (i & 3));
      }
      uint32_t any = 0;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t a = w[i] ^ 0x01010101u, c = w[i] ^ 0x08080808u;
        any |= ((a - 0x01010101u) & ~a & 0x80808080u) | ((c - 0x01010101u) & ~c & 0x80808080u);
      }
      if (any) {
        for (int i = 0; i < 16 && found == kNoPos; ++i) {
          const uint32_t c = (w[i >> 2] >> (8 * (i & 3))) & 0xFFu;
          if ((c == 1 || c == 8) && lo + i >= from && lo + i < to && test(lo + i)) found = lo + i;
        }
      }
    }
    const unsigned hit = __ballot_sync(0xFFFFFFFFu, found != kNoPos);
    if (hit) return __shfl_sync(0xFFFFFFFFu, found, __ffs(hit) - 1);
  }
  return kNoPos;
}

// Frame arithmetic of ONE command whose METHOD frame starts at pos (already known to be a complete, well-ended frame
// that is not a heartbeat): where the command ends if its frames are in order and complete.  Nothing inside the
// payloads is validated beyond what the arithmetic needs (class / method id, bodySize).  Returns false when anything is
// out of the ordinary — wrong frame type, incomplete or badly ended frame, short payload, body overshoot — and the
// caller then lets parse_command decide, so the outcome never depends on this shortcut.
__device__ bool command_end_light(const uint8_t* __restrict__ b, uint64_t pos, uint64_t seg_end, uint64_t* next) {
  FrameHdr f;
  if (read_frame(b, pos, seg_end, f) != kFrameOk || f.type != 1 || f.size < 4) return false;
  const uint32_t cls = (__ldg(b + pos + 7) << 8) | __ldg(b + pos + 8), mth = (__ldg(b + pos + 9) << 8) | __ldg(b + pos + 10);
  pos += 8 + (uint64_t)f.size;
  if (!has_content(cls, mth)) { *next = pos; return true; }
  for (;;) {
    if (read_frame(b, pos, seg_end, f) != kFrameOk) return false;
    if (f.type != 8) break;
    pos += 8 + (uint64_t)f.size;
  }
  if (f.type != 2 || f.size < 14) return false;
  uint64_t bs = 0;
  for (int i = 0; i < 8; ++i) bs = (bs << 8) | __ldg(b + pos + 7 + 4 + i);
  int64_t remaining = (int64_t)bs;
  pos += 8 + (uint64_t)f.size;
  while (remaining > 0) {
    if (read_frame(b, pos, seg_end, f) != kFrameOk) return false;
    if (f.type == 3) {
      remaining -= f.size;
      if (remaining < 0) return false;
    } else if (f.type != 8) return false;
    pos += 8 + (uint64_t)f.size;
  }
  *next = pos;
  return true;
}

// walk_chunk by a whole warp, with the same result.  Lane 0 runs ahead over up to 32 commands with the frame
// arithmetic alone; then every lane validates one of them with the full parse_command (the expensive part: method
// arguments, properties, field tables), all at once.  The first command that fails validation, or that the shortcut
// would not vouch for, ends the batch: lane 0 gives it to parse_command and the walk stops or goes on as it says.
__device__ void walk_chunk_warp(const uint8_t* __restrict__ b, uint64_t pos, uint64_t limit, uint64_t seg_end, uint32_t seg,
                                uint32_t* __restrict__ list, ChunkRes& r, int lane) {
  r.entry = pos; r.count = 0; r.err = 0; r.eoff = 0; r.stopped = 0; r.pad = 0;
  bool done = false;
  while (!done) {
    // ---- lane 0: up to 32 command starts ----
    uint32_t n = 0;
    uint64_t p = pos;
    bool unsure = false;   // the command at p needs the full parser (or is a frame-level stop)
    if (lane == 0) {
      while (n < 32 && p < limit) {
        FrameHdr f;
        const FrameStatus fs = read_frame(b, p, seg_end, f);
        if (fs != kFrameOk) { unsure = true; break; }
        if (f.type == 8) { p += 8 + (uint64_t)f.size; continue; }
        uint64_t nx;
        if (!command_end_light(b, p, seg_end, &nx)) { unsure = true; break; }
        list[r.count + n] = (uint32_t)p;
        ++n;
        p = nx;
      }
    }
    n = __shfl_sync(0xFFFFFFFFu, n, 0);
    p = __shfl_sync(0xFFFFFFFFu, p, 0);
    unsure = __shfl_sync(0xFFFFFFFFu, (int)unsure, 0) != 0;
    __syncwarp();
    // ---- every lane: full validation of one command ----
    bool bad = false;
    if ((uint32_t)lane < n) {
      uint64_t nx, eo;
      int32_t er;
      bad = parse_command(b, list[r.count + lane], seg_end, seg, nullptr, &nx, &er, &eo) != kCmdDone;
    }
    const unsigned bad_mask = __ballot_sync(0xFFFFFFFFu, bad);
    if (bad_mask) {  // the batch ends at its first invalid command, which lane 0 parses again below
      n = (uint32_t)__ffs(bad_mask) - 1;
      p = __shfl_sync(0xFFFFFFFFu, (uint64_t)list[r.count + n], 0);   // (lane 0 reads the list entry; all lanes get it)
      unsure = true;
    }
    r.count += n;
    pos = p;
    if (unsure) {  // the sequential step at pos, exactly as walk_chunk does it
      uint32_t stop = 0, adv = 0;
      if (lane == 0) {
        FrameHdr f;
        const FrameStatus fs = read_frame(b, pos, seg_end, f);
        if (fs == kFrameIncomplete) stop = 1;
        else if (fs == kFrameNegative) { r.err = kErrThrown; r.eoff = pos; stop = 1; }
        else if (fs == kFrameBadEnd) { r.err = kErrFrame; r.eoff = pos; stop = 1; }
        else if (f.type == 8) pos += 8 + (uint64_t)f.size;
        else {
          uint64_t nx = pos;
          const CmdStatus cs = parse_command(b, pos, seg_end, seg, nullptr, &nx, &r.err, &r.eoff);
          if (cs != kCmdDone) { if (cs == kCmdIncomplete) { r.err = 0; r.eoff = 0; } stop = 1; }
          else { list[r.count] = (uint32_t)pos; adv = 1; pos = nx; }
        }
      }
      stop = __shfl_sync(0xFFFFFFFFu, stop, 0);
      adv = __shfl_sync(0xFFFFFFFFu, adv, 0);
      pos = __shfl_sync(0xFFFFFFFFu, pos, 0);
      r.count += adv;
      if (stop) { r.stopped = 1; done = true; }
    }
    if (pos >= limit) done = true;
    __syncwarp();
  }
  r.exit = pos;
}

// chunks per segment -> chunk_base[n_seg + 1] (one block)
__global__ void k_chunk_plan(const uint64_t* __restrict__ seg_off, uint32_t n_seg, uint64_t chunk, uint32_t* __restrict__ chunk_base) {
  __shared__ uint32_t carry;
  __shared__ uint32_t warp_tot[32];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n_seg; base += blockDim.x) {
    const uint32_t i = base + threadIdx.x;
    uint32_t v = 0;
    if (i < n_seg) {
      const uint64_t len = seg_off[i + 1] - seg_off[i];
      v = (uint32_t)max((len + chunk - 1) / chunk, (uint64_t)1);
    }
    uint32_t incl = v;
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if ((threadIdx.x & 31) >= d) incl += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t wbase = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) wbase += warp_tot[w];
    if (i < n_seg) chunk_base[i] = carry + wbase + incl - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry += wbase + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) chunk_base[n_seg] = carry;
}

// one warp (its first lane) per chunk: guess the entry, walk
__global__ void k_chunk_walk(const uint8_t* __restrict__ b, const uint64_t* __restrict__ seg_off, uint32_t n_seg, uint64_t chunk,
                             const uint32_t* __restrict__ chunk_base, uint32_t max_chunks, uint32_t* __restrict__ cmd_off,
                             ChunkRes* __restrict__ res, uint32_t* __restrict__ chunk_seg, uint64_t* __restrict__ chunk_start,
                             uint64_t* __restrict__ chunk_count) {
  // One chunk per WARP: the walk is a serial, branchy byte parser, and 32 of them in one warp run one after the
  // other (measured with a thread per chunk: 6.5 active lanes per instruction, 3 % occupancy, 1.6 ms).  With a warp
  // each they run side by side, and inside the warp lane 0 follows the frame chain while all lanes validate the
  // commands it finds (walk_chunk_warp).
  const int lane = threadIdx.x & 31;
  const uint32_t v = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (v >= max_chunks) return;
  if (v >= chunk_base[n_seg]) {
    if (lane == 0) { chunk_seg[v] = n_seg - 1; chunk_start[v] = 0; chunk_count[v] = 0; }
    return;
  }
  uint32_t lo = 0, hi = n_seg;  // last segment whose first chunk <= v
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (chunk_base[mid] <= v) lo = mid; else hi = mid;
  }
  const uint32_t s = lo, j = v - chunk_base[s];
  const uint64_t begin = seg_off[s], end = seg_off[s + 1];
  const uint64_t cs = begin + (uint64_t)j * chunk, ce = min(cs + chunk, end);
  if (lane == 0) { chunk_seg[v] = s; chunk_start[v] = cs; }
  ChunkRes r;
  uint64_t entry = begin;
  if (j != 0) {
    entry = guess_idle_warp(b, cs, ce, end, s, lane);
  }
  if (entry == kNoPos) { r.entry = kNoPos; r.exit = kNoPos; r.eoff = 0; r.count = 0; r.err = 0; r.stopped = 0; r.pad = 0; }
  else walk_chunk_warp(b, entry, ce, end, s, cmd_off + chunk_list_base(cs, v), r, lane);
  if (lane == 0) res[v] = r;
}

// one warp per segment: the chain of chunk walks, checked 32 links at a time
__global__ void k_chunk_validate(const uint8_t* __restrict__ b, const uint64_t* __restrict__ seg_off, uint32_t n_seg, uint64_t chunk,
                                 const uint32_t* __restrict__ chunk_base, uint32_t* __restrict__ cmd_off, ChunkRes* __restrict__ res,
                                 uint64_t* __restrict__ chunk_count, uint64_t* __restrict__ seg_consumed,
                                 int32_t* __restrict__ seg_err, uint64_t* __restrict__ seg_err_off) {
  const uint32_t s = blockIdx.x;
  const int lane = threadIdx.x;
  const uint32_t c_lo = chunk_base[s], c_hi = chunk_base[s + 1];
  const uint64_t begin = seg_off[s], end = seg_off[s + 1];
  uint64_t entry = begin, eoff = 0;
  int32_t err = 0;
  bool stopped = false;
  uint32_t c = c_lo;
  while (c < c_hi) {
    if (stopped) {  // nothing after the point where the chain ended counts
      for (uint32_t t = c + lane; t < c_hi; t += 32) chunk_count[t] = 0;
      break;
    }
    const uint32_t t = c + lane;
    const bool have = t < c_hi;
    ChunkRes r;
    r.entry = kNoPos; r.exit = kNoPos; r.stopped = 0; r.count = 0; r.err = 0; r.eoff = 0;
    if (have) r = res[t];
    uint64_t prev_exit = __shfl_up_sync(0xFFFFFFFFu, r.exit, 1);
    uint32_t prev_stop = __shfl_up_sync(0xFFFFFFFFu, r.stopped, 1);
    if (lane == 0) { prev_exit = entry; prev_stop = 0; }
    const bool ok = have && r.entry != kNoPos && r.entry == prev_exit && !prev_stop;
    const unsigned good = __ballot_sync(0xFFFFFFFFu, ok);
    const int f = (~good) ? __ffs(~good) - 1 : 32;  // the links before lane f hold
    if (lane < f) chunk_count[t] = r.count;
    if (f > 0) {
      entry = __shfl_sync(0xFFFFFFFFu, r.exit, f - 1);
      if (__shfl_sync(0xFFFFFFFFu, r.stopped, f - 1)) {
        stopped = true;
        err = __shfl_sync(0xFFFFFFFFu, r.err, f - 1);
        eoff = __shfl_sync(0xFFFFFFFFu, r.eoff, f - 1);
      }
    }
    c += f;
    if (stopped || c >= c_hi || f == 32) continue;
    // chunk c was walked from the wrong place (or not at all): do it again from where the chain really is
    const uint64_t cs = begin + (uint64_t)(c - c_lo) * chunk, ce = min(cs + chunk, end);
    ChunkRes fix;
    fix.exit = entry; fix.stopped = 0; fix.err = 0; fix.eoff = 0; fix.count = 0;
    if (lane == 0) {
      if (entry < ce) {
        walk_chunk(b, entry, ce, end, s, cmd_off + chunk_list_base(cs, c), fix);
        res[c] = fix;
      }  // else: a frame spans the whole chunk, no idle position inside it
      chunk_count[c] = fix.count;
    }
    entry = __shfl_sync(0xFFFFFFFFu, fix.exit, 0);
    if (__shfl_sync(0xFFFFFFFFu, fix.stopped, 0)) {
      stopped = true;
      err = __shfl_sync(0xFFFFFFFFu, fix.err, 0);
      eoff = __shfl_sync(0xFFFFFFFFu, fix.eoff, 0);
    }
    ++c;
  }
  if (lane == 0) {
    seg_consumed[s] = entry - begin;
    seg_err[s] = err;
    seg_err_off[s] = err ? eoff : 0;
  }
}

// exclusive prefix sum of the per-segment command counts (n_seg is small: one block)
__global__ void k_seg_scan(const uint64_t* __restrict__ seg_count, uint32_t n_seg, uint64_t* __restrict__ rec_base,
                           uint64_t* __restrict__ n_recs, uint64_t recs_cap, int32_t* __restrict__ status) {
  __shared__ uint64_t carry;
  __shared__ uint64_t warp_tot[32];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < n_seg; base += blockDim.x) {
    const uint32_t i = base + threadIdx.x;
    const uint64_t v = i < n_seg ? seg_count[i] : 0;
    uint64_t incl = v;
    for (int d = 1; d < 32; d <<= 1) {
      const uint64_t t = __shfl_up_sync(0xFFFFFFFFu, incl, d);
      if ((threadIdx.x & 31) >= d) incl += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint64_t wbase = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); ++w) wbase += warp_tot[w];
    if (i < n_seg) rec_base[i] = carry + wbase + incl - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry += wbase + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    rec_base[n_seg] = carry;
    *n_recs = carry;
    *status = carry > recs_cap ? CHQ_E_OVERFLOW : CHQ_OK;
  }
}

// K5: one thread per command (grid-stride: the count lives on the device).  The command lists are per
// segment (chunk_seg == null) or per chunk of a segment (n_seg = number of chunks, chunk_seg / chunk_start
// say which segment a chunk belongs to and where it starts).
__global__ void k_publish_decode(const uint8_t* __restrict__ b, const uint64_t* __restrict__ seg_off, uint32_t n_seg,
                                 const uint32_t* __restrict__ cmd_off, const uint64_t* __restrict__ rec_base,
                                 chq_publish_rec* __restrict__ recs, uint64_t recs_cap, const uint32_t* __restrict__ chunk_seg,
                                 const uint64_t* __restrict__ chunk_start) {
  const uint64_t total = min(rec_base[n_seg], recs_cap);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    uint32_t lo = 0, hi = n_seg;  // last list whose base <= i
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (rec_base[mid] <= i) lo = mid; else hi = mid;
    }
    const uint32_t s = chunk_seg ? chunk_seg[lo] : lo;
    const uint64_t base = chunk_seg ? chunk_list_base(chunk_start[lo], lo) : seg_list_base(seg_off[s], s);
    const uint64_t pos = cmd_off[base + (i - rec_base[lo])];  // the batch is < 4 GiB
    uint64_t next, eoff;
    int32_t err;
    chq_publish_rec rec;
    parse_command(b, pos, seg_off[s + 1], s, &rec, &next, &err, &eoff);  // validated by the scan already
    recs[i] = rec;
  }
}

// extracts (offset, length) of the routing keys of the publish records (kind 1), compacted, for the router
__global__ void k_publish_keys(const chq_publish_rec* __restrict__ recs, const uint64_t* __restrict__ n_recs, uint64_t recs_cap,
                               uint32_t* __restrict__ key_off, uint32_t* __restrict__ key_len, uint32_t* __restrict__ rec_index,
                               uint32_t* __restrict__ n_pub) {
  const uint64_t total = min(*n_recs, recs_cap);
  for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < total; i += (uint64_t)gridDim.x * blockDim.x) {
    const bool pub = recs[i].kind == 1;
    const unsigned m = __ballot_sync(__activemask(), pub);
    if (!pub) continue;
    const unsigned lane = threadIdx.x & 31;
    uint32_t base = 0;
    const int leader = __ffs(m) - 1;
    if ((int)lane == leader) base = atomicAdd(n_pub, (uint32_t)__popc(m));
    base = __shfl_sync(m, base, leader);
    const uint32_t j = base + __popc(m & ((1u << lane) - 1));
    key_off[j] = recs[i].key_off;
    key_len[j] = recs[i].key_len;
    rec_index[j] = (uint32_t)i;
  }
}

thread_local std::string g_dec_err;
int dfail(int code, const std::string& m) { g_dec_err = m; return code; }
#define DCU(call)                                                                                              \
  do {                                                                                                         \
    cudaError_t e__ = (call);                                                                                  \
    if (e__ != cudaSuccess) return dfail(CHQ_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e__));     \
  } while (0)

struct DecodeScratch {  // per device; grown on demand
  uint32_t* cmd_off = nullptr; size_t cmd_cap = 0;      // command offsets found by the scan
  uint32_t* chunk_base = nullptr; size_t cbase_cap = 0; // chunked scan: first chunk of every segment
  ChunkRes* chunk_res = nullptr; size_t cres_cap = 0;
  uint32_t* chunk_seg = nullptr; size_t cseg_cap = 0;
  uint64_t* chunk_start = nullptr; size_t cstart_cap = 0;
  uint64_t* seg_count = nullptr; size_t count_cap = 0;
  uint64_t* rec_base = nullptr; size_t base_cap = 0;
  // host-buffer entry point
  uint8_t* d_bytes = nullptr; size_t bytes_cap = 0;
  uint64_t* d_seg = nullptr; size_t seg_cap = 0;
  chq_publish_rec* d_recs = nullptr; size_t recs_cap = 0;
};
DecodeScratch g_scratch[64];
// One scratch per GPU.  chq_decode (host buffers) holds the lock for its whole call, so concurrent broker threads
// are simply served one after the other; chq_decode_device holds it while it grows the scratch and queues its
// launches — the kernels of two device-level calls on different streams would still share the scratch, so those
// calls must not overlap on one GPU (include/chanamq_gpu.h).
std::recursive_mutex g_scratch_mu[64];

template <typename T>
int grow(T** p, size_t* cap, size_t need) {
  if (need <= *cap) return CHQ_OK;
  if (*p) DCU(cudaFree(*p));
  *p = nullptr; *cap = 0;
  const size_t want = need + need / 8 + 64;
  DCU(cudaMalloc(p, want * sizeof(T)));
  *cap = want;
  return CHQ_OK;
}

}  // namespace

extern "C" {

const char* chq_decode_last_error(void) { return g_dec_err.c_str(); }

int32_t chq_decode_device(const uint8_t* d_bytes, uint64_t len, const uint64_t* d_seg_offsets, uint32_t n_seg,
                          chq_publish_rec* d_recs, uint64_t recs_cap, uint64_t* d_n_recs, uint64_t* d_seg_consumed,
                          int32_t* d_seg_err, uint64_t* d_seg_err_off, int32_t* d_status, void* stream) {
  if (!d_bytes || !d_seg_offsets || !d_recs || !d_n_recs || !d_seg_consumed || !d_seg_err || !d_seg_err_off || !d_status)
    return dfail(CHQ_E_ARG, "chq_decode_device: bad argument");
  if (len >= (1ull << 32)) return dfail(CHQ_E_ARG, "chq_decode_device: at most 4 GiB per call (record offsets are 32 bit)");
  int dev = 0;
  DCU(cudaGetDevice(&dev));
  DecodeScratch& sc = g_scratch[dev & 63];
  std::lock_guard<std::recursive_mutex> lock(g_scratch_mu[dev & 63]);
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  // Segments are cut into chunks of this many bytes, one scanning thread each (CHQ_DECODE_CHUNK overrides the
  // size; 0 = the plain one-thread-per-segment scan).
  const uint64_t chunk = [] {
    const char* e = getenv("CHQ_DECODE_CHUNK");
    const long long v = e ? atoll(e) : (256ll << 10);  // measured best of 16 KB .. 1 MB on B200 (one warp per chunk)
    return (uint64_t)(v <= 0 ? 0 : (v < 64 ? 64 : v));
  }();
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (n_seg == 0) {
    DCU(cudaMemsetAsync(d_n_recs, 0, sizeof(uint64_t), st));
    DCU(cudaMemsetAsync(d_status, 0, sizeof(int32_t), st));
    return CHQ_OK;
  }
  if (chunk == 0) {
    if ((rc = grow(&sc.cmd_off, &sc.cmd_cap, (size_t)(len / 12 + n_seg + 1)))) return rc;
    if ((rc = grow(&sc.seg_count, &sc.count_cap, (size_t)n_seg + 1))) return rc;
    if ((rc = grow(&sc.rec_base, &sc.base_cap, (size_t)n_seg + 1))) return rc;
    k_frame_scan<<<(n_seg + 31) / 32, 32, 0, st>>>(d_bytes, d_seg_offsets, n_seg, sc.cmd_off, sc.seg_count, d_seg_consumed,
                                                    d_seg_err, d_seg_err_off);
    k_seg_scan<<<1, 1024, 0, st>>>(sc.seg_count, n_seg, sc.rec_base, d_n_recs, recs_cap, d_status);
    k_publish_decode<<<sms * 8, 256, 0, st>>>(d_bytes, d_seg_offsets, n_seg, sc.cmd_off, sc.rec_base, d_recs, recs_cap, nullptr,
                                              nullptr);
    chq::note_launches(3);
    DCU(cudaGetLastError());
    return CHQ_OK;
  }
  // the segment boundaries live on the device: size everything for the most chunks they can make
  const uint64_t max_chunks64 = len / chunk + n_seg;
  if (max_chunks64 >= (1ull << 31)) return dfail(CHQ_E_ARG, "chq_decode_device: too many chunks (raise CHQ_DECODE_CHUNK)");
  const uint32_t max_chunks = (uint32_t)max_chunks64;
  if ((rc = grow(&sc.cmd_off, &sc.cmd_cap, (size_t)(len / 12 + max_chunks + 2)))) return rc;
  if ((rc = grow(&sc.seg_count, &sc.count_cap, (size_t)max_chunks + 1))) return rc;
  if ((rc = grow(&sc.rec_base, &sc.base_cap, (size_t)max_chunks + 1))) return rc;
  if ((rc = grow(&sc.chunk_base, &sc.cbase_cap, (size_t)n_seg + 1))) return rc;
  if ((rc = grow(&sc.chunk_res, &sc.cres_cap, (size_t)max_chunks))) return rc;
  if ((rc = grow(&sc.chunk_seg, &sc.cseg_cap, (size_t)max_chunks))) return rc;
  if ((rc = grow(&sc.chunk_start, &sc.cstart_cap, (size_t)max_chunks))) return rc;
  k_chunk_plan<<<1, 1024, 0, st>>>(d_seg_offsets, n_seg, chunk, sc.chunk_base);
  k_chunk_walk<<<(max_chunks + 3) / 4, 128, 0, st>>>(d_bytes, d_seg_offsets, n_seg, chunk, sc.chunk_base, max_chunks, sc.cmd_off,
                                                      sc.chunk_res, sc.chunk_seg, sc.chunk_start, sc.seg_count);
  k_chunk_validate<<<n_seg, 32, 0, st>>>(d_bytes, d_seg_offsets, n_seg, chunk, sc.chunk_base, sc.cmd_off, sc.chunk_res, sc.seg_count,
                                         d_seg_consumed, d_seg_err, d_seg_err_off);
  k_seg_scan<<<1, 1024, 0, st>>>(sc.seg_count, max_chunks, sc.rec_base, d_n_recs, recs_cap, d_status);
  k_publish_decode<<<sms * 8, 256, 0, st>>>(d_bytes, d_seg_offsets, max_chunks, sc.cmd_off, sc.rec_base, d_recs, recs_cap,
                                            sc.chunk_seg, sc.chunk_start);
  chq::note_launches(5);
  DCU(cudaGetLastError());
  return CHQ_OK;
}

int32_t chq_publish_keys_device(const chq_publish_rec* d_recs, const uint64_t* d_n_recs, uint64_t recs_cap, uint32_t* d_key_off,
                                uint32_t* d_key_len, uint32_t* d_rec_index, uint32_t* d_n_pub, void* stream) {
  if (!d_recs || !d_n_recs || !d_key_off || !d_key_len || !d_rec_index || !d_n_pub)
    return dfail(CHQ_E_ARG, "chq_publish_keys_device: bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  int dev = 0, sms = 148;
  DCU(cudaGetDevice(&dev));
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  DCU(cudaMemsetAsync(d_n_pub, 0, sizeof(uint32_t), st));
  k_publish_keys<<<sms * 4, 256, 0, st>>>(d_recs, d_n_recs, recs_cap, d_key_off, d_key_len, d_rec_index, d_n_pub);
  chq::note_launches(1);
  DCU(cudaGetLastError());
  return CHQ_OK;
}

int32_t chq_decode(const uint8_t* bytes, uint64_t len, const uint64_t* seg_offsets, uint32_t n_seg, chq_publish_rec* recs,
                   uint64_t recs_cap, uint64_t* n_recs, uint64_t* seg_consumed, int32_t* seg_err, uint64_t* seg_err_off) {
  if (!bytes || !seg_offsets || !recs || !n_recs || !seg_consumed || !seg_err || !seg_err_off)
    return dfail(CHQ_E_ARG, "chq_decode: bad argument");
  int dev = 0;
  DCU(cudaGetDevice(&dev));  // fails loudly without a GPU: there is no CPU decoder in this library
  DecodeScratch& sc = g_scratch[dev & 63];
  std::lock_guard<std::recursive_mutex> lock(g_scratch_mu[dev & 63]);
  int rc;
  if ((rc = grow(&sc.d_bytes, &sc.bytes_cap, (size_t)len + 16))) return rc;
  // one block: seg offsets (n+1), consumed (n), err_off (n), err (n, in 64-bit slots), misc (2)
  if ((rc = grow(&sc.d_seg, &sc.seg_cap, (size_t)n_seg * 4 + 8))) return rc;
  if ((rc = grow(&sc.d_recs, &sc.recs_cap, (size_t)recs_cap + 1))) return rc;
  uint64_t* d_off = sc.d_seg;
  uint64_t* d_consumed = d_off + n_seg + 1;
  uint64_t* d_err_off = d_consumed + n_seg;
  int32_t* d_err = reinterpret_cast<int32_t*>(d_err_off + n_seg);
  uint64_t* d_misc = d_err_off + 2 * (size_t)n_seg + 2;  // n_recs, status
  cudaStream_t st = nullptr;
  DCU(cudaMemcpyAsync(sc.d_bytes, bytes, len, cudaMemcpyHostToDevice, st));
  DCU(cudaMemcpyAsync(d_off, seg_offsets, ((size_t)n_seg + 1) * 8, cudaMemcpyHostToDevice, st));
  rc = chq_decode_device(sc.d_bytes, len, d_off, n_seg, sc.d_recs, recs_cap, d_misc, d_consumed, d_err, d_err_off,
                         reinterpret_cast<int32_t*>(d_misc + 1), st);
  if (rc) return rc;
  uint64_t misc[2] = {0, 0};
  DCU(cudaMemcpyAsync(misc, d_misc, 16, cudaMemcpyDeviceToHost, st));
  DCU(cudaMemcpyAsync(seg_consumed, d_consumed, (size_t)n_seg * 8, cudaMemcpyDeviceToHost, st));
  DCU(cudaMemcpyAsync(seg_err_off, d_err_off, (size_t)n_seg * 8, cudaMemcpyDeviceToHost, st));
  DCU(cudaMemcpyAsync(seg_err, d_err, (size_t)n_seg * 4, cudaMemcpyDeviceToHost, st));
  DCU(cudaStreamSynchronize(st));
  *n_recs = misc[0];
  if ((int32_t)(uint32_t)misc[1] != CHQ_OK) return dfail(CHQ_E_OVERFLOW, "chq_decode: more commands than recs_cap");
  if (misc[0]) DCU(cudaMemcpy(recs, sc.d_recs, (size_t)misc[0] * sizeof(chq_publish_rec), cudaMemcpyDeviceToHost));
  return CHQ_OK;
}

}  // extern "C"
