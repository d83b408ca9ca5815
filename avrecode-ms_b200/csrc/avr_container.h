// The recode.proto container on memory buffers (reference recode.proto:1-19; block sequence rules recode.cpp:1282-1304,
// 1129-1131; reader rules recode.cpp:1366-1416): writer, reader, and the decompress-side plan that turns a container back into
// the list of coded slices (header fields re-read from the literal bytes, picture flips, block matching as
// decompressor::init / find_next_coded_block do, recode.cpp:1534-1580).  Pure host code without CUDA: avr_files.cpp uses it
// around the device batches, tests/host_emul.cpp around the emulated kernel bodies.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "../../include/avrecode_b200.h"
#include "avr_record.h"

namespace avrc {

// A container under construction: malloc-backed so that the finished bytes become the caller's avr_buffer without a copy.
struct Buf {
  uint8_t* p = nullptr; size_t n = 0, cap = 0; bool bad = false;
  Buf() = default;
  Buf(const Buf&) = delete; Buf& operator=(const Buf&) = delete;
  Buf(Buf&& o) noexcept : p(o.p), n(o.n), cap(o.cap), bad(o.bad) { o.p = nullptr; o.n = o.cap = 0; }
  ~Buf() { free(p); }
  void reserve(size_t c) {
    if (c <= cap) return;
    uint8_t* q = (uint8_t*)realloc(p, c);
    if (!q) { bad = true; return; }
    p = q; cap = c;
  }
  bool room(size_t k) { if (n + k > cap) reserve((n + k) + (n + k) / 2 + 64); return n + k <= cap; }
  void push_back(uint8_t v) { if (room(1)) p[n++] = v; }
  void append(const uint8_t* s, size_t k) { if (k && room(k)) { memcpy(p + n, s, k); n += k; } }
  void clear() { n = 0; }
  size_t size() const { return n; }
};

// ---- proto2 wire writer (fields in tag order, set-but-empty fields are emitted) ----------------
inline void put_varint(Buf& o, uint64_t v) {
  while (v >= 0x80) { o.push_back((uint8_t)(v | 0x80)); v >>= 7; }
  o.push_back((uint8_t)v);
}
inline size_t varint_len(uint64_t v) { size_t n = 1; while (v >= 0x80) { v >>= 7; n++; } return n; }
inline void put_bytes_field(Buf& o, uint8_t tag, const uint8_t* p, size_t n) {
  o.push_back(tag); put_varint(o, n); o.append(p, n);
}
inline void put_block_literal(Buf& o, const uint8_t* p, size_t n) {
  o.push_back(0x12); put_varint(o, 1 + varint_len(n) + n);
  put_bytes_field(o, 0x12, p, n);
}
inline void put_block_skip(Buf& o, uint64_t size) {
  o.push_back(0x12); put_varint(o, 1 + varint_len(size) + 2);
  o.push_back(0x08); put_varint(o, size);
  o.push_back(0x18); o.push_back(0x01);
}
inline void put_block_cabac(Buf& o, uint64_t size, const uint8_t* cabac, size_t n, int parity, const uint8_t* last_byte) {
  size_t body = 1 + varint_len(size) + 1 + varint_len(n) + n + 2 + (last_byte ? 3 : 0);
  o.push_back(0x12); put_varint(o, body);
  o.push_back(0x08); put_varint(o, size);
  put_bytes_field(o, 0x22, cabac, n);
  o.push_back(0x28); o.push_back((uint8_t)parity);
  if (last_byte) { o.push_back(0x32); o.push_back(0x01); o.push_back(*last_byte); }
}

// ---- reader -----------------------------------------------------------------------------------
struct Block {
  bool has_size = false, has_literal = false, has_skip = false, has_cabac = false, has_parity = false, has_last = false;
  uint64_t size = 0; bool skip = false; int parity = 0;
  const uint8_t* literal = nullptr; size_t literal_len = 0;
  const uint8_t* cabac = nullptr; size_t cabac_len = 0;
  const uint8_t* last = nullptr; size_t last_len = 0;
};
inline bool get_varint(const uint8_t* p, size_t n, size_t& i, uint64_t* v) {
  uint64_t r = 0; int sh = 0;
  while (i < n && sh < 64) {
    uint8_t b = p[i++];
    r |= (uint64_t)(b & 0x7F) << sh; sh += 7;
    if (!(b & 0x80)) { *v = r; return true; }
  }
  return false;
}
inline bool skip_field(const uint8_t* p, size_t n, size_t& i, int wt) {
  uint64_t v;
  switch (wt) {
    case 0: return get_varint(p, n, i, &v);
    case 1: if (i + 8 > n) return false; i += 8; return true;
    case 2: if (!get_varint(p, n, i, &v) || v > n - i) return false; i += (size_t)v; return true;
    case 5: if (i + 4 > n) return false; i += 4; return true;
    default: return false;
  }
}
inline bool parse_container(const uint8_t* p, size_t n, std::vector<Block>* blocks) {
  size_t i = 0;
  while (i < n) {
    uint64_t tag, len;
    if (!get_varint(p, n, i, &tag)) return false;
    if ((tag >> 3) != 2 || (tag & 7) != 2) { if (!skip_field(p, n, i, (int)(tag & 7))) return false; continue; }
    if (!get_varint(p, n, i, &len) || len > n - i) return false;
    const uint8_t* b = p + i; size_t bn = (size_t)len, j = 0;
    i += bn;
    Block blk;
    while (j < bn) {
      uint64_t t, v;
      if (!get_varint(b, bn, j, &t)) return false;
      int f = (int)(t >> 3), wt = (int)(t & 7);
      if (wt == 0 && (f == 1 || f == 3 || f == 5)) {
        if (!get_varint(b, bn, j, &v)) return false;
        if (f == 1) { blk.has_size = true; blk.size = v; }
        else if (f == 3) { blk.has_skip = true; blk.skip = v != 0; }
        else { blk.has_parity = true; blk.parity = v != 0; }
      } else if (wt == 2 && (f == 2 || f == 4 || f == 6)) {
        if (!get_varint(b, bn, j, &v) || v > bn - j) return false;
        if (f == 2) { blk.has_literal = true; blk.literal = b + j; blk.literal_len = (size_t)v; }
        else if (f == 4) { blk.has_cabac = true; blk.cabac = b + j; blk.cabac_len = (size_t)v; }
        else { blk.has_last = true; blk.last = b + j; blk.last_len = (size_t)v; }
        j += (size_t)v;
      } else if (!skip_field(b, bn, j, wt)) return false;
    }
    blocks->push_back(blk);
  }
  return true;
}


struct DFile {
  std::vector<Block> blocks;
  std::vector<uint8_t> surrogate;             // literals + 'X' placeholders (recode.cpp:1534-1551 uses marker + 'X')
  std::vector<avr_dslice> slices;             // coded slices in decode order
  std::vector<int> slice_block;               // block index of each
  int rc = AVR_OK;
};

// One container -> blocks, surrogate stream and the coded slices in decode order (out->rc = AVR_OK or why not).
inline void plan_decompress(const uint8_t* in, size_t in_len, DFile* out) {
  DFile& d = *out;
  if (!parse_container(in, in_len, &d.blocks)) { d.rc = AVR_E_CORRUPT; return; }
  std::vector<size_t> coded_pos; std::vector<int> coded_blk;
  size_t total = 0;
  for (auto& b : d.blocks) total += b.has_literal ? b.literal_len : b.has_cabac ? (size_t)b.size : 0;
  d.surrogate.reserve(total);
  for (size_t i = 0; i < d.blocks.size(); i++) {
    const Block& b = d.blocks[i];
    if ((int)b.has_literal + (int)b.has_cabac + (int)b.has_skip != 1) { d.rc = AVR_E_CORRUPT; return; }   // recode.cpp:1371-1373
    if (b.has_literal) d.surrogate.insert(d.surrogate.end(), b.literal, b.literal + b.literal_len);
    else if (b.has_cabac) {
      if (!b.has_size || b.size > (1u << 30)) { d.rc = AVR_E_CORRUPT; return; }                          // recode.cpp:1384-1386
      coded_pos.push_back(d.surrogate.size()); coded_blk.push_back((int)i);
      d.surrogate.insert(d.surrogate.end(), (size_t)b.size, (uint8_t)'X');
    } else if (b.skip) { coded_pos.push_back((size_t)-1); coded_blk.push_back((int)i); }
    else { d.rc = AVR_E_CORRUPT; return; }
  }
  bool any_cabac = false;
  for (auto& b : d.blocks) any_cabac |= b.has_cabac;
  if (!any_cabac) return;
  size_t next = 0;
  bool have = false; int w = 0, h = 0, fnum = 0;
  int nflips = 0; bool last_prev_valid = false;
  int unsupported = 0, last_first_mb = -1;
  int rc = avr::scan_annexb_slices(d.surrogate.data(), d.surrogate.size(), [&](const avr::SliceLoc& loc) -> int {
    const avr::SliceHeaderInfo& sh = loc.sh;
    // update_frame_spec; every picture flips unless AVR_STRICT_DOMAIN; a slice also starts a picture when none has started yet
    // or the size changed under it or first_mb does not advance (avr_record.h starts_picture: the same rule on the compress side)
    const bool starts_picture = sh.sp.first_mb == 0 || !have || sh.sp.mb_w != w || sh.sp.mb_h != h || sh.sp.first_mb <= last_first_mb;
    last_first_mb = sh.sp.first_mb;
    if (starts_picture && !(avr::strict_domain() && have && sh.sp.mb_w == w && sh.sp.mb_h == h && sh.frame_num == fnum)) {
      last_prev_valid = have && sh.sp.mb_w == w && sh.sp.mb_h == h;
      nflips++;
      have = true; w = sh.sp.mb_w; h = sh.sp.mb_h; fnum = sh.frame_num;
    }
    if (next >= coded_blk.size()) return AVR_E_CORRUPT;
    const Block& b = d.blocks[coded_blk[next]];
    const size_t pos = coded_pos[next];
    next++;
    if (b.has_skip) return b.size == loc.size ? 0 : AVR_E_CORRUPT;              // recode.cpp:1573-1575
    if (pos != loc.raw_off || b.size != loc.size) return AVR_E_CORRUPT;          // recode.cpp:1564-1566
    avr_dslice s;
    memset(&s, 0, sizeof s);
    s.payload = b.cabac; s.payload_len = (uint32_t)b.cabac_len; s.size = (uint32_t)b.size;
    s.flip = nflips == 0 ? 0u : (nflips == 1 && last_prev_valid) ? 1u : 2u;
    nflips = 0;
    s.sp.slice_type = sh.sp.slice_type; s.sp.first_mb = sh.sp.first_mb; s.sp.mb_w = sh.sp.mb_w; s.sp.mb_h = sh.sp.mb_h;
    s.sp.num_ref[0] = sh.sp.num_ref[0]; s.sp.num_ref[1] = sh.sp.num_ref[1]; s.sp.cabac_init_idc = sh.sp.cabac_init_idc;
    s.sp.slice_qp = sh.sp.slice_qp; s.sp.transform_8x8_mode = sh.sp.transform_8x8_mode;
    s.sp.chroma_format_idc = sh.sp.chroma_format_idc; s.sp.direct_8x8_inference = sh.sp.direct_8x8_inference;
    d.slices.push_back(s); d.slice_block.push_back(coded_blk[next - 1]);
    return 0;
  }, &unsupported);
  // a container whose coded blocks belong to slices this library passes over (field/MBAFF, 4:2:2+ ... written by the reference's
  // compressor) cannot line up with the scan: say unsupported rather than corrupt
  if (rc) { d.rc = unsupported ? AVR_E_UNSUPPORTED : rc; return; }
  for (; next < coded_blk.size(); next++)
    if (d.blocks[coded_blk[next]].has_cabac) { d.rc = unsupported ? AVR_E_UNSUPPORTED : AVR_E_CORRUPT; return; }   // "Not all blocks were decoded."
}

// Compress side: the block sequence of one file (recode.cpp:1282-1304: literal up to each coded payload, the coded block, a
// size-only block for a slice left uncoded; :1129-1131: the rest as one literal).  span(i, &off, &size, &idx): idx < 0 = uncoded;
// payload(idx, &p, &len) = the range-coded bytes of coded slice idx.
template <class SpanFn, class PayloadFn>
inline void write_container(const uint8_t* in, size_t in_len, uint32_t nspans, SpanFn&& span, PayloadFn&& payload, Buf& o) {
  o.reserve(in_len + in_len / 16 + 64);
  size_t prev_end = 0;
  for (uint32_t i = 0; i < nspans; i++) {
    uint64_t off; uint32_t size; int32_t idx;
    span(i, &off, &size, &idx);
    if (idx < 0) { put_block_skip(o, size); continue; }
    put_block_literal(o, in + prev_end, (size_t)off - prev_end);
    prev_end = (size_t)off + size;
    const uint8_t* p; uint32_t len;
    payload((uint32_t)idx, &p, &len);
    put_block_cabac(o, size, p, len, (int)(size & 1), size > 1 ? in + prev_end - 1 : nullptr);
  }
  put_block_literal(o, in + prev_end, in_len - prev_end);
}

// Decompress side: the output file = literal blocks from the container, coded blocks from slice(k, &p, &len) = the re-encoded
// CABAC bytes of planned slice k, with the x264 padding patch (recode.cpp:1352-1362, 1387-1391: the last byte is replaced, or
// appended when the length parity differs).  The buffer is malloc'ed; false = out of memory.
template <class SliceFn>
inline bool assemble(const DFile& d, SliceFn&& slice, uint8_t** out_data, size_t* out_size) {
  auto patched_len = [](const Block& blk, uint32_t len) -> size_t {
    if (blk.has_parity && blk.has_last && blk.last_len > 0 && blk.parity != (int)(len & 1)) return (size_t)len + 1;
    return len;
  };
  std::vector<int> blk_slice(d.blocks.size(), -1);
  for (size_t k = 0; k < d.slice_block.size(); k++) blk_slice[d.slice_block[k]] = (int)k;
  size_t total = 0;
  for (size_t i = 0; i < d.blocks.size(); i++) {
    const Block& blk = d.blocks[i];
    if (blk.has_literal) total += blk.literal_len;
    else if (blk.has_cabac) {
      const uint8_t* p; uint32_t len;
      slice((uint32_t)blk_slice[i], &p, &len);
      total += patched_len(blk, len);
    }
  }
  *out_size = total;
  *out_data = (uint8_t*)malloc(total ? total : 1);
  if (!*out_data) return false;
  uint8_t* o = *out_data;
  for (size_t i = 0; i < d.blocks.size(); i++) {
    const Block& blk = d.blocks[i];
    if (blk.has_literal) { if (blk.literal_len) memcpy(o, blk.literal, blk.literal_len); o += blk.literal_len; }
    else if (blk.has_cabac) {
      const uint8_t* p; uint32_t len;
      slice((uint32_t)blk_slice[i], &p, &len);
      if (len) memcpy(o, p, len);
      if (blk.has_parity && blk.has_last && blk.last_len > 0) {
        if (blk.parity != (int)(len & 1)) o[len] = blk.last[0];
        else if (len) o[len - 1] = blk.last[0];
      }
      o += patched_len(blk, len);
    }
  }
  return true;
}

}  // namespace avrc
