// Host-side H.264 byte-stream plumbing for the native (libavcodec-free) front end:
// Annex-B NAL scan, emulation-prevention removal, SPS / PPS / slice_header() parse
// (ITU-T H.264 7.3.1, 7.3.2.1, 7.3.2.2, 7.3.3).  Only what is needed to locate the CABAC
// payload of each slice and to fill avr::SliceParams.
//
// Mirrors what the reference gets from libavcodec before its hooks fire:
//   - NAL extent and RBSP unescape  : ffmpeg/libavcodec/h264.c ff_h264_decode_nal
//   - trailing-zero strip + stop bit: ffmpeg/libavcodec/h264.c decode_nal_units/get_bit_length
//   - CABAC start / size            : ffmpeg/libavcodec/h264_slice.c:2357-2364
//     (buf = first byte after the byte-aligned slice header, size = (bits_left+7)/8)
#pragma once
#include <stdint.h>
#include <string.h>
#include <vector>
#include "h264_syntax.h"

namespace avr {

class BitReader {
 public:
  BitReader(const uint8_t* p, size_t nbits) : p_(p), n_(nbits), pos_(0), err_(false) {}
  uint32_t u(int n) {
    uint32_t v = 0;
    for (int i = 0; i < n; i++) {
      if (pos_ >= n_) { err_ = true; return 0; }
      v = (v << 1) | ((p_[pos_ >> 3] >> (7 - (pos_ & 7))) & 1);
      pos_++;
    }
    return v;
  }
  uint32_t ue() {
    int z = 0;
    while (!err_ && u(1) == 0) { if (++z > 31) { err_ = true; return 0; } }
    if (err_) return 0;
    return z ? ((1u << z) - 1 + u(z)) : 0;
  }
  int32_t se() {
    uint32_t k = ue();
    return (k & 1) ? (int32_t)((k + 1) >> 1) : -(int32_t)(k >> 1);
  }
  size_t pos() const { return pos_; }
  size_t left() const { return n_ - pos_; }
  bool err() const { return err_; }
 private:
  const uint8_t* p_; size_t n_, pos_; bool err_;
};

struct Sps {
  bool valid = false;
  int profile_idc = 0, chroma_format_idc = 1, separate_colour_plane = 0;
  int bit_depth_luma = 8, bit_depth_chroma = 8;
  int log2_max_frame_num = 4, poc_type = 0, log2_max_poc_lsb = 4, delta_pic_order_always_zero = 0;
  int mb_w = 0, map_h = 0, frame_mbs_only = 1, mb_aff = 0, direct_8x8_inference = 0;
};
struct Pps {
  bool valid = false;
  int sps_id = 0, cabac = 0, pic_order_present = 0, num_slice_groups = 1;
  int num_ref_default[2] = {1, 1};
  int weighted_pred = 0, weighted_bipred_idc = 0, pic_init_qp = 26;
  int deblocking_control = 0, constrained_intra_pred = 0, redundant_pic_cnt_present = 0;
  int transform_8x8_mode = 0;
};

struct SliceHeaderInfo {
  SliceParams sp;
  int frame_num = 0;
  int nal_unit_type = 0, nal_ref_idc = 0;
  int pps_id = 0;
  bool cabac = false;
  bool supported = true;     // false: SP/SI / field / MBAFF / FMO / high bit depth / 4:2:2+
  size_t header_bits = 0;    // bits consumed by slice_header() in the RBSP (after the NAL header byte)
};

static inline void skip_scaling_list(BitReader& br, int size) {
  int last = 8, next = 8;
  for (int j = 0; j < size; j++) {
    if (next != 0) { int d = br.se(); next = (last + d + 256) % 256; }
    last = next == 0 ? last : next;
  }
}

static inline uint32_t br_ue_after_scp(BitReader& br, uint32_t chroma_format_idc, int* separate_colour_plane) {
  if (chroma_format_idc == 3) *separate_colour_plane = (int)br.u(1);
  return br.ue();                               // bit_depth_luma_minus8
}
// Picture sizes the chain kernels accept (avr_dbatch_create enforces the same bound): larger streams are marked
// unsupported and stored as literal blocks instead of sizing device stores from attacker-chosen values.
enum { AVR_MAX_MB_DIM = 1024 };
// rbsp points AFTER the one-byte NAL header
static inline bool parse_sps(const uint8_t* rbsp, size_t n, Sps* out, int* id_out) {
  BitReader br(rbsp, n * 8);
  Sps s;
  s.profile_idc = br.u(8);
  br.u(8);  // constraint flags + reserved
  br.u(8);  // level_idc
  const uint32_t id = br.ue();                 // unsigned: a 32-bit Exp-Golomb value must not turn into a negative index
  if (id > 31u) return false;
  int p = s.profile_idc;
  if (p == 100 || p == 110 || p == 122 || p == 244 || p == 44 || p == 83 || p == 86 || p == 118 || p == 128 ||
      p == 138 || p == 139 || p == 134 || p == 135) {
    const uint32_t cfi = br.ue(), bdl = br_ue_after_scp(br, cfi, &s.separate_colour_plane), bdc = br.ue();
    if (cfi > 3u || bdl > 6u || bdc > 6u) return false;
    s.chroma_format_idc = (int)cfi;
    s.bit_depth_luma = 8 + (int)bdl;
    s.bit_depth_chroma = 8 + (int)bdc;
    br.u(1);  // qpprime_y_zero_transform_bypass
    if (br.u(1)) {  // seq_scaling_matrix_present
      int cnt = s.chroma_format_idc != 3 ? 8 : 12;
      for (int i = 0; i < cnt; i++)
        if (br.u(1)) skip_scaling_list(br, i < 6 ? 16 : 64);
    }
  }
  const uint32_t l2fn = br.ue(), poc = br.ue();
  if (l2fn > 12u || poc > 2u) return false;       // 7.4.2.1.1 ranges; u(log2_max_frame_num) reads at most 16 bits
  s.log2_max_frame_num = 4 + (int)l2fn;
  s.poc_type = (int)poc;
  if (s.poc_type == 0) {
    const uint32_t l2poc = br.ue();
    if (l2poc > 12u) return false;
    s.log2_max_poc_lsb = 4 + (int)l2poc;
  } else if (s.poc_type == 1) {
    s.delta_pic_order_always_zero = br.u(1);
    br.se(); br.se();
    const uint32_t k = br.ue();
    if (k > 255u) return false;
    for (uint32_t i = 0; i < k && !br.err(); i++) br.se();
  }
  br.ue();  // max_num_ref_frames
  br.u(1);  // gaps_in_frame_num_value_allowed
  const uint32_t wm1 = br.ue(), hm1 = br.ue();
  if (wm1 >= 65535u || hm1 >= 65535u) return false;   // keeps every later product inside 32 bits; > AVR_MAX_MB_DIM goes literal
  s.mb_w = (int)wm1 + 1;
  s.map_h = (int)hm1 + 1;
  s.frame_mbs_only = br.u(1);
  if (!s.frame_mbs_only) s.mb_aff = br.u(1);
  s.direct_8x8_inference = br.u(1);
  // frame cropping / VUI not needed
  if (br.err()) return false;
  s.valid = true;
  *out = s; *id_out = (int)id;
  return true;
}

static inline bool more_rbsp_data(const uint8_t* rbsp, size_t n, size_t bitpos) {
  // true if there is data before the rbsp trailing bits
  while (n > 0 && rbsp[n - 1] == 0) n--;
  if (n == 0) return false;
  int r = 0; uint8_t v = rbsp[n - 1];
  while (!(v & 1)) { v >>= 1; r++; }
  size_t last_data_bit = n * 8 - (size_t)r - 1;  // position of the stop bit
  return bitpos < last_data_bit;
}

static inline bool parse_pps(const uint8_t* rbsp, size_t n, const Sps* sps_tab, Pps* out, int* id_out) {
  BitReader br(rbsp, n * 8);
  Pps p;
  const uint32_t id = br.ue();
  if (id > 255u) return false;
  const uint32_t sid = br.ue();
  if (sid > 31u) return false;
  p.sps_id = (int)sid;
  p.cabac = br.u(1);
  p.pic_order_present = br.u(1);
  const uint32_t nsg = br.ue();
  if (nsg > 7u) return false;
  p.num_slice_groups = (int)nsg + 1;
  if (p.num_slice_groups > 1) { p.valid = true; *out = p; *id_out = (int)id; return !br.err(); }  // FMO: unsupported, flagged later
  const uint32_t nr0 = br.ue(), nr1 = br.ue();
  if (nr0 > 31u || nr1 > 31u) return false;
  p.num_ref_default[0] = (int)nr0 + 1;
  p.num_ref_default[1] = (int)nr1 + 1;
  p.weighted_pred = br.u(1);
  p.weighted_bipred_idc = br.u(2);
  p.pic_init_qp = 26 + br.se();
  br.se();  // pic_init_qs
  br.se();  // chroma_qp_index_offset
  p.deblocking_control = br.u(1);
  p.constrained_intra_pred = br.u(1);
  p.redundant_pic_cnt_present = br.u(1);
  if (!br.err() && more_rbsp_data(rbsp, n, br.pos())) {
    p.transform_8x8_mode = br.u(1);
    if (br.u(1)) {  // pic_scaling_matrix_present
      const Sps& s = sps_tab[p.sps_id];
      int cnt = 6 + ((s.chroma_format_idc != 3) ? 2 : 6) * p.transform_8x8_mode;
      for (int i = 0; i < cnt; i++)
        if (br.u(1)) skip_scaling_list(br, i < 6 ? 16 : 64);
    }
    br.se();  // second_chroma_qp_index_offset
  }
  if (br.err()) return false;
  p.valid = true;
  *out = p; *id_out = (int)id;
  return true;
}

// rbsp points AFTER the NAL header byte; nbits = number of RBSP bits before the stop bit.
static inline bool parse_slice_header(const uint8_t* rbsp, size_t nbits, int nal_unit_type, int nal_ref_idc,
                                      const Sps* sps_tab, const Pps* pps_tab, SliceHeaderInfo* out) {
  BitReader br(rbsp, nbits);
  SliceHeaderInfo h;
  h.nal_unit_type = nal_unit_type; h.nal_ref_idc = nal_ref_idc;
  const uint32_t first_mb_u = br.ue();
  const uint32_t st_u = br.ue();
  if (st_u > 9u || first_mb_u > 0x7FFFFFFFu) return false;
  const int first_mb = (int)first_mb_u;
  const int st = (int)(st_u % 5u);  // 0 P, 1 B, 2 I, 3 SP, 4 SI
  const uint32_t pps_id_u = br.ue();
  if (br.err() || pps_id_u > 255u || !pps_tab[pps_id_u].valid) return false;
  const int pps_id = (int)pps_id_u;
  const Pps& pps = pps_tab[pps_id];
  if (!sps_tab[pps.sps_id].valid) return false;
  const Sps& sps = sps_tab[pps.sps_id];
  h.pps_id = pps_id;
  h.cabac = pps.cabac != 0;
  if (sps.separate_colour_plane) br.u(2);
  h.frame_num = br.u(sps.log2_max_frame_num);
  int field_pic = 0;
  if (!sps.frame_mbs_only) {
    field_pic = br.u(1);
    if (field_pic) br.u(1);
  }
  if (nal_unit_type == 5) br.ue();  // idr_pic_id
  if (sps.poc_type == 0) {
    br.u(sps.log2_max_poc_lsb);
    if (pps.pic_order_present && !field_pic) br.se();
  } else if (sps.poc_type == 1 && !sps.delta_pic_order_always_zero) {
    br.se();
    if (pps.pic_order_present && !field_pic) br.se();
  }
  if (pps.redundant_pic_cnt_present) br.ue();
  if (st == 1) br.u(1);  // direct_spatial_mv_pred
  int num_ref[2] = {pps.num_ref_default[0], pps.num_ref_default[1]};
  if (st == 0 || st == 1 || st == 3) {
    if (br.u(1)) {
      const uint32_t n0 = br.ue();
      if (n0 > 31u) return false;
      num_ref[0] = (int)n0 + 1;
      if (st == 1) { const uint32_t n1 = br.ue(); if (n1 > 31u) return false; num_ref[1] = (int)n1 + 1; }
    }
  }
  if (st != 2 && st != 4) {  // ref_pic_list_modification
    for (int l = 0; l < (st == 1 ? 2 : 1); l++) {
      if (br.u(1)) {
        for (int guard = 0; guard < 100 && !br.err(); guard++) {
          uint32_t idc = br.ue();
          if (idc == 3) break;
          if (idc > 3) return false;
          br.ue();
        }
      }
    }
  }
  if ((pps.weighted_pred && (st == 0 || st == 3)) || (pps.weighted_bipred_idc == 1 && st == 1)) {
    br.ue();  // luma_log2_weight_denom
    if (sps.chroma_format_idc != 0) br.ue();
    for (int l = 0; l < (st == 1 ? 2 : 1); l++)
      for (int i = 0; i < num_ref[l] && !br.err(); i++) {
        if (br.u(1)) { br.se(); br.se(); }
        if (sps.chroma_format_idc != 0 && br.u(1)) { br.se(); br.se(); br.se(); br.se(); }
      }
  }
  if (nal_ref_idc != 0) {  // dec_ref_pic_marking
    if (nal_unit_type == 5) { br.u(1); br.u(1); }
    else if (br.u(1)) {
      for (int guard = 0; guard < 100 && !br.err(); guard++) {
        uint32_t op = br.ue();
        if (op == 0) break;
        if (op == 1 || op == 3) br.ue();
        if (op == 2) br.ue();
        if (op == 3 || op == 6) br.ue();
        if (op == 4) br.ue();
      }
    }
  }
  uint32_t cabac_init_idc = 0;
  if (pps.cabac && st != 2 && st != 4) cabac_init_idc = br.ue();
  int qp = pps.pic_init_qp + br.se();
  if (st == 3 || st == 4) { if (st == 3) br.u(1); br.se(); }
  if (pps.deblocking_control) {
    if (br.ue() != 1) { br.se(); br.se(); }
  }
  // SliceQPY outside 0..51 (8-bit video): libavcodec refuses the slice header (h264_slice.c "QP %u out of range"), so no hook
  // ever sees the slice and the reference's container has no block for it
  if (br.err() || cabac_init_idc > 2u || qp < 0 || qp > 51) return false;
  h.header_bits = br.pos();
  h.sp.slice_type = st;
  h.sp.first_mb = first_mb;
  h.sp.mb_w = sps.mb_w;
  h.sp.mb_h = sps.map_h * (2 - sps.frame_mbs_only);
  h.sp.num_ref[0] = num_ref[0]; h.sp.num_ref[1] = num_ref[1];
  h.sp.cabac_init_idc = (int)cabac_init_idc;
  h.sp.slice_qp = qp;
  h.sp.transform_8x8_mode = pps.transform_8x8_mode;
  h.sp.chroma_format_idc = sps.chroma_format_idc;
  h.sp.direct_8x8_inference = sps.direct_8x8_inference;
  h.supported = (st == 0 || st == 1 || st == 2) && !field_pic && !sps.mb_aff && sps.frame_mbs_only &&
                pps.num_slice_groups == 1 && sps.chroma_format_idc <= 1 && sps.bit_depth_luma == 8 &&
                sps.bit_depth_chroma == 8 && h.sp.mb_w <= AVR_MAX_MB_DIM && h.sp.mb_h <= AVR_MAX_MB_DIM &&
                (int64_t)first_mb < (int64_t)h.sp.mb_w * (int64_t)h.sp.mb_h;
  *out = h;
  return true;
}

// One NAL unit located in an Annex-B byte stream.
struct NalSpan {
  size_t start;   // offset of the NAL header byte
  size_t end;     // one past the last byte of the NAL as libavcodec delimits it
};

// Finds the next NAL at or after `pos`. NAL end = first "00 00 0x" (x<=2... i.e. 00 00 00 / 00 00 01
// / 00 00 02) after the header, like ff_h264_decode_nal, else end of buffer.
static inline bool next_nal(const uint8_t* d, size_t n, size_t pos, NalSpan* out) {
  // start code: first 00 00 01 at or after pos (memchr for the zero, then two compares)
  size_t i = pos;
  for (;;) {
    if (i + 3 > n) return false;
    const uint8_t* z = static_cast<const uint8_t*>(memchr(d + i, 0, n - 2 - i));
    if (!z) return false;
    i = (size_t)(z - d);
    if (d[i + 1] == 0 && d[i + 2] == 1) break;
    i++;
  }
  const size_t s = i + 3;
  if (s >= n) return false;
  // end: first 00 00 0x with x <= 2 at e >= s + 1, e + 2 < n; else the end of the buffer
  size_t e = s + 1;
  for (;;) {
    if (e + 2 >= n) { e = n; break; }
    const uint8_t* z = static_cast<const uint8_t*>(memchr(d + e, 0, n - 2 - e));
    if (!z) { e = n; break; }
    e = (size_t)(z - d);
    if (d[e + 1] == 0 && d[e + 2] <= 2) break;
    e++;
  }
  out->start = s; out->end = e;
  return true;
}

// Unescape NAL payload (after the header byte) into `rbsp`: every 03 that follows two zero bytes is dropped
// (ff_h264_decode_nal).  `esc` (optional) gets the RBSP offset at which each dropped byte sat, ascending, so that
// the raw offset of RBSP byte k is k + (number of entries <= k).  Segments between escapes are copied whole.
static inline void unescape_nal(const uint8_t* p, size_t n, std::vector<uint8_t>* rbsp, std::vector<uint32_t>* esc) {
  rbsp->clear();
  if (esc) esc->clear();
  rbsp->reserve(n);
  size_t seg = 0;
  for (size_t i = 2; i < n;) {
    const uint8_t* q = static_cast<const uint8_t*>(memchr(p + i, 3, n - i));
    if (!q) break;
    i = (size_t)(q - p);
    if (p[i - 1] == 0 && p[i - 2] == 0 && i - 2 >= seg) {     // both zeros were kept bytes of this segment
      rbsp->insert(rbsp->end(), p + seg, p + i);
      if (esc) esc->push_back((uint32_t)rbsp->size());
      seg = i + 1;
    }
    i++;
  }
  rbsp->insert(rbsp->end(), p + seg, p + n);
}
// raw offset (within the NAL payload) of RBSP byte k
static inline size_t rbsp_to_raw(const std::vector<uint32_t>& esc, size_t k) {
  size_t lo = 0, hi = esc.size();
  while (lo < hi) { const size_t mid = (lo + hi) / 2; if (esc[mid] <= k) lo = mid + 1; else hi = mid; }
  return k + lo;
}

}  // namespace avr
