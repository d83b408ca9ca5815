// Standard H.264 CABAC arithmetic decoding / encoding engines (ITU-T H.264 9.3.3.2, 9.3.4),
// host side.  The decoder is the bin source of the compress-side parse: it plays the role of
// ff_get_cabac / ff_get_cabac_bypass / ff_get_cabac_terminate on the private CABACContext copy
// in reference recode.cpp:1156-1159,1182-1199.  The encoder is used by the synthetic clip
// generator only.
//
// Context state bytes use the packed form (pStateIdx << 1) | valMPS, the same form the
// reference's hooks see through `uint8_t* state` (ffmpeg/libavcodec/cabac_functions.h:108-111),
// so a recorder built on this decoder can be installed as CABACCodingHooks::get directly.
#pragma once
#include <stdint.h>
#include <vector>
#include "avr_common.h"
#include "h264_std_tables.h"

namespace avr {

// (m,n) -> initial packed state, 9.3.1.1 ; planes: 0 = I, 1..3 = cabac_init_idc 0..2
inline void cabac_init_states(uint8_t* st /*[1024]*/, int slice_is_intra, int cabac_init_idc, int slice_qp) {
  static const int8_t* const M[4] = {avr_ctx_init_m0, avr_ctx_init_m1, avr_ctx_init_m2, avr_ctx_init_m3};
  static const int8_t* const N[4] = {avr_ctx_init_n0, avr_ctx_init_n1, avr_ctx_init_n2, avr_ctx_init_n3};
  int plane = slice_is_intra ? 0 : 1 + cabac_init_idc;
  int qp = slice_qp < 0 ? 0 : slice_qp > 51 ? 51 : slice_qp;
  for (int i = 0; i < 1024; i++) {
    int pre = ((M[plane][i] * qp) >> 4) + N[plane][i];
    if (pre < 1) pre = 1;
    if (pre > 126) pre = 126;
    // pStateIdx/valMPS per 9.3.1.1, packed
    if (pre <= 63) st[i] = (uint8_t)((63 - pre) << 1);
    else st[i] = (uint8_t)(((pre - 64) << 1) | 1);
  }
}

class CabacDecoder {
 public:
  void init(const uint8_t* buf, size_t size) {
    p_ = begin_ = buf; end_ = buf + size; bitbuf_ = 0; nbits_ = 0; over_ = 0;
    range_ = 510;
    offset_ = read_bits(9);
  }
  inline int decode(uint8_t* state) {
    uint32_t s = *state;
    uint32_t p = s >> 1, mps = s & 1;
    uint32_t lps = avr_range_tab_lps[(p << 2) | ((range_ >> 6) & 3)];
    range_ -= lps;
    int bin;
    if (offset_ >= range_) {
      bin = (int)(mps ^ 1);
      offset_ -= range_;
      range_ = lps;
      if (p == 0) mps ^= 1;
      p = avr_trans_idx_lps[p];
    } else {
      bin = (int)mps;
      if (p < 62) p++;
    }
    *state = (uint8_t)((p << 1) | mps);
    if (range_ < 256) {
      int sh = __builtin_clz(range_) - 23;
      range_ <<= sh;
      offset_ = (offset_ << sh) | read_bits(sh);
    }
    return bin;
  }
  inline int bypass() {
    offset_ = (offset_ << 1) | read_bits(1);
    if (offset_ >= range_) { offset_ -= range_; return 1; }
    return 0;
  }
  inline int terminate() {
    range_ -= 2;
    if (offset_ >= range_) return 1;
    if (range_ < 256) { range_ <<= 1; offset_ = (offset_ << 1) | read_bits(1); }
    return 0;
  }

 private:
  inline uint32_t read_bits(int n) {
    if (n == 0) return 0;
    while (nbits_ < n) {
      uint32_t b = 0;
      if (p_ < end_) b = *p_++; else over_++;
      bitbuf_ = (bitbuf_ << 8) | b;
      nbits_ += 8;
    }
    nbits_ -= n;
    return (uint32_t)((bitbuf_ >> nbits_) & ((1u << n) - 1));
  }
  const uint8_t* p_ = nullptr; const uint8_t* end_ = nullptr; const uint8_t* begin_ = nullptr;
  uint64_t bitbuf_ = 0; int nbits_ = 0;
  size_t over_ = 0;            // zero bytes supplied past the end of the buffer
  uint32_t range_ = 510, offset_ = 0;

 public:
  // Bits the engine has read, the 9-bit codIOffset window included.  After end_of_slice_flag = 1 of a conforming slice this is
  // the length of slice_data() plus one: the encoder's flush (9.3.4.5) ends with WriteBits(((low >> 7) & 3) | 1, 2), whose last
  // bit is rbsp_stop_one_bit, and the window reaches exactly that far.
  size_t bits_consumed() const { return ((size_t)(p_ - begin_) + over_) * 8 - (size_t)nbits_; }
};

// Textbook encoder (9.3.4.2 .. 9.3.4.5) writing into a byte vector.
class CabacEncoder {
 public:
  void init(std::vector<uint8_t>* out) {
    out_ = out; low_ = 0; range_ = 510; outstanding_ = 0; first_ = true; cur_ = 0; ncur_ = 0;
  }
  void encode(int bin, uint8_t* state) {
    uint32_t s = *state, p = s >> 1, mps = s & 1;
    uint32_t lps = avr_range_tab_lps[(p << 2) | ((range_ >> 6) & 3)];
    range_ -= lps;
    if ((uint32_t)bin != mps) {
      low_ += range_;
      range_ = lps;
      if (p == 0) mps ^= 1;
      p = avr_trans_idx_lps[p];
    } else if (p < 62) p++;
    *state = (uint8_t)((p << 1) | mps);
    renorm();
  }
  void bypass(int bin) {
    low_ <<= 1;
    if (bin) low_ += range_;
    if (low_ >= 1024) { put_bit(1); low_ -= 1024; }
    else if (low_ < 512) put_bit(0);
    else { low_ -= 512; outstanding_++; }
  }
  // terminate bin; when bin==1 performs the 9.3.4.5 flush and writes the rbsp stop bit + alignment
  void terminate(int bin) {
    range_ -= 2;
    if (bin) {
      low_ += range_;
      range_ = 2;
      renorm();
      put_bit((low_ >> 9) & 1);
      write_bit((low_ >> 8) & 1);
      write_bit(1);  // ((low >> 7) & 3) | 1 : the last bit written is the rbsp_stop_one_bit
      while (ncur_) write_bit(0);
    } else {
      renorm();
    }
  }

 private:
  void renorm() {
    while (range_ < 256) {
      if (low_ < 256) put_bit(0);
      else if (low_ >= 512) { low_ -= 512; put_bit(1); }
      else { low_ -= 256; outstanding_++; }
      range_ <<= 1;
      low_ <<= 1;
    }
  }
  void put_bit(int b) {
    if (first_) first_ = false; else write_bit(b);
    while (outstanding_ > 0) { write_bit(1 - b); outstanding_--; }
  }
  void write_bit(int b) {
    cur_ = (cur_ << 1) | (uint32_t)(b & 1);
    if (++ncur_ == 8) { out_->push_back((uint8_t)cur_); cur_ = 0; ncur_ = 0; }
  }
  std::vector<uint8_t>* out_ = nullptr;
  uint32_t low_ = 0, range_ = 510; int outstanding_ = 0; bool first_ = true;
  uint32_t cur_ = 0; int ncur_ = 0;
};

}  // namespace avr
