// Device-side primitives of the entropy hot path (host+device so that tests/ can run the very
// same source on the CPU as an emulated grid; the product library only ever calls them from
// kernels, see avr_kernels.cu).
//
//   exact range/total          reference recode.cpp:823-827
//   estimator update/halving   reference recode.cpp:1037-1054
//   64-bit range encoder       reference arithmetic_code.h:106-190  (arithmetic_code<uint64_t,uint8_t>)
//   64-bit range decoder       reference arithmetic_code.h:220-288
//   H.264 CABAC encoder        reference cabac_code.h:30-67 (arithmetic_code<uint32_t,uint16_t,0x200>)
//
// Carry handling: the reference defers digits that a later carry could still change in an
// `overflow` vector (arithmetic_code.h:153-174).  Here digits are written to the per-slice
// scratch at once and a carry is propagated backwards through the bytes already written;
// the byte string produced is identical (a deferred digit is emitted unchanged or
// incremented, which is exactly what the backward increment does).
#pragma once
#include "avr_common.h"
#include "avr_div_tab.h"

#if defined(__CUDA_ARCH__)
#define AVR_MULHI64(a, b) __umul64hi((a), (b))
#define AVR_CLZ32(x) __clz((int)(x))
#else
#define AVR_MULHI64(a, b) ((uint64_t)(((unsigned __int128)(a) * (unsigned __int128)(b)) >> 64))
#define AVR_CLZ32(x) __builtin_clz((unsigned)(x))
#endif

struct AvrDivTab {
  uint64_t m[98];
  uint8_t s[98];
};
#define AVR_DIV_TAB_INIT {AVR_DIV_TAB_M, AVR_DIV_TAB_S}

// range_of_1 = (range / total) * pos, exact (total in 2..97)
static AVR_HD uint64_t avr_range_of_1(uint64_t range, uint32_t pos, uint32_t total, const AvrDivTab* dt) {
  uint64_t q = AVR_MULHI64(range, dt->m[total]) >> dt->s[total];
  return q * (uint64_t)pos;
}

// Estimator cell: low byte = pos, high byte = neg, both start at 1 (0x0101), never exceed 0x60.
#define AVR_EST_INIT 0x0101u
static AVR_HD uint32_t avr_est_update(uint32_t cell, int symbol, uint32_t limit /*0x50 sig-map, 0x60 else*/) {
  uint32_t pos = cell & 0xFF, neg = cell >> 8;
  if (symbol) pos++; else neg++;
  if (pos + neg > limit) { pos = (pos + 1) >> 1; neg = (neg + 1) >> 1; }
  return pos | (neg << 8);
}

// ---------------------------------------------------------------------------------------------
struct AvrRangeEnc {
  uint64_t low, range;
  uint8_t* out;       // per-slice scratch
  uint32_t n, cap;
  int err;
  AVR_HD void init(uint8_t* o, uint32_t c) { low = 0; range = 1ull << 63; out = o; n = 0; cap = c; err = 0; }
  AVR_HD void carry_back() {
    uint32_t i = n;
    while (i > 0) { if (++out[--i] != 0) break; }
  }
  AVR_HD void emit_digit() {
    if (low >> 63) { carry_back(); low &= (1ull << 63) - 1; }
    if (n < cap) out[n] = (uint8_t)(low >> 55); else err = AVR_E_OVERFLOW;
    n++;
    low = (low & ((1ull << 55) - 1)) << 8;
    range <<= 8;
  }
  AVR_HD void put(int symbol, uint64_t range_of_1) {
    const uint64_t range_of_0 = range - range_of_1;
    low += symbol ? range_of_0 : 0ull;                 // selects: lanes that code different slices stay converged
    range = symbol ? range_of_1 : range_of_0;
    if (range < (1ull << 51)) {
      if (range == 0) { err = AVR_E_CORRUPT; range = 1; }
      while (range < (1ull << 55)) emit_digit();
    }
  }
  AVR_HD void finish() {
    for (uint64_t stop = 1ull << 62; stop > 0; stop >>= 1) {
      uint64_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) { range = 1; emit_digit(); }
    range = 0;
  }
};

// Input bytes are fetched eight at a time from a 16-byte aligned, zero-padded pool and the next
// word is requested as soon as the current one is opened, so the load latency is off the per-bin
// dependency chain.  Bytes at or beyond `size` read as zero (arithmetic_code.h:283).
struct AvrRangeDec {
  uint64_t low, range;
  const uint8_t* in;
  uint32_t pos, size;
  uint32_t next_digit;
  uint64_t w_cur, w_next;
  AVR_HD uint64_t load_word(uint32_t off) const {
    if (off >= size) return 0;
    uint64_t w;
#if defined(__CUDA_ARCH__)
    w = *reinterpret_cast<const uint64_t*>(in + off);
#else
    w = 0;
    for (int k = 7; k >= 0; k--) w = (w << 8) | in[off + (uint32_t)k];   // pool is padded: in[off..off+7] is readable
#endif
    uint32_t valid = size - off;
    if (valid < 8) w &= (1ull << (8 * valid)) - 1;
    return w;
  }
  AVR_HD uint32_t aligned() {
    if ((pos & 7) == 0 && pos) { w_cur = w_next; w_next = load_word(pos + 8); }
    uint32_t b = (uint32_t)(w_cur >> (8 * (pos & 7))) & 0xFF;
    pos++;
    return b;
  }
  AVR_HD void consume() {
    uint32_t d = aligned();
    uint32_t digit = ((next_digit << 7) | (d >> 1)) & 0xFF;
    next_digit = d;
    low = (low << 8) + digit;
    range <<= 8;
  }
  // p must be 8-byte aligned and readable up to ((sz + 15) & ~7) bytes
  AVR_HD void init(const uint8_t* p, uint32_t sz) {
    in = p; pos = 0; size = sz;
    w_cur = load_word(0); w_next = load_word(8);
    next_digit = aligned();
    low = next_digit >> 1;
    range = 128;
    while (range < (1ull << 63)) consume();
  }
  AVR_HD int get(uint64_t range_of_1) {
    uint64_t range_of_0 = range - range_of_1;
    int symbol = low >= range_of_0;
    if (symbol) { low -= range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < (1ull << 51)) {
      if (range == 0) return symbol;   // corrupt input; caller bounds the walk
      while (range < (1ull << 55)) consume();
    }
    return symbol;
  }
};

// ---------------------------------------------------------------------------------------------
// CABAC encoder.  `lps_tab`/`trans_tab` = ITU-T H.264 Table 9-44 / 9-45 (avr_range_tab_lps,
// avr_trans_idx_lps); state byte = (pStateIdx << 1) | valMPS, the form libavcodec's hooks pass.
struct AvrCabacEnc {
  uint32_t low, range;
  uint8_t* out;
  uint32_t n, cap;
  int err;
  AVR_HD void init(uint8_t* o, uint32_t c) { low = 0; range = (1u << 22) * 0x1FEu; out = o; n = 0; cap = c; err = 0; }
  AVR_HD void carry_back() {
    uint32_t i = n;
    while (i > 0) { if (++out[--i] != 0) break; }
  }
  AVR_HD void push(uint32_t b) { if (n < cap) out[n] = (uint8_t)b; else err = AVR_E_OVERFLOW; n++; }
  AVR_HD void emit16() {
    if (low >> 31) { carry_back(); low &= 0x7FFFFFFFu; }
    uint32_t digit = low >> 15;
    push(digit >> 8); push(digit & 0xFF);
    low = (low & 0x7FFFu) << 16;
    range <<= 16;
  }
  AVR_HD void emit8() {
    if (low >> 31) { carry_back(); low &= 0x7FFFFFFFu; }
    push(low >> 23);
    low = (low & 0x7FFFFFu) << 8;
    range <<= 8;
  }
  AVR_HD void put_raw(int symbol, uint32_t range_of_1) {
    uint32_t range_of_0 = range - range_of_1;
    if (symbol) { low += range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < 0x200u) {
      if (range == 0) { err = AVR_E_CORRUPT; range = 1; }
      while (range < 0x8000u) emit16();
    }
  }
  AVR_HD void put(int symbol, uint8_t* state, const uint8_t* lps_tab, const uint8_t* trans_tab) {
    uint32_t s = *state, p = s >> 1, mps = s & 1;
    int lps = symbol != (int)mps;
    int normalize = 23 - AVR_CLZ32(range);          // floor(log2(range / 0x100)), range >= 0x200
    uint32_t approx = range >> (normalize - 1);
    uint32_t rlps = (uint32_t)lps_tab[(p << 2) | ((approx >> 7) & 3)] << normalize;
    put_raw(lps, rlps);
    if (lps) { if (p == 0) mps ^= 1; p = trans_tab[p]; } else if (p < 62) p++;
    *state = (uint8_t)((p << 1) | mps);
  }
  AVR_HD void put_bypass(int symbol) { put_raw(symbol, range >> 1); }
  // returns 1 when the slice is finished (bytes final, trailing 0x80 stripped: recode.cpp:1508-1515)
  AVR_HD int put_terminate(int eos) {
    int normalize = 23 - AVR_CLZ32(range);
    put_raw(eos, 2u << normalize);
    if (!eos) return 0;
    for (uint32_t stop = 1u << 30; stop > 0; stop >>= 1) {
      uint32_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) { range = 1; emit8(); }
    range = 0;
    if (n > 0 && n <= cap && out[n - 1] == 0x80) n--;
    return 1;
  }
};
