// Independent bin streams through the generic range coder (BASELINE configs[0]: the reference's coder test,
// test/arithmetic_code.cpp:93-110, S streams at a time instead of one).
//
//   coder      arithmetic_code<uint64_t, Digit> with Digit = uint16_t (the test, test/arithmetic_code.cpp:93)
//              or uint8_t (production, recode.cpp:322-323), output iterator over uint8_t:
//              put-time digits are Digit wide and leave MSB first (arithmetic_code.h:185-190), finish()
//              flushes in 8-bit digits (arithmetic_code.h:139-142)
//   model      FIXED_HALF: p1 = range/2 (test/arithmetic_code.cpp:97,105)
//              ADAPTIVE:   one estimator per stream, p1 = (range/(pos+neg))*pos, counts start at 1 and halve
//                          above 0x60 (recode.cpp:823-827, 1037-1054, 1064)
//
// One stream per THREAD: streams share nothing, every lane runs the same straight-line step, and the two
// data-dependent parts (digit output, carry) are rare.  Host+device so tests/host_emul.cpp runs the same
// bodies on the CPU; the library only calls them from the kernels in avr_streams.cu.
#pragma once
#include "avr_device.h"

#ifndef AVR_STREAM_FIXED_HALF
#define AVR_STREAM_FIXED_HALF 0
#define AVR_STREAM_ADAPTIVE 1
#endif

template <int DIGIT_BITS>
struct AvrStreamEnc {
  static constexpr int kTop = 63 - DIGIT_BITS;                  // most_significant_digit = 2^kTop (arithmetic_code.h:151)
  static constexpr uint64_t kMinRange = 1ull << (kTop - 4);     // (fixed_one / digit_base) / 16 (arithmetic_code.h:61-62)
  uint64_t low, range;
  uint8_t* out;
  uint32_t n, cap;
  int err;
  // Output bytes wait in registers and leave eight at a time as one aligned store: `acc` collects the newest bytes
  // (oldest on top), `held` is the last complete group, written only when the next one is complete.  A carry
  // (arithmetic_code.h:153-159) then ends in registers unless eight or more pending bytes are all 0xFF; without the
  // held group every carry right after a store was a read-modify-write of global memory on the coder's critical path
  // (about every other 16-bit digit of an incompressible stream).  n counts all bytes pushed; the scratch of a stream
  // starts 16-byte aligned.
  uint64_t acc, held;
  AVR_HD void init(uint8_t* o, uint32_t c) { low = 0; range = 1ull << 63; out = o; n = 0; cap = c; err = 0; acc = 0; held = 0; }
  AVR_HD void store_group(uint32_t at, uint64_t w) {             // eight bytes, oldest first, at byte offset `at`
    if (at + 8 > cap) { err = AVR_E_OVERFLOW; return; }
    w = ((w & 0x00FF00FF00FF00FFull) << 8) | ((w >> 8) & 0x00FF00FF00FF00FFull);
    w = ((w & 0x0000FFFF0000FFFFull) << 16) | ((w >> 16) & 0x0000FFFF0000FFFFull);
    w = (w << 32) | (w >> 32);
    *reinterpret_cast<uint64_t*>(out + at) = w;
  }
  AVR_HD void push(uint32_t b) {
    acc = (acc << 8) | (uint64_t)(b & 0xFF);
    n++;
    if ((n & 7) == 0) {
      if (n >= 16) store_group(n - 16, held);
      held = acc; acc = 0;
    }
  }
  AVR_HD void carry_back() {
    const uint32_t pend = n & 7;
    if (pend) {
      acc += 1;
      if ((acc >> (8 * pend)) == 0) return;
      acc &= (1ull << (8 * pend)) - 1;                           // all pending bytes were 0xFF
    }
    if (n >= 8) { if (++held != 0) return; }                     // the held group; all 0xFF again: on into memory
    uint32_t i = n >= 8 ? (n & ~7u) - 8 : 0; if (i > cap) i = cap & ~7u;
    while (i > 0) { if (++out[--i] != 0) break; }
  }
  AVR_HD void flush() {                                          // after finish(): the held group, then the bytes still in acc
    const uint32_t pend = n & 7, base = n - pend;
    if (n >= 8) store_group(base - 8, held);
    for (uint32_t k = 0; k < pend; k++) {
      if (base + k < cap) out[base + k] = (uint8_t)(acc >> (8 * (pend - 1 - k))); else err = AVR_E_OVERFLOW;
    }
  }
  AVR_HD void emit_put_digit() {
    if (low >> 63) { carry_back(); low &= (1ull << 63) - 1; }
    const uint32_t digit = (uint32_t)(low >> kTop);
    if (DIGIT_BITS == 16) push(digit >> 8);
    push(digit & 0xFF);
    low = (low & ((1ull << kTop) - 1)) << DIGIT_BITS;
    range <<= DIGIT_BITS;
  }
  AVR_HD void put(int symbol, uint64_t range_of_1) {            // arithmetic_code.h:106-126
    const uint64_t range_of_0 = range - range_of_1;
    low += symbol ? range_of_0 : 0ull;                          // selects: the lanes of a warp code different symbols
    range = symbol ? range_of_1 : range_of_0;
    if (range < kMinRange) {
      if (range == 0) { err = AVR_E_CORRUPT; range = 1; }       // "emitted a zero-probability symbol"
      while (range < (1ull << kTop)) emit_put_digit();
    }
  }
  AVR_HD void finish() {                                        // arithmetic_code.h:128-144
    for (uint64_t stop = 1ull << 62; stop > 0; stop >>= 1) {
      const uint64_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) {
      if (low >> 63) { carry_back(); low &= (1ull << 63) - 1; }
      push((uint32_t)(low >> 55));
      low = (low & ((1ull << 55) - 1)) << 8;
    }
    range = 0;
    flush();
  }
};

template <int DIGIT_BITS>
struct AvrStreamDec {
  static constexpr int kTop = 63 - DIGIT_BITS;
  static constexpr uint64_t kMinRange = 1ull << (kTop - 4);
  uint64_t low, range;
  const uint8_t* in;
  uint32_t pos, size, next_digit;
  AVR_HD uint32_t byte_or_zero() { const uint32_t b = pos < size ? in[pos] : 0u; pos++; return b; }   // arithmetic_code.h:281-285
  AVR_HD uint32_t aligned() {
    uint32_t d = byte_or_zero();
    if (DIGIT_BITS == 16) d = (d << 8) | byte_or_zero();
    return d;
  }
  AVR_HD void consume() {                                       // digit_alignment = 2: reads are one bit off (arithmetic_code.h:259-275)
    const uint32_t d = aligned();
    const uint32_t digit = ((next_digit << (DIGIT_BITS - 1)) | (d >> 1)) & ((1u << DIGIT_BITS) - 1);
    next_digit = d;
    low = (low << DIGIT_BITS) + digit;
    range <<= DIGIT_BITS;
  }
  AVR_HD void init(const uint8_t* p, uint32_t sz) {             // arithmetic_code.h:220-230
    in = p; pos = 0; size = sz;
    next_digit = aligned();
    low = next_digit >> 1;
    range = 1ull << (DIGIT_BITS - 1);
    while (range < (1ull << 63)) consume();
  }
  AVR_HD int get(uint64_t range_of_1) {                         // arithmetic_code.h:232-248
    const uint64_t range_of_0 = range - range_of_1;
    const int symbol = low >= range_of_0;
    low -= symbol ? range_of_0 : 0ull;
    range = symbol ? range_of_1 : range_of_0;
    if (range < kMinRange) {
      if (range == 0) return symbol;                            // corrupt input: the caller's loop is bounded by nbins
      while (range < (1ull << kTop)) consume();
    }
    return symbol;
  }
};

struct AvrStreamModel {
  uint32_t cell;                                                // pos | neg << 8
  AVR_HD void init() { cell = AVR_EST_INIT; }
  template <int MODEL>
  AVR_HD uint64_t p1(uint64_t range, const AvrDivTab* dt) const {
    if (MODEL == AVR_STREAM_FIXED_HALF) return range >> 1;
    const uint32_t pos = cell & 0xFF;
    return avr_range_of_1(range, pos, pos + (cell >> 8), dt);
  }
  template <int MODEL>
  AVR_HD void update(int symbol) { if (MODEL == AVR_STREAM_ADAPTIVE) cell = avr_est_update(cell, symbol, 0x60); }
};

// stream s of a batch: bits[s * nbins .. +nbins) (one byte per bin, 0/1) -> scratch[s * cap .. ), len[s]
template <int DIGIT_BITS, int MODEL>
static AVR_HD void avr_body_stream_encode(uint32_t s, const uint8_t* bits, uint32_t nbins, const AvrDivTab* dt, uint8_t* scratch,
                                          uint32_t cap, uint32_t* len, int* status) {
  const uint8_t* b = bits + (size_t)s * nbins;
  AvrStreamEnc<DIGIT_BITS> enc; enc.init(scratch + (size_t)s * cap, cap);
  AvrStreamModel m; m.init();
  uint32_t i = 0;
  // rows that start 8-byte aligned are read a word at a time
  if ((((size_t)s * nbins) & 7) == 0 && (((uintptr_t)bits) & 7) == 0) {
    for (; i + 8 <= nbins; i += 8) {
      uint64_t w = *reinterpret_cast<const uint64_t*>(b + i);
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const int sym = (int)(w & 0xFF) != 0; w >>= 8;
        enc.put(sym, m.template p1<MODEL>(enc.range, dt));
        m.template update<MODEL>(sym);
      }
    }
  }
  for (; i < nbins; i++) {
    const int sym = b[i] != 0;
    enc.put(sym, m.template p1<MODEL>(enc.range, dt));
    m.template update<MODEL>(sym);
  }
  enc.finish();
  len[s] = enc.n; status[s] = enc.err;
}

template <int DIGIT_BITS, int MODEL>
static AVR_HD void avr_body_stream_decode(uint32_t s, const uint8_t* coded, const uint64_t* coded_off, uint32_t nbins, const AvrDivTab* dt,
                                          uint8_t* bits_out) {
  AvrStreamDec<DIGIT_BITS> dec; dec.init(coded + coded_off[s], (uint32_t)(coded_off[s + 1] - coded_off[s]));
  AvrStreamModel m; m.init();
  uint8_t* o = bits_out + (size_t)s * nbins;
  uint32_t i = 0;
  if ((((size_t)s * nbins) & 7) == 0 && (((uintptr_t)bits_out) & 7) == 0) {
    for (; i + 8 <= nbins; i += 8) {
      uint64_t w = 0;
#pragma unroll
      for (int k = 0; k < 8; k++) {
        const int sym = dec.get(m.template p1<MODEL>(dec.range, dt));
        m.template update<MODEL>(sym);
        w |= (uint64_t)sym << (8 * k);
      }
      *reinterpret_cast<uint64_t*>(o + i) = w;
    }
  }
  for (; i < nbins; i++) {
    const int sym = dec.get(m.template p1<MODEL>(dec.range, dt));
    m.template update<MODEL>(sym);
    o[i] = (uint8_t)sym;
  }
}

// scratch bytes one stream can need: FIXED_HALF costs exactly one bit per bin; the adaptive estimator never goes
// below 1/97 (< 6.6 bits per bin); plus the finish() flush (arithmetic_code.h:139-142, at most 8 digits) and slack
static inline uint32_t avr_stream_cap(uint32_t nbins, int model) {
  const uint64_t c = model == AVR_STREAM_FIXED_HALF ? (uint64_t)nbins / 8 + 32 : (uint64_t)nbins + 32;
  return (uint32_t)((c + 15) & ~15ull);
}
