// Per-file sequential model ("chain"): the state the reference keeps in one h264_model
// instance for the life of a file (recode.cpp:625-1066) laid out for the device.
//   estimators : u16 cell per dense key (avr_common.h), file lifetime, never cleared
//                (reference std::map<model_key, estimator>, recode.cpp:1064-1065)
//   nnz store  : two ping-pong frames of 64 B per macroblock: [0..50] = num_nonzeros[n],
//                [51] = is_8x8 (reference FrameBuffer/BlockMeta, framebuffer.h, block.h:21-23;
//                only these fields are live)
// Used by the decompress kernel (one chain per file, inherently sequential: SURVEY.md 7.0)
// and by the single-thread compress kernel that serves as the on-device cross-check of the
// slice-parallel compress pipeline.
//
// Domain: every (macroblock, n) is visited at most once per picture and consecutive pictures
// differ in frame_num, i.e. the domain in which the reference itself round-trips (SURVEY.md
// 4.3); the recorder refuses other streams, so the 816 per-MB significance flags the
// reference keeps (block.h:5) reduce to the count of the block being coded.
#pragma once
#include "avr_device.h"

#define AVR_MB_STRIDE 64
#define AVR_MB_IS8X8 51

// significant_coeff_flag ctxIdxInc for frame-coded 8x8 blocks (ITU-T H.264 Table 9-43),
// the reference's sig_coeff_flag_offset_8x8[0] (recode.cpp:691-696), 4 bits per entry.
static AVR_HD int avr_sig8x8_off(int zz) {
  const int q = zz >> 4;   // selects, not an indexed stack array
  const uint64_t t = q == 0 ? 0x5544433445543210ull : q == 1 ? 0x789A987776334444ull : q == 2 ? 0xB689AE9876BDCB67ull : 0x0CAEBDCB9AE96BDCull;
  return (int)((t >> ((zz & 15) * 4)) & 15);
}
// zigzag offset used in the sig-map key (recode.cpp:704-712)
static AVR_HD int avr_zz_off(int zz, int size, int is_dc, int c422) {
  if (is_dc && c422) return zz < 2 ? 0 : zz < 4 ? 1 : 2;    // sig_coeff_offset_dc, recode.cpp:704
  if (size > 32) return avr_sig8x8_off(zz);
  return zz;
}

// get_neighbor_sub_mb (recode.cpp:426-478) on block coordinates instead of the scan_8 /
// reverse_scan_8 tables (:270-319): each of the three planes is a 4x4 grid of 4x4 blocks
// (n = 16*plane + blkIdx); the left neighbour of column 0 is column 3 of MB (x-1,y), the one
// above row 0 is row 3 of MB (x,y-1); blocks of 32+ coefficients round the neighbour down to
// a multiple of 4; DC blocks (n >= 48) use the same n of the neighbouring MB.
// Returns the neighbour's n; *cross = 1 if it lies in the neighbouring macroblock.
static AVR_HD int avr_neighbor_n(int n, int above, int size, int* cross) {
  if (n >= 48) { *cross = 1; return n; }
  int plane = n >> 4, i = n & 15;
  int x = (i & 1) | ((i >> 1) & 2), y = ((i >> 1) & 1) | ((i >> 2) & 2);
  *cross = 0;
  if (above) { if (y == 0) { *cross = 1; y = 3; } else y--; }
  else       { if (x == 0) { *cross = 1; x = 3; } else x--; }
  int nn = (plane << 4) + ((x & 1) | ((y & 1) << 1) | ((x >> 1) << 2) | ((y >> 1) << 3));
  if (size >= 32) nn &= ~3;
  return nn;
}

// Neighbour terms of the nnz side-channel key (recode.cpp:865-891), gathered once per block.
struct AvrNzCtx {
  int has_left, left_nz, above_nz, prev_nz, third;
};
static AVR_HD uint32_t avr_nz_bit_key(const AvrNzCtx& c, int i, int so_far) {
  int bit = 1 << i;
  int lb = c.has_left ? (c.left_nz >= bit) : 2;
  int ab = c.above_nz ? (c.above_nz >= bit) : 2;      // sic: tests the count, not has_above (recode.cpp:888)
  return avr_key_nz(i, so_far & (bit - 1), c.prev_nz >= bit, lb, ab, c.third);   // so_far < 2^i by construction; the mask keeps prefetch keys inside the line
}
// cur / prev = base of the 64-byte-per-MB stores of the current and previous picture (prev may
// be null = all zero).  Only valid after the block's own is_8x8 contribution has been stored.
static AVR_HD void avr_nz_ctx_x(AvrNzCtx* c, const uint8_t* cur, const uint8_t* prev, uint32_t mb_addr, int mb_x, int mb_w,
                                int n, int size, int cat, int is_dc, int c422) {
  int cross, nn;
  nn = avr_neighbor_n(n, 0, size, &cross);
  c->has_left = !(cross && mb_x == 0);
  c->left_nz = c->has_left ? cur[(size_t)(mb_addr - (cross ? 1u : 0u)) * AVR_MB_STRIDE + nn] : 0;
  nn = avr_neighbor_n(n, 1, size, &cross);
  int has_above = !(cross && mb_addr < (uint32_t)mb_w);
  c->above_nz = has_above ? cur[(size_t)(mb_addr - (cross ? (uint32_t)mb_w : 0u)) * AVR_MB_STRIDE + nn] : 0;
  c->prev_nz = prev ? prev[(size_t)mb_addr * AVR_MB_STRIDE + n] : 0;
  int is8 = cur[(size_t)mb_addr * AVR_MB_STRIDE + AVR_MB_IS8X8] || size > 32;
  c->third = is8 + 2 * is_dc + c422 + 4 * cat;
}

static AVR_HD void avr_nz_ctx(AvrNzCtx* c, const uint8_t* cur, const uint8_t* prev, uint32_t mb_addr, int mb_w,
                              int n, int size, int cat, int is_dc, int c422) {
  avr_nz_ctx_x(c, cur, prev, mb_addr, (int)(mb_addr % (uint32_t)mb_w), mb_w, n, size, cat, is_dc, c422);
}

struct AvrChain {
  uint16_t* est;        // [AVR_KEY_COUNT] (global memory)
  uint16_t* est_def;    // [AVR_NUM_CTX_KEYS] cells of the 1026 default keys: shared memory in the
                        // decompress kernel (most bins use them), == est otherwise
  uint8_t* store[2];    // [max_mbs * 64] each
  uint32_t max_mbs;
  int cur;
  int mb_w;
  const AvrDivTab* dt;

  AVR_HD void init(uint16_t* e, uint8_t* s0, uint8_t* s1, uint32_t mbs, const AvrDivTab* d) {
    est = e; est_def = e; store[0] = s0; store[1] = s1; max_mbs = mbs; cur = 0; mb_w = 1; dt = d;
  }
  AVR_HD uint16_t* cell(uint32_t key) { return key < AVR_NUM_CTX_KEYS ? est_def + key : est + key; }
  // update_frame_spec (recode.cpp:831-850) when it flips: the new current picture is zeroed;
  // with a size change both are (prev_valid = 0).
  AVR_HDN void flip(int mb_w_, int mb_h_, int prev_valid) {
    cur ^= 1;
    mb_w = mb_w_;
    zero_stores(store[cur], prev_valid ? nullptr : store[cur ^ 1], mb_w_, mb_h_, 0, 1);
  }
  // lane `lane` of `nlanes` zeroes its share (the decompress kernel calls it with the whole warp)
  static AVR_HDN void zero_stores(uint8_t* a8, uint8_t* b8, int mb_w_, int mb_h_, int lane, int nlanes) {
    size_t nb = (size_t)mb_w_ * mb_h_ * AVR_MB_STRIDE / 16;
    struct alignas(16) Q { uint64_t a, b; };
    Q z; z.a = z.b = 0;
    Q* a = reinterpret_cast<Q*>(a8);
    for (size_t i = lane; i < nb; i += nlanes) a[i] = z;
    if (b8) {
      Q* b = reinterpret_cast<Q*>(b8);
      for (size_t i = lane; i < nb; i += nlanes) b[i] = z;
    }
  }
  AVR_HD uint8_t* mb(uint32_t mb_addr) { return store[cur] + (size_t)mb_addr * AVR_MB_STRIDE; }

  AVR_HD void encode(AvrRangeEnc& enc, uint32_t key, int symbol, uint32_t limit) {
    uint16_t* c = cell(key);
    uint32_t cell = *c;
    uint32_t pos = cell & 0xFF, tot = pos + (cell >> 8);
    enc.put(symbol, avr_range_of_1(enc.range, pos, tot, dt));
    *c = (uint16_t)avr_est_update(cell, symbol, limit);
  }
  AVR_HD int decode(AvrRangeDec& dec, uint32_t key, uint32_t limit) {
    uint16_t* c = cell(key);
    uint32_t cell = *c;
    uint32_t pos = cell & 0xFF, tot = pos + (cell >> 8);
    int symbol = dec.get(avr_range_of_1(dec.range, pos, tot, dt));
    *c = (uint16_t)avr_est_update(cell, symbol, limit);
    return symbol;
  }
};
