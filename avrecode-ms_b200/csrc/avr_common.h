// Shared host/device definitions for the avrecode entropy hot path (B200 build).
//
// Record format: what the host-side parse (libavcodec-hooks recorder, or the native
// slice_data parser in h264_syntax.h) hands to the device compress path.  It restates, as
// flat arrays, the event stream the reference's hooks deliver one callback at a time
// (reference ffmpeg/libavcodec/coding_hooks.h:27-55; recode.cpp:150-215).
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define AVR_HD __host__ __device__ __forceinline__
#define AVR_HDN __host__ __device__
#else
#define AVR_HD inline
#define AVR_HDN
#endif

// ---- bin record (u16) --------------------------------------------------------------------
// bit 0      : symbol
// bits 1..11 : context code: 0..1023 = ctxIdx (offset of the state byte in
//              H264SliceContext::cabac_state, reference ffmpeg/libavcodec/h264.h:511),
//              1024 = bypass, 1025 = terminate  (reference recode.cpp:1056,1190,1197)
// bits 12..13: kind: 0 = coded immediately (default key), 1 = significance-map MAP bin,
//              2 = significance-map EOB bin (queued on compress; never range-coded)
#define AVR_CTX_BYPASS 1024
#define AVR_CTX_TERMINATE 1025
#define AVR_NUM_CTX_KEYS 1026
#define AVR_KIND_DEFAULT 0
#define AVR_KIND_SIG_MAP 1
#define AVR_KIND_SIG_EOB 2
#define AVR_BIN(sym, ctx, kind) ((uint16_t)(((sym) & 1) | ((ctx) << 1) | ((kind) << 12)))
#define AVR_BIN_SYM(b) ((b) & 1)
#define AVR_BIN_CTX(b) (((b) >> 1) & 0x7FF)
#define AVR_BIN_KIND(b) (((b) >> 12) & 3)

// ---- residual block record (16 B): one per begin_sub_mb..end_sub_mb (h264_cabac.c:1652-1779)
typedef struct AvrBlock {
  uint32_t bin_start;   // index (file-global) of the block's first significance-map bin
  uint32_t mb_addr;     // mb_x + mb_y * mb_width   (mb_xy hook, h264_slice.c:2380)
  uint32_t frame;       // index into AvrFrame[] (one per frame-buffer flip, recode.cpp:831-850)
  uint8_t n;            // scan8 block index 0..50 (h264.h:1001-1002)
  uint8_t cat_flags;    // bits0-3 ctxBlockCat 0..13, bit4 is_dc, bit5 chroma422
  uint8_t max_coeff;    // 4, 8, 15, 16 or 64
  uint8_t nbins;        // number of significance-map bins (MAP + EOB) recorded, <= 126
} AvrBlock;

// ---- coded slice record: one per init_decoder whose block was found (recode.cpp:1143-1164)
typedef struct AvrSlice {
  uint32_t bin_start, bin_end;      // [start,end) in the file's bin array (terminate(1) included)
  uint32_t block_start, block_end;  // [start,end) in the file's AvrBlock array
} AvrSlice;

// ---- frame record: one per update_frame_spec() that flipped (recode.cpp:831-850)
typedef struct AvrFrame {
  uint16_t mb_w, mb_h;
  uint32_t prev_valid;  // 1 if frames[!cur] holds the previous flip with the same dims
  uint32_t mb_base;     // macroblocks of the file's earlier pictures (slot of this picture's store)
} AvrFrame;

// ---- estimator key space (dense re-indexing of reference model_key, recode.cpp:325,683-822,891)
// default keys   : [0, 1026)                    key = ctx code
// sig-map keys   : AVR_KEY_SIG_BASE + (row_base(cat) + row(stored, obs)) * 16 + zz_off, zz_off < 16
//                  one (cat, stored, obs) row = 16 cells = one 32-byte sector; a block walks its rows in order (obs -> obs + 1
//                  is the next row).  Only rows the model can reach exist (round 1 kept 14 x 64 x 64 rows = 1.8 MB per file,
//                  most of them unreachable; a chain's working set then missed L1 and the batch's pools missed L2):
//                    cats with <= 16 coefficients (all but 5, 9, 13): stored in 1..16, obs in 0..16 -> 17 x 17 rows
//                    8x8 cats (5, 9, 13): stored = 0 is a FULL block (six-bit count, recode.cpp:928-931) with obs in 0..63;
//                    stored in 1..63 has obs < stored (the EOB bin that follows one number `stored` is inferred and ends the
//                    block, recode.cpp:1444-1445) -> 64 + 63*64/2 = 2080 rows
// nnz-bit keys   : AVR_KEY_NZ_BASE + ((third*18 + nb18)*64 + (2^i + so_far))
//                  third = is_8x8 + 2*is_dc + c422 + 4*cat  (<64), nb18 = prev + 2*left + 6*above (each of
//                  left/above in 0..2), i<6, so_far<2^i: level i of a count's bit tree is the 2^i cells that start at
//                  cell 2^i of the line (naturally aligned; cell 0 unused); one 128-byte line per neighbour class
// Bases are multiples of 64 cells and the per-file pool is a multiple of 128 bytes, so rows never straddle sectors.
// 18 key bits: a compress sort key is (file << 19) | (key << 1) | symbol, which leaves 13 file bits in 32.
#define AVR_SIG_ROWS_SMALL (17u * 17u)
#define AVR_SIG_ROWS_8X8 (64u + 63u * 32u)
#define AVR_KEY_SIG_BASE 1088u
#define AVR_KEY_SIG_COUNT ((11u * AVR_SIG_ROWS_SMALL + 3u * AVR_SIG_ROWS_8X8) * 16u)
#define AVR_KEY_NZ_BASE ((AVR_KEY_SIG_BASE + AVR_KEY_SIG_COUNT + 63u) & ~63u)
#define AVR_KEY_NZ_COUNT (64u * 18u * 64u)
#define AVR_KEY_COUNT (AVR_KEY_NZ_BASE + AVR_KEY_NZ_COUNT)
#define AVR_KEY_BITS 18
#define AVR_KEY_MASK ((1u << AVR_KEY_BITS) - 1u)
#define AVR_MAX_BATCH_FILES 4096          /* sort key: 13 file bits; kernels are sized for this many chains per launch */

static AVR_HD int avr_cat_is_8x8(int cat) { return cat == 5 || cat == 9 || cat == 13; }
// first row of a category
static AVR_HD uint32_t avr_sig_row_base(int cat) {
  const uint32_t n8 = (uint32_t)(cat > 5) + (uint32_t)(cat > 9);
  return ((uint32_t)cat - n8) * AVR_SIG_ROWS_SMALL + n8 * AVR_SIG_ROWS_8X8;
}
// row of (stored, obs) inside its category; out-of-model arguments are folded into the category (memory safety only)
static AVR_HD uint32_t avr_sig_row(int cat, int stored, int obs) {
  uint32_t s = (uint32_t)stored & 63u, o = (uint32_t)obs & 63u;
  if (avr_cat_is_8x8(cat)) {
    if (s == 0) return o;
    if (o >= s) o = s - 1;
    return 64u + s * (s - 1u) / 2u + o;
  }
  if (s > 16u) s = 16u;
  if (o > 16u) o = 16u;
  return s * 17u + o;
}
static AVR_HD uint32_t avr_key_sig(int cat, int nnz_stored, int obs, int zz_off) {
  return AVR_KEY_SIG_BASE + (avr_sig_row_base(cat) + avr_sig_row(cat, nnz_stored, obs)) * 16u + ((uint32_t)zz_off & 15u);
}
static AVR_HD uint32_t avr_key_nz(int i, int so_far, int prev, int left, int above, int third) {
  return AVR_KEY_NZ_BASE + (((uint32_t)third * 18u + (uint32_t)(prev + 2 * left + 6 * above)) * 64u + ((1u << i) + (uint32_t)so_far));
}
// halving threshold by key class (reference recode.cpp:1048-1049): sig-map keys 0x50, all others 0x60
static AVR_HD int avr_key_is_sig(uint32_t key) { return key >= AVR_KEY_SIG_BASE && key < AVR_KEY_NZ_BASE; }

// nnz side-channel width (reference recode.cpp:865)
static AVR_HD int avr_nz_bits(int max_coeff) { return max_coeff > 16 ? 6 : max_coeff > 4 ? 4 : 2; }

// error codes (C-ABI, include/avrecode_b200.h)
#define AVR_OK 0
#define AVR_E_ARG -1
#define AVR_E_CUDA -2
#define AVR_E_OVERFLOW -3
#define AVR_E_UNSUPPORTED -4
#define AVR_E_CORRUPT -5
#define AVR_E_NOMEM -6
