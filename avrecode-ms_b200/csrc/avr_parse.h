// Compress-side front end ON THE DEVICE: standard CABAC decode of the original slice bytes +
// slice_data() walk, one slice per warp (slices are independent for CABAC: states re-initialise
// per slice, ffmpeg h264_slice.c:2368), emitting the same record arrays the host recorder emits
// (avr_record.h).  In the reference this is libavcodec's parse with compressor::cabac_decoder's
// private ff_get_cabac (recode.cpp:1156-1159, 1182-1199) -- host work ("next"
// row 8f-1).  It moved here because at 27 Mbins/s per host core it was 90 % of the end-to-end
// time once the coding itself ran on the GPU; the host keeps only NAL scanning, header parsing and
// the block-finder memmem (avr_files.cpp).  The host parser (avr_record.cpp) stays as the
// fallback for slices whose records do not fit the per-slice capacity.
//
// Include only from avr_kernels.cu and tests/host_emul.cpp.
#pragma once
#include "avr_bodies.h"

// One coded slice to parse.
struct AvrParseSlice {
  uint64_t raw_off;        // first CABAC byte in the raw-file pool
  uint32_t size;           // CABAC payload bytes (what libavcodec reports)
  uint8_t tail[4];         // unescaped bytes that follow the payload in libavcodec's buffer (reader may peek)
  uint64_t bin_off;        // where this slice's bins go in the bins pool (file-local index = bin_off - fd.bin_off)
  uint32_t bin_cap;
  uint32_t blk_cap;
  uint64_t blk_off;        // where this slice's blocks go in the capacity block pool
  uint32_t file_bin_base;  // file-local index of the slice's first bin (AvrBlock::bin_start is file-local)
  uint32_t frame;          // picture (flip) index within the file
  avr::SliceParams sp;
};
struct AvrParseResult {
  uint32_t nbins, nblocks, emitted;
  int status;
};

struct AvrParShared {
  uint8_t st[1024];
  uint8_t range_lps[256];
  uint8_t trans_lps[64];
  // standard CABAC decoding engine (ITU-T H.264 9.3.3.2)
  uint32_t range, offset;
  uint64_t bitbuf; int nbits;          // unread bits, right-aligned
  const uint8_t* in; uint32_t pos, size; uint32_t tail;
  // recorder (avr_record.h Recorder)
  uint16_t* bins; uint32_t nb, bin_cap, file_bin_base;
  AvrBlock* blocks; uint32_t nblk, blk_cap;
  uint32_t emitted, frame, mb_addr; int mb_w;
  int cat, n, max_coeff, is_dc;
  int in_sigmap, expect_eob, overflow;
};

#if defined(__CUDACC__)
__shared__ AvrParShared g_par;
#define AVR_PAR_FN static __device__ __noinline__
#define AVR_PAR_INL static __device__ __forceinline__
#define AVR_PAR_M __device__ __forceinline__
#else
static AvrParShared g_par;
#define AVR_PAR_FN static
#define AVR_PAR_INL static inline
#define AVR_PAR_M inline
#endif

AVR_PAR_INL uint32_t avr_par_byte(uint32_t pos) {
  if (pos < g_par.size) return g_par.in[pos];
  const uint32_t k = pos - g_par.size;
  return k < 4 ? (g_par.tail >> (8 * k)) & 0xFF : 0u;
}
AVR_PAR_INL uint32_t avr_par_bits(int n) {
  if (n == 0) return 0;
  while (g_par.nbits < n) {
    g_par.bitbuf = (g_par.bitbuf << 8) | avr_par_byte(g_par.pos++);
    g_par.nbits += 8;
  }
  g_par.nbits -= n;
  return (uint32_t)(g_par.bitbuf >> g_par.nbits) & ((1u << n) - 1);
}
AVR_PAR_INL void avr_par_emit(int ctx, int sym) {
  int kind = AVR_KIND_DEFAULT;
  if (g_par.in_sigmap) {
    kind = g_par.expect_eob ? AVR_KIND_SIG_EOB : AVR_KIND_SIG_MAP;
    g_par.expect_eob = !g_par.expect_eob && sym;
    if (g_par.nblk <= g_par.blk_cap && g_par.nblk > 0) g_par.blocks[g_par.nblk - 1].nbins++;
  }
  if (g_par.nb < g_par.bin_cap) g_par.bins[g_par.nb] = AVR_BIN(sym, ctx, kind); else g_par.overflow = 1;
  g_par.nb++;
  if (kind != AVR_KIND_SIG_EOB) g_par.emitted++;
}
AVR_PAR_FN int avr_par_get(int ctx) {
  uint32_t s = g_par.st[ctx], p = s >> 1, mps = s & 1;
  uint32_t range = g_par.range, offset = g_par.offset;
  const uint32_t lps = g_par.range_lps[(p << 2) | ((range >> 6) & 3)];
  range -= lps;
  int bin;
  if (offset >= range) {
    bin = (int)(mps ^ 1); offset -= range; range = lps;
    if (p == 0) mps ^= 1;
    p = g_par.trans_lps[p];
  } else {
    bin = (int)mps;
    if (p < 62) p++;
  }
  g_par.st[ctx] = (uint8_t)((p << 1) | mps);
  if (range < 256) {
    const int sh = AVR_CLZ32(range) - 23;
    range <<= sh;
    offset = (offset << sh) | avr_par_bits(sh);
  }
  g_par.range = range; g_par.offset = offset;
  avr_par_emit(ctx, bin);
  return bin;
}
AVR_PAR_FN int avr_par_bypass() {
  uint32_t offset = (g_par.offset << 1) | avr_par_bits(1);
  int bin = 0;
  if (offset >= g_par.range) { offset -= g_par.range; bin = 1; }
  g_par.offset = offset;
  avr_par_emit(AVR_CTX_BYPASS, bin);
  return bin;
}
AVR_PAR_FN int avr_par_terminate() {
  uint32_t range = g_par.range - 2;
  int bin;
  if (g_par.offset >= range) bin = 1;
  else {
    bin = 0;
    if (range < 256) { range <<= 1; g_par.offset = (g_par.offset << 1) | avr_par_bits(1); }
  }
  g_par.range = range;
  avr_par_emit(AVR_CTX_TERMINATE, bin);
  return bin;
}
AVR_PAR_FN void avr_par_sigmap_begin() {
  if (g_par.nblk < g_par.blk_cap) {
    AvrBlock b;
    b.bin_start = g_par.file_bin_base + g_par.nb;
    b.mb_addr = g_par.mb_addr;
    b.frame = g_par.frame;
    b.n = (uint8_t)g_par.n;
    b.cat_flags = (uint8_t)(g_par.cat | (g_par.is_dc << 4));
    b.max_coeff = (uint8_t)g_par.max_coeff;
    b.nbins = 0;
    g_par.blocks[g_par.nblk] = b;
  } else g_par.overflow = 1;
  g_par.nblk++;
  g_par.in_sigmap = 1; g_par.expect_eob = 0;
  g_par.emitted += (uint32_t)avr_nz_bits(g_par.max_coeff);
}

struct AvrParsePolicy {
  static constexpr bool kFastResidual = false;
  AVR_PAR_M int residual_block(int, int, int, int, int, int, int) { return 0; }
  AVR_PAR_M int get(int ctx) { return avr_par_get(ctx); }
  AVR_PAR_M int bypass(int) { return avr_par_bypass(); }
  AVR_PAR_M int terminate(int) { return avr_par_terminate(); }
  AVR_PAR_M void mb_begin(int x, int y) { g_par.mb_addr = (uint32_t)(x + y * g_par.mb_w); }
  AVR_PAR_M void sub_mb_begin(int cat, int n, int mc, int dc, int) { g_par.cat = cat; g_par.n = n; g_par.max_coeff = mc; g_par.is_dc = dc; }
  AVR_PAR_M void sub_mb_end() {}
  AVR_PAR_M void sigmap_begin() { avr_par_sigmap_begin(); }
  AVR_PAR_M void sigmap_end() { g_par.in_sigmap = 0; }
  AVR_PAR_M void hint(int, int) {}
};

// One slice on one thread.  g_par.st / tables must be initialised by the caller (cooperatively in
// the kernel).  row = [mb_w] MbInfo scratch private to this slice (neighbours never cross slices).
AVR_PAR_FN void avr_parse_slice(const AvrParseSlice& ps, const uint8_t* raw, uint16_t* bins_pool, AvrBlock* blk_pool,
                                avr::MbInfo* row, AvrParseResult* res) {
  g_par.in = raw + ps.raw_off; g_par.pos = 0; g_par.size = ps.size;
  g_par.tail = (uint32_t)ps.tail[0] | ((uint32_t)ps.tail[1] << 8) | ((uint32_t)ps.tail[2] << 16) | ((uint32_t)ps.tail[3] << 24);
  g_par.bitbuf = 0; g_par.nbits = 0;
  g_par.range = 510; g_par.offset = avr_par_bits(9);
  g_par.bins = bins_pool + ps.bin_off; g_par.nb = 0; g_par.bin_cap = ps.bin_cap; g_par.file_bin_base = ps.file_bin_base;
  g_par.blocks = blk_pool + ps.blk_off; g_par.nblk = 0; g_par.blk_cap = ps.blk_cap;
  g_par.emitted = 0; g_par.frame = ps.frame; g_par.mb_addr = 0; g_par.mb_w = ps.sp.mb_w;
  g_par.in_sigmap = 0; g_par.expect_eob = 0; g_par.overflow = 0;
  for (int x = 0; x < ps.sp.mb_w; x++) row[x].uid = 0;
  AvrParsePolicy io;
  avr::SliceParser<AvrParsePolicy> parser(io, ps.sp, row, 1);
  const int rc = parser.run();
  res->nbins = g_par.nb; res->nblocks = g_par.nblk; res->emitted = g_par.emitted;
  res->status = g_par.overflow ? AVR_E_OVERFLOW : rc == avr::SYN_E_PCM ? AVR_E_UNSUPPORTED : rc < 0 ? AVR_E_CORRUPT : AVR_OK;
}
