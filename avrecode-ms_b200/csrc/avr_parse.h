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
#include <stddef.h>
#include "avr_bodies.h"
#include "avr_mem.h"

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
  uint32_t mbs;            // macroblocks the slice covers (the host checks that slices of one picture do not overlap)
};

// before the parse: the bytes that follow a payload in the pool become the bytes libavcodec's buffer holds there
static AVR_HD void avr_body_patch_tail(const AvrParseSlice& ps, uint8_t* raw) {
  for (int k = 0; k < 4; k++) raw[ps.raw_off + ps.size + (uint32_t)k] = ps.tail[k];
}

struct AvrParShared {
  uint64_t lut[64];                    // per pStateIdx: rangeTabLPS row (4 bytes) | transIdxLPS << 32, one load per decision
  uint64_t self;                       // shared-window address of this object, read back as an opaque value (avr_decode.h avr_sbase)
  uint8_t st[1024];
  uint8_t range_lps[256];              // [pStateIdx][qRangeIdx]; read as one 32-bit row per state
  uint8_t trans_lps[64];
  // CABAC decoding engine (ITU-T H.264 9.3.3.2) in left-aligned form: val bits 31..23 = codIOffset, the bits
  // below are read-ahead; "offset >= range" is val >= range << 23 and renormalisation is a plain shift
  uint32_t val, range, nextw;          // nextw = the 16 bits that follow the valid ones (loaded one refill ahead)
  int nvalid;                          // valid bits in val, from the top (>= 17 between bins)
  const uint8_t* in; uint32_t pos, size;
  // recorder (avr_record.h Recorder)
  uint16_t* bins; uint32_t nb, bin_cap, file_bin_base;
  AvrBlock* blocks; uint32_t nblk, blk_cap;
  uint32_t emitted, frame, mb_addr; int mb_w;
  int cat, n, max_coeff, is_dc;
};
#define AVR_P_OFF(member) ((avr_sptr)offsetof(AvrParShared, member))

#if defined(__CUDACC__)
// One AvrParShared per LANE in dynamic shared memory (avr_k_parse: slices shorter than the batch's long ones are packed several to
// a warp, each lane walking its own slice on its own instance).  The stride is 2 mod 32 words (the 64-bit table rows need 8-byte
// alignment): lanes that touch the same member spread over 16 banks.
#define AVR_PAR_STRIDE ((((uint32_t)sizeof(AvrParShared) + 127u) & ~127u) + 8u)
extern __shared__ __align__(16) uint8_t avr_par_dyn[];
static __device__ __forceinline__ AvrParShared* avr_par_self() {
  return reinterpret_cast<AvrParShared*>(avr_par_dyn + threadIdx.x * AVR_PAR_STRIDE);
}
#define g_par (*avr_par_self())
#define AVR_PAR_FN static __device__ __noinline__
#define AVR_PAR_INL static __device__ __forceinline__
#define AVR_PAR_M __device__ __forceinline__
#else
static AvrParShared g_par;
#define AVR_PAR_FN static
#define AVR_PAR_INL static inline
#define AVR_PAR_M inline
#endif

// tables of the engine; lane/nlanes split the work when a warp initialises them
AVR_PAR_INL void avr_par_tables(const AvrTables* tab, int lane, int nlanes) {
  for (int i = lane; i < 256; i += nlanes) g_par.range_lps[i] = tab->hot.range_lps[i];
  for (int i = lane; i < 64; i += nlanes) {
    g_par.trans_lps[i] = tab->hot.trans_lps[i];
    uint64_t row = 0;
    for (int q = 0; q < 4; q++) row |= (uint64_t)tab->hot.range_lps[(i << 2) | q] << (8 * q);
    g_par.lut[i] = row | ((uint64_t)tab->hot.trans_lps[i] << 32);
  }
  if (lane == 0) g_par.self = (uint64_t)avr_sp(&g_par);
}

// Engine registers: live in g_par between the parser's single-bin calls, in registers across a residual block.
struct AvrParRegs { uint32_t val, range, nextw, pos, nb, lim; int nvalid; const uint8_t* in; };
AVR_PAR_INL void avr_par_load(AvrParRegs& r) {
  r.val = g_par.val; r.range = g_par.range; r.nextw = g_par.nextw; r.pos = g_par.pos; r.nb = g_par.nb; r.nvalid = g_par.nvalid;
  r.in = g_par.in; r.lim = g_par.size + 6;
}
AVR_PAR_INL void avr_par_store(const AvrParRegs& r) {
  g_par.val = r.val; g_par.range = r.range; g_par.nextw = r.nextw; g_par.pos = r.pos; g_par.nb = r.nb; g_par.nvalid = r.nvalid;
}
// Two stream bytes at pos, big-endian.  The four bytes after the payload have been overwritten in the pool with what
// libavcodec's unescaped buffer holds there (avr_body_patch_tail), so the reader needs no end-of-payload cases; reads are
// clamped a few bytes past the end (a broken stream must not walk out of the pool; a valid one never gets there).
AVR_PAR_INL uint32_t avr_par_word(const AvrParRegs& r, uint32_t pos) {
  const uint32_t p = pos < r.lim ? pos : r.lim;
  return ((uint32_t)r.in[p] << 8) | r.in[p + 1];
}
// Rare (once per ~20 bins) and deliberately a call: inlined, ptxas predicates the body into every decision and
// schedules the state prefetches behind the whole decision chain (tools/sass_model.py: 115 -> 220 cycles per bin).
struct AvrParFill { uint32_t val, nextw, pos; };
AVR_PAR_FN AvrParFill avr_par_refill_slow(uint32_t val, uint32_t nextw, uint32_t pos, int nvalid, uint32_t lim, const uint8_t* in) {
  AvrParFill f;
  f.val = val | (nextw << (16 - nvalid));
  f.pos = pos + 2;
  const uint32_t p = f.pos < lim ? f.pos : lim;
  f.nextw = ((uint32_t)in[p] << 8) | in[p + 1];          // used one refill later: its latency stays off the bin chain
  return f;
}
AVR_PAR_INL void avr_par_refill(AvrParRegs& r) {
  if (r.nvalid <= 16) {
    const AvrParFill f = avr_par_refill_slow(r.val, r.nextw, r.pos, r.nvalid, r.lim, r.in);
    r.val = f.val; r.nextw = f.nextw; r.pos = f.pos;
    r.nvalid += 16;
  }
}
AVR_PAR_INL void avr_par_engine_init(AvrParRegs& r) {
  r.in = g_par.in; r.lim = g_par.size + 6;
  r.val = (avr_par_word(r, 0) << 16) | avr_par_word(r, 2);
  r.nvalid = 32; r.pos = 4; r.nextw = avr_par_word(r, 4);
  r.range = 510; r.nb = 0;
}
AVR_PAR_INL void avr_par_put(AvrParRegs& r, uint32_t rec) {
  const uint32_t i = r.nb < g_par.bin_cap ? r.nb : g_par.bin_cap - 1;   // past the capacity: keep counting, status says overflow
  g_par.bins[i] = (uint16_t)rec;
  r.nb++;
}
// DecodeDecision (9.3.3.2.1) on the state byte (pStateIdx << 1 | valMPS)
AVR_PAR_INL int avr_par_decision(AvrParRegs& r, int ctx) {
  const uint32_t s = g_par.st[ctx];
  const uint32_t row = reinterpret_cast<const uint32_t*>(g_par.range_lps)[s >> 1];
  const uint32_t lps = (row >> ((r.range >> 3) & 24)) & 0xFF;
  uint32_t range = r.range - lps;
  const uint32_t scaled = range << 23;
  uint32_t bin = s & 1, ns;
  if (r.val >= scaled) {
    r.val -= scaled; range = lps; bin ^= 1;
    ns = ((uint32_t)g_par.trans_lps[s >> 1] << 1) | ((s & 1) ^ (uint32_t)((s >> 1) == 0));
  } else {
    ns = s < 124 ? s + 2 : s;
  }
  g_par.st[ctx] = (uint8_t)ns;
  const int sh = AVR_CLZ32(range) - 23;          // 0 when range >= 256
  r.range = range << sh; r.val <<= sh; r.nvalid -= sh;
  avr_par_refill(r);
  return (int)bin;
}
AVR_PAR_INL int avr_par_bypass_r(AvrParRegs& r) {
  const uint32_t scaled = r.range << 22;         // (offset << 1 | bit) >= range  <=>  val >= range << 22
  int bin = 0;
  if (r.val >= scaled) { r.val -= scaled; bin = 1; }
  r.val <<= 1; r.nvalid -= 1;
  avr_par_refill(r);
  return bin;
}
AVR_PAR_INL int avr_par_terminate_r(AvrParRegs& r) {
  const uint32_t range = r.range - 2;
  if (r.val >= (range << 23)) { r.range = range; return 1; }
  const int sh = range < 256 ? 1 : 0;
  r.range = range << sh; r.val <<= sh; r.nvalid -= sh;
  avr_par_refill(r);
  return 0;
}

// ---- single-bin entry points for the syntax walk outside residual blocks ---------------------
AVR_PAR_FN int avr_par_get(int ctx) {
  AvrParRegs r; avr_par_load(r);
  const int bin = avr_par_decision(r, ctx);
  avr_par_put(r, AVR_BIN(bin, ctx, AVR_KIND_DEFAULT));
  avr_par_store(r);
  g_par.emitted++;
  return bin;
}
AVR_PAR_FN int avr_par_bypass() {
  AvrParRegs r; avr_par_load(r);
  const int bin = avr_par_bypass_r(r);
  avr_par_put(r, AVR_BIN(bin, AVR_CTX_BYPASS, AVR_KIND_DEFAULT));
  avr_par_store(r);
  g_par.emitted++;
  return bin;
}
AVR_PAR_FN int avr_par_terminate() {
  AvrParRegs r; avr_par_load(r);
  const int bin = avr_par_terminate_r(r);
  avr_par_put(r, AVR_BIN(bin, AVR_CTX_TERMINATE, AVR_KIND_DEFAULT));
  avr_par_store(r);
  g_par.emitted++;
  return bin;
}

// ---- one residual block (h264_cabac.c:1583-1780; recorder: Recorder::begin_sub_mb / sigmap in avr_record.h) ----
// Engine and recorder cursors stay in registers for the whole block; the block record is written once.
// State byte and table row of the NEXT significance context are requested while the current bin is decided
// (the context sequence of the map does not depend on the bins, only its length does).
struct AvrParCtx { uint32_t s; uint64_t lut; };          // state byte + table entry of its pStateIdx
AVR_PAR_INL void avr_par_ctx_load(avr_sptr s_st, avr_sptr s_lut, uint32_t ctx, AvrParCtx& c) {
  c.s = avr_s_ld8(s_st + ctx);
  c.lut = avr_s_ld64_ro(s_lut + ((c.s >> 1) << 3));
}
// DecodeDecision with the state already in registers; returns the bin, ns = new state byte
AVR_PAR_INL int avr_par_decide(AvrParRegs& r, const AvrParCtx& c, uint32_t& ns) {
  const uint32_t s = c.s, row = (uint32_t)c.lut, tl = (uint32_t)(c.lut >> 32);
  const uint32_t lps = (row >> ((r.range >> 3) & 24)) & 0xFF;
  const uint32_t rmps = r.range - lps;
  const uint32_t scaled = rmps << 23;
  const bool is_lps = r.val >= scaled;
  const uint32_t ns_lps = (tl << 1) | ((s & 1) ^ (uint32_t)(s < 2));   // pStateIdx 0: valMPS flips
  const uint32_t ns_mps = s < 124 ? s + 2 : s;
  ns = is_lps ? ns_lps : ns_mps;
  r.val -= is_lps ? scaled : 0u;
  const uint32_t range = is_lps ? lps : rmps;
  const int sh = AVR_CLZ32(range) - 23;                  // 0 when range >= 256
  r.range = range << sh; r.val <<= sh; r.nvalid -= sh;
  avr_par_refill(r);
  return (int)((s & 1) ^ (uint32_t)is_lps);
}
AVR_PAR_INL void avr_par_put_r(uint16_t* bins, uint32_t cap_m1, uint32_t& nb, uint32_t rec) {
  bins[nb < cap_m1 ? nb : cap_m1] = (uint16_t)rec;       // past the capacity: keep counting, status says overflow
  nb++;
}

template <bool IS64>
AVR_PAR_INL int avr_par_sigmap(AvrParRegs& r, avr_sptr s_st, avr_sptr s_lut, uint16_t* bins, uint32_t cap_m1, int max_coeff,
                               int sig_base, int last_base, int* last_out) {
  const int coefs = max_coeff - 1;
  int coeff_count = 0, last = 0;
  uint32_t cs = (uint32_t)(sig_base + (IS64 ? avr_sig8x8_off(0) : 0));
  AvrParCtx c; avr_par_ctx_load(s_st, s_lut, cs, c);
  for (;;) {
    const int nlast = last + 1;
    const uint32_t ncs = (uint32_t)(sig_base + (IS64 ? avr_sig8x8_off(nlast & 63) : nlast));   // a valid state index even past the end
    const uint32_t cl = (uint32_t)(last_base + (!IS64 ? last : last == 0 ? 0 : last < 16 ? 1 : last < 32 ? 2 : last < 40 ? 3 : last < 48 ? 4 : last < 52 ? 5 : last < 56 ? 6 : last < 60 ? 7 : 8));
    AvrParCtx cn, cq; avr_par_ctx_load(s_st, s_lut, ncs, cn); avr_par_ctx_load(s_st, s_lut, cl, cq);
    uint32_t ns;
    const int sig = avr_par_decide(r, c, ns);
    avr_s_st8(s_st + cs, ns);
    avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(sig, cs, AVR_KIND_SIG_MAP));
    if (IS64 && ncs == cs) { cn.s = ns; cn.lut = avr_s_ld64_ro(s_lut + ((ns >> 1) << 3)); }   // 8x8 offsets repeat: what was read is stale
    if (sig) {
      coeff_count++;
      uint32_t nq;
      const int eob = avr_par_decide(r, cq, nq);
      avr_s_st8(s_st + cl, nq);
      avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(eob, cl, AVR_KIND_SIG_EOB));
      if (eob) { last = max_coeff; break; }
    }
    last = nlast;
    if (nlast >= coefs) break;
    c = cn; cs = ncs;
  }
  *last_out = last;
  return coeff_count;
}

AVR_PAR_FN int avr_par_residual_block(int cat, int n, int max_coeff, int is_dc, int sig_base, int last_base, int abs_base) {
  AvrParRegs r; avr_par_load(r);
  uint16_t* const bins = g_par.bins;
  const uint32_t cap_m1 = g_par.bin_cap - 1;
  const avr_sptr s_base = (avr_sptr)g_par.self, s_st = s_base + AVR_P_OFF(st), s_lut = s_base + AVR_P_OFF(lut);
  const uint32_t nb0 = r.nb;
  int coeff_count = 0, last = 0;
  if (max_coeff > 1)
    coeff_count = max_coeff == 64 ? avr_par_sigmap<true>(r, s_st, s_lut, bins, cap_m1, max_coeff, sig_base, last_base, &last)
                                  : avr_par_sigmap<false>(r, s_st, s_lut, bins, cap_m1, max_coeff, sig_base, last_base, &last);
  const uint32_t map_bins = r.nb - nb0;
  uint32_t emitted = (uint32_t)avr_nz_bits(max_coeff) + map_bins - (uint32_t)coeff_count;   // EOB bins are inferred, not coded
  if (last == max_coeff - 1) coeff_count++;
  if (g_par.nblk < g_par.blk_cap) {
    AvrBlock b;
    b.bin_start = g_par.file_bin_base + nb0; b.mb_addr = g_par.mb_addr; b.frame = g_par.frame;
    b.n = (uint8_t)n; b.cat_flags = (uint8_t)(cat | (is_dc << 4)); b.max_coeff = (uint8_t)max_coeff;
    b.nbins = (uint8_t)map_bins;
    g_par.blocks[g_par.nblk] = b;
  }
  g_par.nblk++;
  // coeff_abs_level_minus1 + sign, reverse scan order (h264_cabac.c:1727-1768)
  const uint32_t lvl_start = r.nb;
  int node_ctx = 0;
  const int gt1_max = cat == 3 ? 8 : 9;
  for (int c = coeff_count; c > 0; c--) {
    const uint32_t ctx1 = (uint32_t)(abs_base + (node_ctx < 4 ? node_ctx + 1 : 0));
    const uint32_t ctx2 = (uint32_t)(abs_base + (node_ctx < 4 ? 5 : (node_ctx + 2 > gt1_max ? gt1_max : node_ctx + 2)));
    AvrParCtx c1, c2; avr_par_ctx_load(s_st, s_lut, ctx1, c1); avr_par_ctx_load(s_st, s_lut, ctx2, c2);   // never the same state byte
    uint32_t ns;
    const int b0 = avr_par_decide(r, c1, ns);
    avr_s_st8(s_st + ctx1, ns);
    avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(b0, ctx1, AVR_KIND_DEFAULT));
    if (b0 == 0) {
      node_ctx = node_ctx < 3 ? node_ctx + 1 : node_ctx;
    } else {
      node_ctx = node_ctx < 4 ? 4 : (node_ctx < 7 ? node_ctx + 1 : 7);
      int coeff_abs = 2;
      while (coeff_abs < 15) {                           // the ctx2 state stays in registers across its run
        const int b = avr_par_decide(r, c2, ns);
        c2.s = ns; c2.lut = avr_s_ld64_ro(s_lut + ((ns >> 1) << 3));
        avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(b, ctx2, AVR_KIND_DEFAULT));
        if (!b) break;
        coeff_abs++;
      }
      avr_s_st8(s_st + ctx2, c2.s);
      if (coeff_abs >= 15) {
        int j = 0;
        for (;;) {
          const int b = avr_par_bypass_r(r);
          avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(b, AVR_CTX_BYPASS, AVR_KIND_DEFAULT));
          if (!b || j >= 30) break;
          j++;
        }
        while (j--) avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(avr_par_bypass_r(r), AVR_CTX_BYPASS, AVR_KIND_DEFAULT));
      }
    }
    avr_par_put_r(bins, cap_m1, r.nb, AVR_BIN(avr_par_bypass_r(r), AVR_CTX_BYPASS, AVR_KIND_DEFAULT));
  }
  emitted += r.nb - lvl_start;
  g_par.emitted += emitted;
  avr_par_store(r);
  return coeff_count;
}

struct AvrParsePolicy {
  static constexpr bool kFastResidual = true;
  AVR_PAR_M int residual_block(int cat, int n, int mc, int dc, int sig, int last, int abs_) { return avr_par_residual_block(cat, n, mc, dc, sig, last, abs_); }
  AVR_PAR_M int get(int ctx) { return avr_par_get(ctx); }
  AVR_PAR_M int bypass(int) { return avr_par_bypass(); }
  AVR_PAR_M int terminate(int) { return avr_par_terminate(); }
  AVR_PAR_M void mb_begin(int x, int y) { g_par.mb_addr = (uint32_t)(x + y * g_par.mb_w); }
  AVR_PAR_M void sub_mb_begin(int, int, int, int, int) {}
  AVR_PAR_M void sub_mb_end() {}
  AVR_PAR_M void sigmap_begin() {}
  AVR_PAR_M void sigmap_end() {}
  AVR_PAR_M void hint(int, int) {}
};

// One slice on one thread.  g_par.st / tables must be initialised by the caller (cooperatively in
// the kernel).  row = [mb_w] MbInfo scratch private to this slice (neighbours never cross slices).
template <bool HAS_B>
AVR_PAR_FN void avr_parse_slice(const AvrParseSlice& ps, const uint8_t* raw, uint16_t* bins_pool, AvrBlock* blk_pool,
                                avr::MbInfo* row, AvrParseResult* res) {
  g_par.in = raw + ps.raw_off; g_par.size = ps.size;
  g_par.bins = bins_pool + ps.bin_off; g_par.bin_cap = ps.bin_cap; g_par.file_bin_base = ps.file_bin_base;
  g_par.blocks = blk_pool + ps.blk_off; g_par.nblk = 0; g_par.blk_cap = ps.blk_cap;
  g_par.emitted = 0; g_par.frame = ps.frame; g_par.mb_addr = 0; g_par.mb_w = ps.sp.mb_w;
  { AvrParRegs r; avr_par_engine_init(r); avr_par_store(r); }
  for (int x = 0; x < ps.sp.mb_w; x++) row[x].uid = 0;
  AvrParsePolicy io;
  int rc;
  { avr::SliceParser<AvrParsePolicy, HAS_B> parser(io, ps.sp, row, 1); rc = parser.run(); }
  // the walk must end on rbsp_stop_one_bit (avr_record.cpp record_annexb, cabac_std.h bits_consumed): bits in val count from the
  // top, the 9-bit codIOffset window is read
  const uint32_t consumed = 8u * g_par.pos - (uint32_t)g_par.nvalid + 9u;
  if (rc >= 0 && (consumed + 6u) / 8u != ps.size) rc = avr::SYN_E_LENGTH;
  // ... and through end_of_slice_flag = 1, not by running out of macroblocks: the decompress chain could not close such a slice
  if (rc >= 0 && g_par.nb >= 1 && g_par.nb <= g_par.bin_cap && g_par.bins[g_par.nb - 1] != AVR_BIN(1, AVR_CTX_TERMINATE, AVR_KIND_DEFAULT))
    rc = avr::SYN_E_LENGTH;
  res->nbins = g_par.nb; res->nblocks = g_par.nblk; res->emitted = g_par.emitted; res->mbs = rc > 0 ? (uint32_t)rc : 0u;
  const int over = g_par.nb > g_par.bin_cap || g_par.nblk > g_par.blk_cap;
  res->status = over ? AVR_E_OVERFLOW : rc == avr::SYN_E_PCM ? AVR_E_UNSUPPORTED : rc < 0 ? AVR_E_CORRUPT : AVR_OK;
}
