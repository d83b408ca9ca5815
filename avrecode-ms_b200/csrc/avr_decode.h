// Decompress chain: one per file, strictly sequential per bin (reference recode.cpp:1442-1499:
// range-decode or infer the bin, re-encode it as CABAC, update the model, hand the symbol back
// to the slice_data() walk which picks the next context).
//
// Layout decision (from the first ncu capture, profiles/r01_*): with the bin I/O inlined at each
// of the ~150 call sites of the syntax walk the kernel was 51 k SASS instructions (instruction
// cache misses, `stall_no_inst`) and the coder state lived behind generic pointers (20 LD + 10 ST
// + 32 R2UR per bin).  So: the whole per-chain state is ONE shared-memory object per block
// (explicit LDS/STS, constant addresses), and get / bypass / terminate / sigmap_begin are single
// non-inlined functions that work on it; the syntax walk calls them through an empty policy.
//
// Include only from avr_kernels.cu (device: state in __shared__, one chain per block) and from
// tests/host_emul.cpp (host: the same code on a static object).
#pragma once
#include "avr_bodies.h"

struct AvrDecShared {
  AvrHotTabs hot;                              // per-bin constant tables
  uint16_t est_def[AVR_NUM_CTX_KEYS + 2];      // estimators of the 1026 default keys (file lifetime)
  uint8_t st[1024];                            // CABAC context states of the current slice
  // range decoder (arithmetic_code.h:220-288)
  uint64_t low, range, w_cur, w_next;
  const uint8_t* in;
  uint32_t pos, size, next_digit;
  // CABAC encoder (cabac_code.h:30-67)
  uint32_t cab_low, cab_range, cab_n, cab_cap;
  uint8_t* cab_out;
  int cab_err;
  // model (h264_model, recode.cpp:625-1066)
  uint16_t* est;                               // global: sig-map and nnz-bit estimators
  uint8_t* store[2];
  int cur, mb_w;
  uint32_t mb_addr;
  int cat, n, size_blk, is_dc, c422;
  int in_sigmap, expect_eob, zz, observed, stored, ones, ended_by_eob;
  uint32_t bins;
  int done;
};

#if defined(__CUDACC__)
__shared__ AvrDecShared g_dec;
#define AVR_DEC_FN static __device__ __noinline__
#define AVR_DEC_INL static __device__ __forceinline__
#define AVR_DEC_M __device__ __forceinline__
#else
static AvrDecShared g_dec;
#define AVR_DEC_FN static
#define AVR_DEC_INL static inline
#define AVR_DEC_M inline
#endif

// ---- range decoder on g_dec -----------------------------------------------------------------
AVR_DEC_INL uint64_t avr_dec_load_word(uint32_t off) {
  if (off >= g_dec.size) return 0;
  uint64_t w;
#if defined(__CUDA_ARCH__)
  w = __ldg(reinterpret_cast<const unsigned long long*>(g_dec.in + off));
#else
  w = 0;
  for (int k = 7; k >= 0; k--) w = (w << 8) | g_dec.in[off + (uint32_t)k];
#endif
  uint32_t valid = g_dec.size - off;
  if (valid < 8) w &= (1ull << (8 * valid)) - 1;
  return w;
}
AVR_DEC_INL void avr_dec_consume(uint64_t& low, uint64_t& range) {
  uint32_t pos = g_dec.pos;
  if ((pos & 7) == 0 && pos) { g_dec.w_cur = g_dec.w_next; g_dec.w_next = avr_dec_load_word(pos + 8); }
  uint32_t d = (uint32_t)(g_dec.w_cur >> (8 * (pos & 7))) & 0xFF;
  g_dec.pos = pos + 1;
  uint32_t digit = ((g_dec.next_digit << 7) | (d >> 1)) & 0xFF;
  g_dec.next_digit = d;
  low = (low << 8) + digit;
  range <<= 8;
}
AVR_DEC_INL void avr_dec_init_range(const uint8_t* p, uint32_t sz) {
  g_dec.in = p; g_dec.pos = 0; g_dec.size = sz;
  g_dec.w_cur = avr_dec_load_word(0); g_dec.w_next = avr_dec_load_word(8);
  g_dec.next_digit = (uint32_t)g_dec.w_cur & 0xFF; g_dec.pos = 1;
  uint64_t low = g_dec.next_digit >> 1, range = 128;
  while (range < (1ull << 63)) avr_dec_consume(low, range);
  g_dec.low = low; g_dec.range = range;
}
// decode one symbol with estimator cell `cell`; returns the symbol, updates low/range
AVR_DEC_INL int avr_dec_symbol(uint32_t cell) {
  uint64_t low = g_dec.low, range = g_dec.range;
  const uint32_t pos = cell & 0xFF, tot = pos + (cell >> 8);
  const uint64_t r1 = (AVR_MULHI64(range, g_dec.hot.div.m[tot]) >> g_dec.hot.div.s[tot]) * (uint64_t)pos;
  const uint64_t r0 = range - r1;
  const int symbol = low >= r0;
  if (symbol) { low -= r0; range = r1; } else { range = r0; }
  if (range < (1ull << 51) && range != 0)
    while (range < (1ull << 55)) avr_dec_consume(low, range);
  g_dec.low = low; g_dec.range = range;
  return symbol;
}
AVR_DEC_INL int avr_dec_default(uint32_t key) {       // key < AVR_NUM_CTX_KEYS: estimator in shared memory
  const uint32_t cell = g_dec.est_def[key];
  const int symbol = avr_dec_symbol(cell);
  g_dec.est_def[key] = (uint16_t)avr_est_update(cell, symbol, 0x60);
  return symbol;
}
AVR_DEC_INL int avr_dec_global(uint32_t key, uint32_t limit) {
  uint16_t* c = g_dec.est + key;
  const uint32_t cell = *c;
  const int symbol = avr_dec_symbol(cell);
  *c = (uint16_t)avr_est_update(cell, symbol, limit);
  return symbol;
}

// ---- CABAC encoder on g_dec -----------------------------------------------------------------
AVR_DEC_INL void avr_cab_push(uint32_t b) {
  if (g_dec.cab_n < g_dec.cab_cap) g_dec.cab_out[g_dec.cab_n] = (uint8_t)b; else g_dec.cab_err = AVR_E_OVERFLOW;
  g_dec.cab_n++;
}
AVR_DEC_INL void avr_cab_carry() {
  uint32_t i = g_dec.cab_n < g_dec.cab_cap ? g_dec.cab_n : g_dec.cab_cap;
  while (i > 0) { if (++g_dec.cab_out[--i] != 0) break; }
}
AVR_DEC_INL void avr_cab_put_raw(int symbol, uint32_t range_of_1) {
  uint32_t low = g_dec.cab_low, range = g_dec.cab_range;
  const uint32_t range_of_0 = range - range_of_1;
  if (symbol) { low += range_of_0; range = range_of_1; } else { range = range_of_0; }
  if (range < 0x200u) {
    if (range == 0) { g_dec.cab_err = AVR_E_CORRUPT; range = 1; }
    while (range < 0x8000u) {                               // 16-bit digits, arithmetic_code.h:113-122
      if (low >> 31) { avr_cab_carry(); low &= 0x7FFFFFFFu; }
      const uint32_t digit = low >> 15;
      avr_cab_push(digit >> 8); avr_cab_push(digit & 0xFF);
      low = (low & 0x7FFFu) << 16;
      range <<= 16;
    }
  }
  g_dec.cab_low = low; g_dec.cab_range = range;
}
AVR_DEC_INL void avr_cab_put(int symbol, int ctx) {
  const uint32_t s = g_dec.st[ctx];
  uint32_t p = s >> 1, mps = s & 1;
  const int lps = symbol != (int)mps;
  const uint32_t range = g_dec.cab_range;
  const int normalize = 23 - AVR_CLZ32(range);
  const uint32_t approx = range >> (normalize - 1);
  const uint32_t rlps = (uint32_t)g_dec.hot.range_lps[(p << 2) | ((approx >> 7) & 3)] << normalize;
  avr_cab_put_raw(lps, rlps);
  if (lps) { if (p == 0) mps ^= 1; p = g_dec.hot.trans_lps[p]; } else if (p < 62) p++;
  g_dec.st[ctx] = (uint8_t)((p << 1) | mps);
}
AVR_DEC_INL int avr_cab_put_terminate(int eos) {
  const int normalize = 23 - AVR_CLZ32(g_dec.cab_range);
  avr_cab_put_raw(eos, 2u << normalize);
  if (!eos) return 0;
  uint32_t low = g_dec.cab_low, range = g_dec.cab_range;
  for (uint32_t stop = 1u << 30; stop > 0; stop >>= 1) {     // finish(), arithmetic_code.h:128-144
    uint32_t x = (low | stop) & ~(stop - 1);
    if (stop < range && low <= x && x < low + range) { low = x; break; }
  }
  while (low != 0) {                                          // 8-bit digits in finish()
    if (low >> 31) { avr_cab_carry(); low &= 0x7FFFFFFFu; }
    avr_cab_push(low >> 23);
    low = (low & 0x7FFFFFu) << 8;
  }
  g_dec.cab_low = 0; g_dec.cab_range = 0;
  if (g_dec.cab_n > 0 && g_dec.cab_n <= g_dec.cab_cap && g_dec.cab_out[g_dec.cab_n - 1] == 0x80) g_dec.cab_n--;   // recode.cpp:1508-1515
  return 1;
}

// ---- the four bin entry points (one copy each in the kernel) ----------------------------------
AVR_DEC_FN int avr_dec_get(int ctx) {
  g_dec.bins++;
  int sym;
  if (g_dec.in_sigmap) {
    if (g_dec.expect_eob) {
      sym = (g_dec.stored == g_dec.observed);                 // inferred, recode.cpp:1444-1445, 809-816
      g_dec.expect_eob = 0;
      if (sym) g_dec.ended_by_eob = 1; else g_dec.zz++;
    } else {
      const uint32_t key = avr_key_sig(g_dec.cat, g_dec.stored, g_dec.observed,
                                       avr_zz_off(g_dec.zz, g_dec.size_blk, g_dec.is_dc, g_dec.c422));
      sym = avr_dec_global(key, 0x50);
      g_dec.observed += sym; g_dec.ones += sym;
      if (sym) g_dec.expect_eob = 1; else g_dec.zz++;
    }
  } else {
    sym = avr_dec_default((uint32_t)ctx);
  }
  avr_cab_put(sym, ctx);
  return sym;
}
AVR_DEC_FN int avr_dec_bypass() {
  g_dec.bins++;
  const int sym = avr_dec_default(AVR_CTX_BYPASS);
  avr_cab_put_raw(sym, g_dec.cab_range >> 1);
  return sym;
}
AVR_DEC_FN int avr_dec_terminate() {
  g_dec.bins++;
  const int sym = avr_dec_default(AVR_CTX_TERMINATE);
  if (avr_cab_put_terminate(sym)) g_dec.done = 1;
  return sym;
}
// nnz side channel first (recode.cpp:1483-1499); the count is recovered completely (a block that
// codes as 0 in 2/4 bits is a full one) so the store holds what the COMPRESSOR held.
AVR_DEC_FN void avr_dec_sigmap_begin() {
  AvrNzCtx nc;
  avr_nz_ctx(&nc, g_dec.store[g_dec.cur], g_dec.store[g_dec.cur ^ 1], g_dec.mb_addr, g_dec.mb_w, g_dec.n, g_dec.size_blk,
             g_dec.cat, g_dec.is_dc, g_dec.c422);
  const int nbits = avr_nz_bits(g_dec.size_blk);
  int v = 0;
  for (int k = 0; k < nbits; k++) v |= avr_dec_global(avr_nz_bit_key(nc, k, v), 0x60) << k;
  if (v == 0 && g_dec.size_blk <= 16) v = g_dec.size_blk;
  g_dec.stored = v;
  g_dec.in_sigmap = 1; g_dec.expect_eob = 0; g_dec.zz = 0; g_dec.observed = 0; g_dec.ones = 0; g_dec.ended_by_eob = 0;
}
AVR_DEC_FN void avr_dec_sigmap_end() {
  uint8_t* mb = g_dec.store[g_dec.cur] + (size_t)g_dec.mb_addr * AVR_MB_STRIDE;
  if (g_dec.size_blk > 32) mb[AVR_MB_IS8X8] = 1;
  mb[g_dec.n] = (uint8_t)((g_dec.ones + !g_dec.ended_by_eob) & 63);
  g_dec.in_sigmap = 0;
}

// Bin I/O policy for avr::SliceParser: stateless, forwards to the functions above.
struct AvrDecPolicy {
  AVR_DEC_M int get(int ctx) { return avr_dec_get(ctx); }
  AVR_DEC_M int bypass(int) { return avr_dec_bypass(); }
  AVR_DEC_M int terminate(int) { return avr_dec_terminate(); }
  AVR_DEC_M void mb_begin(int x, int y) { g_dec.mb_addr = (uint32_t)(x + y * g_dec.mb_w); }
  AVR_DEC_M void sub_mb_begin(int cat, int n, int mc, int dc, int c422) { g_dec.cat = cat; g_dec.n = n; g_dec.size_blk = mc; g_dec.is_dc = dc; g_dec.c422 = c422; }
  AVR_DEC_M void sub_mb_end() {}
  AVR_DEC_M void sigmap_begin() { avr_dec_sigmap_begin(); }
  AVR_DEC_M void sigmap_end() { avr_dec_sigmap_end(); }
  AVR_DEC_M void hint(int, int) {}
};

// ---- per-file / per-slice driver ---------------------------------------------------------------
// lane 0 only
AVR_DEC_INL void avr_dec_file_begin(uint16_t* est, uint8_t* s0, uint8_t* s1) {
  g_dec.est = est; g_dec.store[0] = s0; g_dec.store[1] = s1; g_dec.cur = 0; g_dec.mb_w = 1;
}
// update_frame_spec flip (recode.cpp:831-850): lane 0 toggles, then every lane zeroes its share
AVR_DEC_INL void avr_dec_flip_toggle(const AvrDecSlice& ds) { g_dec.cur ^= 1; g_dec.mb_w = ds.sp.mb_w; }
AVR_DEC_INL void avr_dec_flip_zero(const AvrDecSlice& ds, int lane, int nlanes) {
  AvrChain::zero_stores(g_dec.store[g_dec.cur], ds.flip == 1 ? nullptr : g_dec.store[g_dec.cur ^ 1], ds.sp.mb_w, ds.sp.mb_h, lane, nlanes);
}
// One coded slice on ONE thread.  g_dec.st must hold the slice's initial CABAC states.
// Returns AVR_OK or an error; after an error the model state of the file is meaningless.
AVR_DEC_FN int avr_dec_slice(const AvrDecSlice& ds, const uint8_t* payloads, avr::MbInfo* row, uint32_t uid, uint8_t* out,
                             uint32_t* out_len, uint32_t* out_bins) {
  if (uid == 1) for (int x = 0; x < ds.sp.mb_w; x++) row[x].uid = 0;
  avr_dec_init_range(payloads + ds.payload_off, ds.payload_len);
  g_dec.cab_low = 0; g_dec.cab_range = (1u << 22) * 0x1FEu; g_dec.cab_out = out + ds.out_off; g_dec.cab_n = 0;
  g_dec.cab_cap = ds.out_cap; g_dec.cab_err = 0;
  g_dec.in_sigmap = 0; g_dec.expect_eob = 0; g_dec.bins = 0; g_dec.done = 0; g_dec.mb_addr = 0;
  AvrDecPolicy io;
  avr::SliceParser<AvrDecPolicy> parser(io, ds.sp, row, uid);
  const int rc = parser.run();
  *out_len = g_dec.cab_n; *out_bins = g_dec.bins;
  return rc < 0 ? AVR_E_CORRUPT : g_dec.cab_err ? g_dec.cab_err : !g_dec.done ? AVR_E_CORRUPT
         : g_dec.pos > ds.payload_len + 16 ? AVR_E_CORRUPT : AVR_OK;
}

#if !defined(__CUDACC__)
// Whole file on one host thread (tests/host_emul.cpp); the kernel in avr_kernels.cu runs the same
// steps with the bulk work spread over the warp.
static void avr_decompress_file_host(const AvrFileDesc& fd, const AvrDecSlice* slices_all, const uint8_t* payloads,
                                     uint16_t* est, uint8_t* store0, uint8_t* store1, avr::MbInfo* row, const AvrTables* tab,
                                     uint8_t* out, uint32_t* out_len, uint32_t* out_bins, int* status) {
  g_dec.hot = tab->hot;
  for (uint32_t i = 0; i < AVR_NUM_CTX_KEYS; i++) g_dec.est_def[i] = AVR_EST_INIT;
  avr_dec_file_begin(est, store0, store1);
  uint32_t uid = 0;
  int failed = 0;
  for (uint32_t s = 0; s < fd.nslices; s++) {
    const AvrDecSlice& ds = slices_all[fd.slice_off + s];
    const uint32_t g = fd.slice_off + s;
    if (failed) { out_len[g] = 0; out_bins[g] = 0; status[g] = AVR_E_CORRUPT; continue; }
    if (ds.flip) { avr_dec_flip_toggle(ds); avr_dec_flip_zero(ds, 0, 1); }
    avr_cabac_init_states(g_dec.st, tab, ds.sp.slice_type == avr::SLICE_I, ds.sp.cabac_init_idc, ds.sp.slice_qp, 0, 1);
    status[g] = avr_dec_slice(ds, payloads, row, ++uid, out, &out_len[g], &out_bins[g]);
    if (status[g]) failed = 1;
  }
}
#endif
