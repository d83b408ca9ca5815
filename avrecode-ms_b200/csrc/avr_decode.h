// Decompress chain: one per file, strictly sequential per bin (reference recode.cpp:1442-1499:
// range-decode or infer the bin, re-encode it as CABAC, update the model, hand the symbol back
// to the slice_data() walk which picks the next context).
//
// Layout decisions, each taken from an ncu capture (profiles/r01_decompress_chain_*):
//  * v1 inlined the bin I/O at each of the ~150 call sites of the syntax walk: 51 k SASS
//    instructions (`stall_no_inst`), coder state behind generic pointers (20 LD + 10 ST + 32 R2UR
//    per bin).  -> the per-chain state is ONE shared-memory object per block (explicit LDS/STS)
//    and get / bypass / terminate are single non-inlined functions working on it.
//  * v2 spent ~40 of 168 instructions per bin in the CABAC re-encode, a dependent chain the next
//    bin does not need.  -> the chain thread only posts (ctx, symbol) records into a
//    shared-memory ring; a second warp of the block is the CABAC encoder (cabac_code.h:30-67)
//    with its low/range/cursor in registers.
//  * 80 % of all bins are inside residual blocks.  -> SliceParser hands whole residual blocks to
//    AvrDecPolicy::residual_block(), which keeps the range decoder's low/range in registers
//    across the block instead of a shared-memory round trip per bin.
//
// Include only from avr_kernels.cu (device: state in __shared__, one chain per block) and from
// tests/host_emul.cpp (host: same code on a static object, the encoder runs inline).
#pragma once
#include <stddef.h>
#include "avr_bodies.h"
#include "avr_mem.h"

#define AVR_RING_LOG 10
#define AVR_RING (1u << AVR_RING_LOG)
#define AVR_RING_EMPTY 0xFFFFFFFFu
#define AVR_REC_ABORT 2047u          // ctx code: chain failed, encoder must stop
#define AVR_REC_NOP 2046u            // ctx code: padding up to the end of a batch, ignored
#if defined(AVR_ENC_BATCH)
#define AVR_RING_BATCH 32u           // the encoder takes records in aligned batches of this many (one probe per batch)
#endif

// diagnostic builds only (tools/gpu_ab2.sh): AVR_DIAG_NOENC = the encoder warp does nothing and the chain never waits for it
#if defined(AVR_DIAG_NOENC)
#define AVR_DIAG_WAIT(c) false
#else
#define AVR_DIAG_WAIT(c) (c)
#endif

struct AvrCabRegs {                  // CABAC encoder state (registers of the encoder thread)
  uint32_t low, range, n, cap;
  uint8_t* out;
  int err;
};

struct alignas(16) AvrMs { uint64_t m, s; };               // divisor constants of one total, one 16-byte load (avr_div_tab.h)
struct AvrDecShared {
  AvrHotTabs hot;                              // per-bin constant tables
  AvrMs ms[98];                                // hot.div again, one entry per total
  uint16_t est_def[AVR_NUM_CTX_KEYS + 2];      // estimators of the 1026 default keys (file lifetime)
  uint8_t st[1024];                            // CABAC context states of the current slice (encoder thread)
  alignas(16) uint32_t ring[AVR_RING];         // (ctx << 1) | symbol, or AVR_RING_EMPTY; the consumer empties what it has read
  uint32_t ring_head, ring_tail_seen;          // producer side
  volatile uint32_t ring_tail;                 // consumer progress (back-pressure)
#if defined(AVR_ENC_MBAR)
  alignas(8) uint64_t mbar;                    // experiment: the encoder waits on an mbarrier (one arrival per record) instead of polling
  uint32_t mbar_phase;
#endif
  // range decoder (arithmetic_code.h:220-288)
  uint64_t low, range, w_cur, w_next;
  const uint8_t* in;
  uint32_t pos, size, next_digit;
  // CABAC encoder result of the slice, written by the encoder thread at the end
  uint32_t cab_n; int cab_err;
  // model (h264_model, recode.cpp:625-1066)
  uint16_t* est;                               // global: sig-map and nnz-bit estimators
  uint8_t* store[2];
  int cur, mb_w;
  uint32_t mb_addr, mb_x;
  int done;
  uint64_t self;                               // shared-window address of this object, read back as an opaque value (see avr_sbase)
};

#if defined(__CUDACC__)
__shared__ AvrDecShared g_dec;
#define AVR_DEC_FN static __device__ __noinline__
#define AVR_DEC_INL static __device__ __forceinline__
#define AVR_DEC_M __device__ __forceinline__
#else
static AvrDecShared g_dec;
static AvrCabRegs g_cab_host;        // host emulation: the encoder runs inline at post time
#define AVR_DEC_FN static
#define AVR_DEC_INL static inline
#define AVR_DEC_M inline
#endif

// ---- CABAC encoder (runs on the encoder thread; state in registers) ----------------------------
AVR_DEC_INL void avr_cab_push(AvrCabRegs& c, uint32_t b) {
  if (c.n < c.cap) c.out[c.n] = (uint8_t)b; else c.err = AVR_E_OVERFLOW;
  c.n++;
}
AVR_DEC_INL void avr_cab_carry(AvrCabRegs& c) {
  uint32_t i = c.n < c.cap ? c.n : c.cap;
  while (i > 0) { if (++c.out[--i] != 0) break; }
}
AVR_DEC_INL void avr_cab_put_raw(AvrCabRegs& c, int symbol, uint32_t range_of_1) {
  const uint32_t range_of_0 = c.range - range_of_1;
  if (symbol) { c.low += range_of_0; c.range = range_of_1; } else { c.range = range_of_0; }
  if (c.range < 0x200u) {
    if (c.range == 0) { c.err = AVR_E_CORRUPT; c.range = 1; }
    while (c.range < 0x8000u) {                             // 16-bit digits, arithmetic_code.h:113-122
      if (c.low >> 31) { avr_cab_carry(c); c.low &= 0x7FFFFFFFu; }
      const uint32_t digit = c.low >> 15;
      avr_cab_push(c, digit >> 8); avr_cab_push(c, digit & 0xFF);
      c.low = (c.low & 0x7FFFu) << 16;
      c.range <<= 16;
    }
  }
}
AVR_DEC_INL void avr_cab_finish(AvrCabRegs& c) {
  for (uint32_t stop = 1u << 30; stop > 0; stop >>= 1) {     // finish(), arithmetic_code.h:128-144
    uint32_t x = (c.low | stop) & ~(stop - 1);
    if (stop < c.range && c.low <= x && x < c.low + c.range) { c.low = x; break; }
  }
  while (c.low != 0) {                                        // 8-bit digits in finish()
    if (c.low >> 31) { avr_cab_carry(c); c.low &= 0x7FFFFFFFu; }
    avr_cab_push(c, c.low >> 23);
    c.low = (c.low & 0x7FFFFFu) << 8;
  }
  c.range = 0;
  if (c.n > 0 && c.n <= c.cap && c.out[c.n - 1] == 0x80) c.n--;   // recode.cpp:1508-1515
}
// one record; returns 1 when the slice is complete (end_of_slice or abort)
AVR_DEC_INL int avr_cab_consume(AvrCabRegs& c, uint32_t rec) {
  const uint32_t ctx = (rec >> 1) & 0x7FF;
  const int symbol = (int)(rec & 1);
  if (ctx < 1024) {                                           // cabac::encoder::put, cabac_code.h:33-49
    const uint32_t s = g_dec.st[ctx];
    uint32_t p = s >> 1, mps = s & 1;
    const int lps = symbol != (int)mps;
    const int normalize = 23 - AVR_CLZ32(c.range);
    const uint32_t approx = c.range >> (normalize - 1);
    const uint32_t rlps = (uint32_t)g_dec.hot.range_lps[(p << 2) | ((approx >> 7) & 3)] << normalize;
    avr_cab_put_raw(c, lps, rlps);
    if (lps) { if (p == 0) mps ^= 1; p = g_dec.hot.trans_lps[p]; } else if (p < 62) p++;
    g_dec.st[ctx] = (uint8_t)((p << 1) | mps);
    return 0;
  }
  if (ctx == AVR_CTX_BYPASS) { avr_cab_put_raw(c, symbol, c.range >> 1); return 0; }   // put_bypass
  if (ctx == AVR_CTX_TERMINATE) {                             // put_terminate, cabac_code.h:57-67
    const int normalize = 23 - AVR_CLZ32(c.range);
    avr_cab_put_raw(c, symbol, 2u << normalize);
    if (!symbol) return 0;
    avr_cab_finish(c);
    return 1;
  }
  return ctx == AVR_REC_NOP ? 0 : 1;                          // padding / AVR_REC_ABORT
}

// ---- ring -----------------------------------------------------------------------------------------
#if defined(AVR_ENC_MBAR) && defined(__CUDACC__)
// producer: one arrival per posted record completes a phase (expected count 1); release semantics order the record store before it
AVR_DEC_INL void avr_mbar_arrive() {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(&g_dec.mbar);
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared.b64 st, [%0]; }" ::"r"(a) : "memory");
}
// consumer: hardware-suspended wait for the phase of parity `ph` (returns after a time limit at the latest)
AVR_DEC_INL bool avr_mbar_try_wait(uint32_t ph) {
  const uint32_t a = (uint32_t)__cvta_generic_to_shared(&g_dec.mbar);
  uint32_t ok;
  asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared.b64 p, [%1], %2, %3; selp.u32 %0, 1, 0, p; }"
               : "=r"(ok) : "r"(a), "r"(ph), "r"(2000u) : "memory");
  return ok != 0;
}
#define AVR_MBAR_ARRIVE() avr_mbar_arrive()
#else
#define AVR_MBAR_ARRIVE() do {} while (0)
#endif
AVR_DEC_INL void avr_ring_post(uint32_t ctx, int symbol) {
#if defined(__CUDACC__)
  const uint32_t head = g_dec.ring_head;
  if (head - g_dec.ring_tail_seen >= AVR_RING - 1) {          // full: wait for the encoder
    uint32_t t;
    do { t = g_dec.ring_tail; } while (AVR_DIAG_WAIT(head - t >= AVR_RING - 1));
    g_dec.ring_tail_seen = t;
  }
  reinterpret_cast<volatile uint32_t*>(g_dec.ring)[head & (AVR_RING - 1)] = (ctx << 1) | (uint32_t)symbol;
  AVR_MBAR_ARRIVE();
  g_dec.ring_head = head + 1;
#else
  avr_cab_consume(g_cab_host, (ctx << 1) | (uint32_t)symbol);
  g_dec.ring_head++;
#endif
}

// same, with the producer cursor kept in the caller's registers (residual blocks)
AVR_DEC_INL void avr_ring_post_r(uint32_t& head, uint32_t& tail_seen, uint32_t ctx, int symbol) {
#if defined(__CUDACC__)
  if (head - tail_seen >= AVR_RING - 1) {
    uint32_t t;
    do { t = g_dec.ring_tail; } while (AVR_DIAG_WAIT(head - t >= AVR_RING - 1));
    tail_seen = t;
  }
  reinterpret_cast<volatile uint32_t*>(g_dec.ring)[head & (AVR_RING - 1)] = (ctx << 1) | (uint32_t)symbol;
  AVR_MBAR_ARRIVE();
  head++;
#else
  (void)tail_seen; head++;
  avr_cab_consume(g_cab_host, (ctx << 1) | (uint32_t)symbol);
#endif
}

#if !defined(AVR_ENC_SLEEP)
#define AVR_ENC_SLEEP 32
#endif
#if defined(__CUDACC__)
// Encoder thread: consumes the slice's records until end_of_slice / abort.
static __device__ __noinline__ void avr_cab_encoder_thread(uint8_t* out, uint32_t cap) {
  AvrCabRegs c;
  c.low = 0; c.range = (1u << 22) * 0x1FEu; c.n = 0; c.cap = cap; c.out = out; c.err = 0;   // cabac_code.h:30
  uint32_t tail = g_dec.ring_tail;
#if defined(AVR_DIAG_NOENC)
  g_dec.cab_n = 0; g_dec.cab_err = 0; (void)tail; (void)c;
  return;
#endif
#if defined(AVR_ENC_BATCH)
  // Records are taken in aligned batches of AVR_RING_BATCH: slots are filled in order, so ONE probe of the batch's last
  // slot says that the whole batch is there, and the encoder sleeps (instead of spinning on every slot) while the slower
  // chain fills it.  The chain pads a slice's last batch with NOP records.
  volatile uint32_t* ring = reinterpret_cast<volatile uint32_t*>(g_dec.ring);
  bool finished = false;
  while (!finished) {
    const uint32_t probe = (tail + AVR_RING_BATCH - 1) & (AVR_RING - 1);
    while (ring[probe] == AVR_RING_EMPTY) __nanosleep(AVR_ENC_SLEEP);
    for (uint32_t k = 0; k < AVR_RING_BATCH; k += 4) {
      const uint32_t sa = (uint32_t)__cvta_generic_to_shared(&g_dec.ring[(tail + k) & (AVR_RING - 1)]);
      uint4 q;
      asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(q.x), "=r"(q.y), "=r"(q.z), "=r"(q.w) : "r"(sa) : "memory");
      asm volatile("st.volatile.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(sa), "r"(AVR_RING_EMPTY) : "memory");
      if (!finished) finished = avr_cab_consume(c, q.x) != 0;
      if (!finished) finished = avr_cab_consume(c, q.y) != 0;
      if (!finished) finished = avr_cab_consume(c, q.z) != 0;
      if (!finished) finished = avr_cab_consume(c, q.w) != 0;
    }
    tail += AVR_RING_BATCH;
    g_dec.ring_tail = tail;            // after the slots were emptied: the producer reuses a slot only after it has seen the tail pass it
  }
#else
  volatile uint32_t* ring = reinterpret_cast<volatile uint32_t*>(g_dec.ring);
#if defined(AVR_ENC_MBAR)
  uint32_t mph = g_dec.mbar_phase;
#endif
  for (;;) {
    const uint32_t slot = tail & (AVR_RING - 1);
    uint32_t rec = ring[slot];
#if defined(AVR_ENC_MBAR)
    // the slot is the truth; the barrier only parks the warp.  A phase missed while consuming makes one wait return at once or
    // end on its time limit, never lose a record.
    while (rec == AVR_RING_EMPTY) { if (avr_mbar_try_wait(mph)) mph ^= 1u; rec = ring[slot]; }
#else
    while (rec == AVR_RING_EMPTY) { __nanosleep(AVR_ENC_SLEEP); rec = ring[slot]; }
#endif
    ring[slot] = AVR_RING_EMPTY;       // before the tail moves: the producer reuses a slot only after it has seen the tail pass it
    tail++;
    if ((tail & 63) == 0) g_dec.ring_tail = tail;
    if (avr_cab_consume(c, rec)) break;
  }
  g_dec.ring_tail = tail;
#if defined(AVR_ENC_MBAR)
  g_dec.mbar_phase = mph;
#endif
#endif
  g_dec.cab_n = c.n; g_dec.cab_err = c.err;
}
#endif

// ---- range decoder: low/range are passed by reference (registers of the caller) ----------------
AVR_DEC_INL uint64_t avr_dec_load_word(uint32_t off) {
  if (off >= g_dec.size) return 0;
  uint64_t w;
#if defined(__CUDACC__)
  w = __ldg(reinterpret_cast<const unsigned long long*>(g_dec.in + off));
#else
  w = 0;
  for (int k = 7; k >= 0; k--) w = (w << 8) | g_dec.in[off + (uint32_t)k];
#endif
  const uint32_t valid = g_dec.size - off;
  if (valid < 8) w &= (1ull << (8 * valid)) - 1;
  return w;
}
// input bytes come eight at a time; the next word is requested when the current one is opened
// so its latency is off the per-bin dependency chain.  Rare (about once per 8 bins): not inlined.
struct AvrLowRange { uint64_t low, range; };
// by value in, by value out: a reference parameter of a non-inlined function would pin the
// caller's low/range to local memory (seen as LDL/STL per bin in profiles/r01_*_v3)
AVR_DEC_FN AvrLowRange avr_dec_refill(uint64_t low, uint64_t range, uint64_t target) {
  while (range < target) {
    const uint32_t pos = g_dec.pos;
    if ((pos & 7) == 0 && pos) { g_dec.w_cur = g_dec.w_next; g_dec.w_next = avr_dec_load_word(pos + 8); }
    const uint32_t d = (uint32_t)(g_dec.w_cur >> (8 * (pos & 7))) & 0xFF;
    g_dec.pos = pos + 1;
    const uint32_t digit = ((g_dec.next_digit << 7) | (d >> 1)) & 0xFF;   // 1-bit misaligned reads, arithmetic_code.h:259-275
    g_dec.next_digit = d;
    low = (low << 8) + digit;
    range <<= 8;
  }
  AvrLowRange r; r.low = low; r.range = range;
  return r;
}
AVR_DEC_INL void avr_dec_init_range(const uint8_t* p, uint32_t sz) {
  g_dec.in = p; g_dec.size = sz;
  g_dec.w_cur = avr_dec_load_word(0); g_dec.w_next = avr_dec_load_word(8);
  g_dec.next_digit = (uint32_t)g_dec.w_cur & 0xFF; g_dec.pos = 1;
  uint64_t low = g_dec.next_digit >> 1, range = 128;
  const AvrLowRange r = avr_dec_refill(low, range, 1ull << 63);   // 7 digits: range 2^7 -> 2^63 (arithmetic_code.h:226-228)
  g_dec.low = r.low; g_dec.range = r.range;
}
// one symbol with estimator cell `cell` (recode.cpp:823-827 + arithmetic_code.h:232-248)
AVR_DEC_INL int avr_dec_symbol(uint64_t& low, uint64_t& range, uint32_t cell) {
  const uint32_t pos = cell & 0xFF, tot = pos + (cell >> 8);
  const uint64_t r1 = (AVR_MULHI64(range, g_dec.hot.div.m[tot]) >> g_dec.hot.div.s[tot]) * (uint64_t)pos;
  const uint64_t r0 = range - r1;
  const int symbol = low >= r0;
  if (symbol) { low -= r0; range = r1; } else { range = r0; }
  if (range < (1ull << 51) && range != 0) { const AvrLowRange r = avr_dec_refill(low, range, 1ull << 55); low = r.low; range = r.range; }
  return symbol;
}
// estimator update on the packed cell (pos | neg << 8), recode.cpp:1037-1054
AVR_DEC_INL uint32_t avr_cell_update(uint32_t cell, int symbol, uint32_t limit) {
  cell += symbol ? 1u : 0x100u;
  if ((cell & 0xFF) + (cell >> 8) > limit) cell = ((cell + 0x0101u) >> 1) & 0x7F7Fu;
  return cell;
}
AVR_DEC_INL int avr_dec_default(uint64_t& low, uint64_t& range, uint32_t key) {   // key < 1026: shared memory
  const uint32_t cell = g_dec.est_def[key];
  const int symbol = avr_dec_symbol(low, range, cell);
  g_dec.est_def[key] = (uint16_t)avr_cell_update(cell, symbol, 0x60);
  return symbol;
}
AVR_DEC_INL int avr_dec_global(uint64_t& low, uint64_t& range, uint16_t* est, uint32_t key, uint32_t limit) {
  uint16_t* c = est + key;
  const uint32_t cell = *c;
  const int symbol = avr_dec_symbol(low, range, cell);
  *c = (uint16_t)avr_cell_update(cell, symbol, limit);
  return symbol;
}

// address-space explicit accessors: avr_mem.h
AVR_DEC_INL void avr_s_ld_ms(avr_sptr a, uint64_t& m, uint32_t& sh) { uint64_t t; avr_s_ld_2x64_ro(a, m, t); sh = (uint32_t)t; }
AVR_DEC_INL void avr_ms_fill() {     // chain thread, once per file
  for (uint32_t t = 0; t < 98; t++) { g_dec.ms[t].m = g_dec.hot.div.m[t]; g_dec.ms[t].s = g_dec.hot.div.s[t]; }
}

// ---- one residual block (ffmpeg h264_cabac.c:1583-1780 walk; reference order recode.cpp:1483-1499)
// nnz side channel -> significance map (EOB bins inferred: recode.cpp:1444-1445) -> levels/signs.
// Returns coeff_count.
//
// One chain thread issues in order, so what bounds a bin is the length of the DEPENDENT instruction
// sequence plus the fixed cost of every branch (predicate latency ~13 cycles + BRA 5 + BSSY/BSYNC),
// not the instruction count.  The only true recurrence is low/range -> symbol -> low/range; the rest is
// arranged so that it does not wait on it:
//   * cells and divisor constants (m, s) of BOTH possible next keys are requested before the current
//     bin is decoded, and selects pick after;
//   * the bypass estimator and the coeff_abs_level (ctx2) estimator stay in registers across their
//     runs, with the divisor constants of total+1 requested before the symbol is known;
//   * one branch per significance-map position covers everything unusual (renormalisation, end of
//     block, 8x8 offset repeat); ring space is reserved per block / per coefficient, not per bin; the
//     EOB record is stored unconditionally (AVR_RING_EMPTY when there is none).
struct AvrEstReg { uint32_t cell; uint64_t m; uint32_t s; };
AVR_DEC_INL void avr_est_consts(avr_sptr s_ms, AvrEstReg& e) {
  const uint32_t tot = (e.cell & 0xFF) + (e.cell >> 8);
  avr_s_ld_ms(s_ms + tot * 16u, e.m, e.s);
}
AVR_DEC_INL int avr_dec_core(uint64_t& low, uint64_t& range, const AvrEstReg& e) {   // arithmetic_code.h:232-248 without the renormalisation
  const uint64_t r1 = (AVR_MULHI64(range, e.m) >> e.s) * (uint64_t)(e.cell & 0xFF);
  const uint64_t r0 = range - r1;
  const int symbol = low >= r0;
  low -= symbol ? r0 : 0ull;
  range = symbol ? r1 : r0;
  return symbol;
}
#define AVR_DEC_RENORM(low_, range_) do { if ((range_) < (1ull << 51)) { const AvrLowRange r_ = avr_dec_refill((low_), (range_), 1ull << 55); (low_) = r_.low; (range_) = r_.range; } } while (0)
// register-resident estimator: decode one symbol, update, and have the constants of the new total ready
AVR_DEC_INL int avr_dec_reg(uint64_t& low, uint64_t& range, avr_sptr s_ms, AvrEstReg& e, uint32_t limit) {
  const uint32_t tot1 = (e.cell & 0xFF) + (e.cell >> 8) + 1;
  uint64_t m1; uint32_t s1; avr_s_ld_ms(s_ms + tot1 * 16u, m1, s1);          // independent of the symbol
  const int symbol = avr_dec_core(low, range, e);
  e.cell += symbol ? 1u : 0x100u; e.m = m1; e.s = s1;
  if (tot1 > limit) { e.cell = ((e.cell + 0x0101u) >> 1) & 0x7F7Fu; avr_est_consts(s_ms, e); }   // halving: once per ~50 uses of a key
  return symbol;
}
AVR_DEC_INL void avr_ring_reserve(uint32_t head, uint32_t& tail_seen, uint32_t n) {
#if defined(__CUDACC__)
  if (head + n - 1 - tail_seen >= AVR_RING - 1) {
    uint32_t t;
    do { t = g_dec.ring_tail; } while (AVR_DIAG_WAIT(head + n - 1 - t >= AVR_RING - 1));
    tail_seen = t;
  }
#else
  (void)head; (void)tail_seen; (void)n;
#endif
}
// space reserved by the caller; rec = (ctx << 1) | symbol, or AVR_RING_EMPTY (a no-op on a free slot)
AVR_DEC_INL void avr_ring_store(avr_sptr s_ring, uint32_t slot, uint32_t rec) {
#if defined(__CUDACC__)
  avr_s_st32(s_ring + ((slot & (AVR_RING - 1)) << 2), rec);
  AVR_MBAR_ARRIVE();
#else
  (void)s_ring; (void)slot;
  if (rec != AVR_RING_EMPTY) avr_cab_consume(g_cab_host, rec);
#endif
}

// both outcomes of the estimator update, computed while the symbol is still unknown (recode.cpp:1037-1054)
AVR_DEC_INL void avr_cell_update2(uint32_t cell, uint32_t limit, uint32_t& u0, uint32_t& u1) {
  const bool halve = (cell & 0xFF) + (cell >> 8) + 1 > limit;
  u0 = cell + 0x100u; u1 = cell + 1u;
  if (halve) { u0 = ((u0 + 0x0101u) >> 1) & 0x7F7Fu; u1 = ((u1 + 0x0101u) >> 1) & 0x7F7Fu; }
}
// The shared-window address of g_dec, loaded from g_dec itself: an address ptxas cannot rematerialise from
// SR_CgaCtaId inside the loops (it did, one S2UR + ULEA per bin, when the address came from cvta).
AVR_DEC_INL avr_sptr avr_sbase() { return (avr_sptr)g_dec.self; }
#define AVR_S_OFF(member) ((avr_sptr)offsetof(AvrDecShared, member))

// ---- bin entry points outside residual blocks (one copy each in the kernel) --------------------
AVR_DEC_INL int avr_dec_single(uint32_t key) {
  const avr_sptr s_base = avr_sbase();
  const avr_sptr s_ms = s_base + AVR_S_OFF(ms), s_ring = s_base + AVR_S_OFF(ring), a = s_base + AVR_S_OFF(est_def) + 2u * key;
  uint64_t low = g_dec.low, range = g_dec.range;
  uint32_t head = g_dec.ring_head, tail_seen = g_dec.ring_tail_seen;
  AvrEstReg e; e.cell = avr_s_ld16(a); avr_est_consts(s_ms, e);
  uint32_t u0, u1; avr_cell_update2(e.cell, 0x60, u0, u1);
  const int sym = avr_dec_core(low, range, e);
  avr_s_st16(a, sym ? u1 : u0);
#if defined(__CUDACC__)
  if (head - tail_seen >= AVR_RING - 1) { avr_ring_reserve(head, tail_seen, 1); g_dec.ring_tail_seen = tail_seen; }   // rare: ring full
#endif
  avr_ring_store(s_ring, head, (key << 1) | (uint32_t)sym);
  AVR_DEC_RENORM(low, range);
  g_dec.low = low; g_dec.range = range;
  g_dec.ring_head = head + 1;                                    // every bin posts one record: bins of the slice = head delta
  return sym;
}
AVR_DEC_FN int avr_dec_get(int ctx) { return avr_dec_single((uint32_t)ctx); }
AVR_DEC_FN int avr_dec_bypass() { return avr_dec_single(AVR_CTX_BYPASS); }
AVR_DEC_FN int avr_dec_terminate() {
  const int sym = avr_dec_single(AVR_CTX_TERMINATE);
  if (sym) g_dec.done = 1;
  return sym;
}

template <bool IS64>
AVR_DEC_INL int avr_dec_sigmap(uint64_t& low, uint64_t& range, uint32_t& head, uint16_t* est, avr_sptr s_ms, avr_sptr s_ring,
                               uint32_t key0, int stored, int max_coeff, int sig_base, int last_base, int* last_out) {
  const int coefs = max_coeff - 1;
  int count = 0, last = 0;
  int zz = IS64 ? avr_sig8x8_off(0) : 0;
  uint32_t rowk = key0;                                          // key of (count, offset 0)
  uint32_t k = rowk + (uint32_t)zz;                              // the current position's key
  AvrEstReg e; e.cell = avr_g_ld16(est + k); avr_est_consts(s_ms, e);
  for (;;) {
    // both candidates for the next position, requested before this bin is decoded (always inside the pool)
    const int nlast = last + 1;
    const int nzz = IS64 ? avr_sig8x8_off(nlast) : nlast;
    const uint32_t ka = rowk + (uint32_t)nzz, kb = ka + 16u;
    AvrEstReg ea, eb; ea.cell = avr_g_ld16(est + ka); eb.cell = avr_g_ld16(est + kb);
    avr_est_consts(s_ms, ea); avr_est_consts(s_ms, eb);
    // everything that depends on the symbol only through a select, before the symbol
    const int last_ctx = last_base + (!IS64 ? last : last == 0 ? 0 : last < 16 ? 1 : last < 32 ? 2 : last < 40 ? 3 : last < 48 ? 4 : last < 52 ? 5 : last < 56 ? 6 : last < 60 ? 7 : 8);
    uint32_t u0, u1; avr_cell_update2(e.cell, 0x50, u0, u1);
    const bool last_pos = nlast >= coefs;
    const bool eob1 = stored == count + 1;                       // the EOB bin that follows a 1 is inferred, never range-coded
    const uint32_t rec1 = (uint32_t)(sig_base + zz) << 1;
    const uint32_t rec2 = ((uint32_t)last_ctx << 1) | (uint32_t)eob1;
    const bool stop1 = eob1 | last_pos;
    const bool stop0 = last_pos | (IS64 && nzz == zz);           // 8x8 offsets repeat: the next key would be the cell just written
    // this bin
    const int sig = avr_dec_core(low, range, e);
    avr_g_st16(est + k, sig ? u1 : u0);
    avr_ring_store(s_ring, head, rec1 | (uint32_t)sig);
    avr_ring_store(s_ring, head + 1, sig ? rec2 : AVR_RING_EMPTY);
    head += 1u + (uint32_t)sig;
    count += sig;
    last = nlast;
    if ((sig ? stop1 : stop0) | (range < (1ull << 51))) {        // one branch for everything unusual
      AVR_DEC_RENORM(low, range);
      if (sig && eob1) { last = max_coeff; break; }
      if (last_pos) break;
      if (IS64 && !sig && nzz == zz) { e.cell = u0; avr_est_consts(s_ms, e); continue; }
    }
    e = sig ? eb : ea; k = sig ? kb : ka; rowk += sig ? 16u : 0u; zz = nzz;
  }
  *last_out = last;
  return count;
}

AVR_DEC_FN int avr_dec_residual_block(int cat, int n, int max_coeff, int is_dc, int sig_base, int last_base, int abs_base) {
  uint64_t low = g_dec.low, range = g_dec.range;
  uint32_t head = g_dec.ring_head, tail_seen = g_dec.ring_tail_seen;
  uint16_t* const est = g_dec.est;
  const avr_sptr s_base = avr_sbase();
  const avr_sptr s_ms = s_base + AVR_S_OFF(ms), s_ring = s_base + AVR_S_OFF(ring), s_def = s_base + AVR_S_OFF(est_def);
  // nnz bits first.  The count is recovered completely (a block that codes as 0 in 2/4 bits is a
  // full one) so that the store holds what the COMPRESSOR held.
  AvrNzCtx nc;
  avr_nz_ctx_x(&nc, g_dec.store[g_dec.cur], g_dec.store[g_dec.cur ^ 1], g_dec.mb_addr, (int)g_dec.mb_x, g_dec.mb_w, n, max_coeff, cat, is_dc, 0);
  const int nbits = avr_nz_bits(max_coeff);
  int stored = 0;
  {
    // the keys of bit i+1 for both values of bit i are known before bit i is decoded: their cells (sparse in the
    // 2.7 MB pool, L2 more often than L1) and divisor constants are requested ahead
    uint32_t k = avr_nz_bit_key(nc, 0, 0);
    AvrEstReg e; e.cell = avr_g_ld16(est + k); avr_est_consts(s_ms, e);
    for (int i = 0; i < nbits; i++) {
      const int ni = i + 1 < nbits ? i + 1 : i;                  // past the last bit: any valid key
      const uint32_t ka = avr_nz_bit_key(nc, ni, stored), kb = avr_nz_bit_key(nc, ni, stored | (1 << i));
      AvrEstReg ea, eb; ea.cell = avr_g_ld16(est + ka); eb.cell = avr_g_ld16(est + kb);
      avr_est_consts(s_ms, ea); avr_est_consts(s_ms, eb);
      uint32_t u0, u1; avr_cell_update2(e.cell, 0x60, u0, u1);
      const int bit = avr_dec_core(low, range, e);
      avr_g_st16(est + k, bit ? u1 : u0);
      AVR_DEC_RENORM(low, range);
      stored |= bit << i;
      e = bit ? eb : ea; k = bit ? kb : ka;
    }
  }
  if (stored == 0 && max_coeff <= 16) stored = max_coeff;
  // significance map
  const uint32_t key0 = avr_key_sig(cat, stored, 0, 0);
  int coeff_count = 0, last = 0;
  if (max_coeff > 1) {
    avr_ring_reserve(head, tail_seen, 2u * (uint32_t)max_coeff);
    coeff_count = max_coeff == 64 ? avr_dec_sigmap<true>(low, range, head, est, s_ms, s_ring, key0, stored, max_coeff, sig_base, last_base, &last)
                                  : avr_dec_sigmap<false>(low, range, head, est, s_ms, s_ring, key0, stored, max_coeff, sig_base, last_base, &last);
  }
  if (last == max_coeff - 1) coeff_count++;
  {
    uint8_t* mb = g_dec.store[g_dec.cur] + (size_t)g_dec.mb_addr * AVR_MB_STRIDE;
    if (max_coeff > 32) mb[AVR_MB_IS8X8] = 1;
    mb[n] = (uint8_t)(coeff_count & 63);
  }
  // coeff_abs_level_minus1 + sign, reverse scan order (h264_cabac.c:1727-1768)
  const avr_sptr a_byp = s_def + 2u * AVR_CTX_BYPASS, s_abs = s_def + 2u * (uint32_t)abs_base;
  AvrEstReg byp; byp.cell = avr_s_ld16(a_byp); avr_est_consts(s_ms, byp);
  int node_ctx = 0;
  const int gt1_max = cat == 3 ? 8 : 9;
  avr_sptr a1 = s_abs + 2u;                                      // ctx1 of node_ctx 0
  AvrEstReg e1; e1.cell = avr_s_ld16(a1); avr_est_consts(s_ms, e1);
  const uint32_t rec_abs = (uint32_t)abs_base << 1;
  for (int c = coeff_count; c > 0; c--) {
    avr_ring_reserve(head, tail_seen, 80);                       // one coefficient posts at most 1 + 13 + 61 + 1 records
    const int i1 = node_ctx < 4 ? node_ctx + 1 : 0;              // ctx1 - abs_base in 0..4, ctx2 - abs_base in 5..9: never the same cell
    const int i2 = node_ctx < 4 ? 5 : (node_ctx + 2 > gt1_max ? gt1_max : node_ctx + 2);
    const avr_sptr a2 = s_abs + 2u * (uint32_t)i2;
    AvrEstReg e2; e2.cell = avr_s_ld16(a2); avr_est_consts(s_ms, e2);   // used only if b0 = 1; its latency hides behind b0
    // node_ctx after this coefficient and the address of the next b0 estimator, for both outcomes (selects, no branches)
    const int nn0 = node_ctx < 3 ? node_ctx + 1 : node_ctx;
    const int nn1 = node_ctx < 4 ? 4 : (node_ctx < 7 ? node_ctx + 1 : 7);
    const avr_sptr a10 = s_abs + 2u * (uint32_t)(nn0 < 4 ? nn0 + 1 : 0), a11 = s_abs + 2u * (uint32_t)(nn1 < 4 ? nn1 + 1 : 0);
    uint32_t u0, u1; avr_cell_update2(e1.cell, 0x60, u0, u1);
    const int b0 = avr_dec_core(low, range, e1);
    avr_s_st16(a1, b0 ? u1 : u0);
    avr_ring_store(s_ring, head++, (rec_abs + 2u * (uint32_t)i1) | (uint32_t)b0);
    // node_ctx and ctx1 of the next coefficient are known now: request that cell behind the store above
    node_ctx = b0 ? nn1 : nn0;
    a1 = b0 ? a11 : a10;
    e1.cell = avr_s_ld16(a1); avr_est_consts(s_ms, e1);
    AVR_DEC_RENORM(low, range);
    if (b0) {
      const uint32_t rec2 = rec_abs + 2u * (uint32_t)i2;
      int coeff_abs = 2;
      while (coeff_abs < 15) {
        const int b = avr_dec_reg(low, range, s_ms, e2, 0x60);
        avr_ring_store(s_ring, head++, rec2 | (uint32_t)b);
        AVR_DEC_RENORM(low, range);
        if (!b) break;
        coeff_abs++;
      }
      avr_s_st16(a2, e2.cell);
      if (coeff_abs >= 15) {
        int j = 0;
        for (;;) {
          const int b = avr_dec_reg(low, range, s_ms, byp, 0x60);
          avr_ring_store(s_ring, head++, (AVR_CTX_BYPASS << 1) | (uint32_t)b);
          AVR_DEC_RENORM(low, range);
          if (!b || j >= 30) break;
          j++;
        }
        while (j--) {
          const int b = avr_dec_reg(low, range, s_ms, byp, 0x60);
          avr_ring_store(s_ring, head++, (AVR_CTX_BYPASS << 1) | (uint32_t)b);
          AVR_DEC_RENORM(low, range);
        }
      }
    }
    const int sgn = avr_dec_reg(low, range, s_ms, byp, 0x60);
    avr_ring_store(s_ring, head++, (AVR_CTX_BYPASS << 1) | (uint32_t)sgn);
    AVR_DEC_RENORM(low, range);
  }
  avr_s_st16(a_byp, byp.cell);
  g_dec.low = low; g_dec.range = range;
  g_dec.ring_head = head; g_dec.ring_tail_seen = tail_seen;
  return coeff_count;
}

// Bin I/O policy for avr::SliceParser: stateless, forwards to the functions above.
struct AvrDecPolicy {
  static constexpr bool kFastResidual = true;
  AVR_DEC_M int residual_block(int cat, int n, int mc, int dc, int sig, int last, int abs_) { return avr_dec_residual_block(cat, n, mc, dc, sig, last, abs_); }
  AVR_DEC_M int get(int ctx) { return avr_dec_get(ctx); }
  AVR_DEC_M int bypass(int) { return avr_dec_bypass(); }
  AVR_DEC_M int terminate(int) { return avr_dec_terminate(); }
  AVR_DEC_M void mb_begin(int x, int y) { g_dec.mb_addr = (uint32_t)(x + y * g_dec.mb_w); g_dec.mb_x = (uint32_t)x; }
  AVR_DEC_M void sub_mb_begin(int, int, int, int, int) {}
  AVR_DEC_M void sub_mb_end() {}
  AVR_DEC_M void sigmap_begin() {}
  AVR_DEC_M void sigmap_end() {}
  AVR_DEC_M void hint(int, int) {}
};

// ---- per-file / per-slice driver ---------------------------------------------------------------
AVR_DEC_INL void avr_dec_file_begin(uint16_t* est, uint8_t* s0, uint8_t* s1) {   // chain thread only
  g_dec.self = (uint64_t)avr_sp(&g_dec);
  g_dec.est = est; g_dec.store[0] = s0; g_dec.store[1] = s1; g_dec.cur = 0; g_dec.mb_w = 1;
  g_dec.ring_head = 0; g_dec.ring_tail_seen = 0; g_dec.ring_tail = 0;
}
// update_frame_spec flip (recode.cpp:831-850): the chain thread toggles, then every thread zeroes its share
AVR_DEC_INL void avr_dec_flip_toggle(const AvrDecSlice& ds) { g_dec.cur ^= 1; g_dec.mb_w = ds.sp.mb_w; }
AVR_DEC_INL void avr_dec_flip_zero(const AvrDecSlice& ds, int lane, int nlanes) {
  AvrChain::zero_stores(g_dec.store[g_dec.cur], ds.flip == 1 ? nullptr : g_dec.store[g_dec.cur ^ 1], ds.sp.mb_w, ds.sp.mb_h, lane, nlanes);
}
// One coded slice on the chain thread.  g_dec.st must hold the slice's initial CABAC states and the
// encoder thread must be consuming the ring.  Returns AVR_OK or an error (after an error the model
// state of the file is meaningless).  The CABAC byte count / overflow flag come from the encoder.
// HAS_B: a kernel that holds BOTH parser instances runs I/P slices 26 % slower (one chain) than one that holds only the I/P
// instance (same-box A/B, profiles/r02_summary.md), so the instance is chosen per KERNEL and the host picks the kernel per batch.
template <bool HAS_B>
AVR_DEC_FN int avr_dec_slice(const AvrDecSlice& ds, const uint8_t* payloads, avr::MbInfo* row, uint32_t uid, uint32_t* out_bins) {
  if (uid == 1) for (int x = 0; x < ds.sp.mb_w; x++) row[x].uid = 0;
  avr_dec_init_range(payloads + ds.payload_off, ds.payload_len);
  g_dec.done = 0; g_dec.mb_addr = 0;
  const uint32_t head0 = g_dec.ring_head;                      // every bin posts exactly one record
  AvrDecPolicy io;
  int rc;
  { avr::SliceParser<AvrDecPolicy, HAS_B> parser(io, ds.sp, row, uid); rc = parser.run(); }
  *out_bins = g_dec.ring_head - head0;
  if (!g_dec.done) avr_ring_post(AVR_REC_ABORT, 0);            // release the encoder thread
#if defined(AVR_ENC_BATCH)
  while (g_dec.ring_head & (AVR_RING_BATCH - 1)) avr_ring_post(AVR_REC_NOP, 0);   // the encoder takes whole batches
#endif
  return rc < 0 ? AVR_E_CORRUPT : !g_dec.done ? AVR_E_CORRUPT : g_dec.pos > ds.payload_len + 16 ? AVR_E_CORRUPT : AVR_OK;
}

#if !defined(__CUDACC__)
// Whole file on one host thread (tests/host_emul.cpp); the kernel in avr_kernels.cu runs the same
// steps with the bulk work spread over the block and the CABAC encoder on its own warp.
static void avr_decompress_file_host(const AvrFileDesc& fd, const AvrDecSlice* slices_all, const uint8_t* payloads,
                                     uint16_t* est, uint8_t* store0, uint8_t* store1, avr::MbInfo* row, const AvrTables* tab,
                                     uint8_t* out, uint32_t* out_len, uint32_t* out_bins, int* status) {
  g_dec.hot = tab->hot; avr_ms_fill();
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
    g_cab_host.low = 0; g_cab_host.range = (1u << 22) * 0x1FEu; g_cab_host.n = 0; g_cab_host.cap = ds.out_cap;
    g_cab_host.out = out + ds.out_off; g_cab_host.err = 0;
    int rc = avr_dec_slice<true>(ds, payloads, row, ++uid, &out_bins[g]);
    out_len[g] = g_cab_host.n;
    status[g] = rc ? rc : g_cab_host.err;
    if (status[g]) failed = 1;
  }
}
#endif
