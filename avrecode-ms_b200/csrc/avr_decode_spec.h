// Decompress chain: one per file, strictly sequential per bin (reference recode.cpp:1442-1499:
// range-decode or infer the bin, re-encode it as CABAC, update the model, hand the symbol back
// to the slice_data() walk which picks the next context).
//
// Round 2 design: ONE WARP per chain, lanes as speculative paths.
//   The only true recurrence of the range decoder is range -> r1 = (range / total) * pos -> range'
//   (arithmetic_code.h:232-248): `low` merely says, afterwards, which branch was the real one.  Inside a residual
//   block the key of the next bin is a function of the symbols decoded so far (nnz-bit tree, significance-map
//   position/count, coeff_abs_level state machine), so lane p of the warp ASSUMES the next D <= 5 symbols are the
//   bits of p, fetches the D estimator cells of that path at once, and runs the D-step range recurrence with no
//   load, no select on a decoded symbol and no branch in it; each step only checks that `low` agrees with the
//   assumption.  One ballot finds the lane whose assumptions all held (there is always exactly one path, shared by
//   the lanes that agree on its prefix); its low/range/step count are broadcast, it alone writes the estimator
//   cells and posts the CABAC records.  A path ends early where the block phase ends or the range needs a new
//   digit (renormalisation is done once, warp-uniformly, by the commit).
//   Round 1 (one thread per chain, both next candidates prefetched) measured 372 cycles per bin: ~100 dependent
//   instructions at 3.7 cycles each.  Here the dependent sequence per bin is the ~55-cycle recurrence plus the
//   round's set-up and commit shared by up to 5 decoded bins (+ their inferred EOB bins).
//   Everything outside residual blocks (mb_type, mvd, cbp, ...) runs warp-uniformly: all lanes execute the same
//   slice_data() walk on the same values.
//   The CABAC re-encode (cabac_code.h:30-67) is a second warp of the block, fed through a shared-memory ring; its
//   per-context entry holds the state AND the four rangeTabLPS values of that state, the next entry comes from one
//   transition-table load, and the entry of the next record is requested while the current one is coded.
//   The nnz stores of the macroblock, its left/above neighbours and the previous picture are staged in shared
//   memory once per macroblock by 32 lanes (64 B each) instead of one global load per block and neighbour.
//
// Include only from avr_kernels.cu (device: state in __shared__, one chain per block) and from
// tests/host_emul.cpp (host: same code, lanes as a loop, the encoder runs inline).
#pragma once
#include <stddef.h>
#include "avr_bodies.h"
#include "avr_mem.h"

#define AVR_RING_LOG 10
#define AVR_RING (1u << AVR_RING_LOG)
#define AVR_REC_ABORT 2047u          // ctx code: chain failed, encoder must stop
#define AVR_SPEC_D 5                 // speculation depth: 2^5 paths = one warp

struct AvrCabRegs {                  // CABAC encoder state (registers of the encoder thread)
  uint32_t low, range, n, cap;
  uint8_t* out;
  int err;
};

struct alignas(16) AvrMs { uint64_t m, s; };               // divisor constants of one total, one 16-byte load (avr_div_tab.h)
struct AvrDecShared {
  AvrMs ms[98];                                // exact range/total constants, one entry per total
  uint64_t cab_trans[256];                     // [(state << 1) | bin] -> next entry
  uint64_t cab_ent[1024];                      // per context: bits 0..31 rangeTabLPS[pStateIdx][0..3], bits 32..38 state
  uint16_t est_def[AVR_NUM_CTX_KEYS + 2];      // estimators of the 1026 default keys (file lifetime)
  uint16_t ring[AVR_RING];                     // (ctx << 1) | symbol
  uint8_t nb[4][64];                           // nnz rows staged per macroblock: 0 current, 1 left, 2 above, 3 previous picture
  volatile uint32_t ring_pub;                  // records published by the chain warp
  volatile uint32_t ring_tail;                 // consumer progress (back-pressure)
  uint32_t ring_head, ring_tail_seen;          // producer side
  // range decoder (arithmetic_code.h:220-288): digits are the payload read one bit late (:259-275); `win` holds the
  // next win_n digits, most significant byte first; w_next is the following 8 payload bytes, requested ahead
  uint64_t low, range, win, w_next;
  const uint8_t* in;
  uint32_t pos, size, win_n, carry;
  // CABAC encoder result of the slice, written by the encoder thread at the end
  uint32_t cab_n; int cab_err;
  // model (h264_model, recode.cpp:625-1066)
  uint16_t* est;                               // global: sig-map and nnz-bit estimators
  uint8_t* store[2];
  int cur, mb_w;
  uint32_t mb_addr, mb_x;
  int nb_staged;
  int done, err;
  uint64_t self;                               // shared-window address of this object, read back as an opaque value (see avr_sbase)
};

#if defined(__CUDACC__)
__shared__ AvrDecShared g_dec;
#define AVR_DEC_FN static __device__ __noinline__
#define AVR_DEC_INL static __device__ __forceinline__
#define AVR_DEC_M __device__ __forceinline__
#define AVR_LANE_LOOP(l) for (int l = (int)(threadIdx.x & 31u), l##_it = 0; l##_it < 1; l##_it++)
#define AVR_LANE0 ((threadIdx.x & 31u) == 0)
#define AVR_WARP_SYNC() __syncwarp()
template <class T> struct AvrPerLane { T v; __device__ __forceinline__ T& operator[](int) { return v; } };
static __device__ __forceinline__ uint32_t avr_lane_ballot(AvrPerLane<uint32_t>& p) { return __ballot_sync(0xFFFFFFFFu, p.v != 0); }
static __device__ __forceinline__ uint32_t avr_lane_get(AvrPerLane<uint32_t>& x, int w) { return __shfl_sync(0xFFFFFFFFu, x.v, w); }
static __device__ __forceinline__ uint64_t avr_lane_get(AvrPerLane<uint64_t>& x, int w) {
  const uint32_t lo = __shfl_sync(0xFFFFFFFFu, (uint32_t)x.v, w), hi = __shfl_sync(0xFFFFFFFFu, (uint32_t)(x.v >> 32), w);
  return ((uint64_t)hi << 32) | lo;
}
#define AVR_POPC(x) __popc(x)
#define AVR_FFS(x) __ffs((int)(x))
#else
static AvrDecShared g_dec;
static AvrCabRegs g_cab_host;        // host emulation: the encoder runs inline at post time
#define AVR_DEC_FN static
#define AVR_DEC_INL static inline
#define AVR_DEC_M inline
#define AVR_LANE_LOOP(l) for (int l = 0; l < 32; l++)
#define AVR_LANE0 true
#define AVR_WARP_SYNC() do {} while (0)
template <class T> struct AvrPerLane { T v[32]; T& operator[](int l) { return v[l]; } };
static inline uint32_t avr_lane_ballot(AvrPerLane<uint32_t>& p) { uint32_t m = 0; for (int l = 0; l < 32; l++) m |= (uint32_t)(p.v[l] != 0) << l; return m; }
static inline uint32_t avr_lane_get(AvrPerLane<uint32_t>& x, int w) { return x.v[w]; }
static inline uint64_t avr_lane_get(AvrPerLane<uint64_t>& x, int w) { return x.v[w]; }
#define AVR_POPC(x) __builtin_popcount(x)
#define AVR_FFS(x) __builtin_ffs((int)(x))
#endif

// ---- CABAC encoder (runs on the encoder thread; state in registers) ----------------------------
// Entry of a context: what cabac::encoder::put needs from the state (cabac_code.h:37-47) in one 8-byte load.
AVR_DEC_INL uint64_t avr_cab_entry(uint32_t state, const uint8_t* range_lps) {
  const uint32_t p = state >> 1;
  const uint32_t row = (uint32_t)range_lps[p * 4] | ((uint32_t)range_lps[p * 4 + 1] << 8) | ((uint32_t)range_lps[p * 4 + 2] << 16) |
                       ((uint32_t)range_lps[p * 4 + 3] << 24);
  return (uint64_t)row | ((uint64_t)state << 32);
}
AVR_DEC_INL uint32_t avr_cab_next_state(uint32_t state, uint32_t bin, const uint8_t* trans_lps) {
  uint32_t p = state >> 1, mps = state & 1;
  if (bin != mps) { if (p == 0) mps ^= 1; p = trans_lps[p]; } else if (p < 62) p++;
  return (p << 1) | mps;
}
// lane `lane` of `nlanes` fills its share of the transition table (once per file)
AVR_DEC_INL void avr_cab_fill_trans(const AvrTables* tab, int lane, int nlanes) {
  for (int i = lane; i < 256; i += nlanes)
    g_dec.cab_trans[i] = avr_cab_entry(avr_cab_next_state((uint32_t)i >> 1, (uint32_t)i & 1u, tab->hot.trans_lps), tab->hot.range_lps);
}
// per-slice context entries (ITU-T H.264 9.3.1.1; reference ffmpeg h264_cabac.c:1267-1286)
AVR_DEC_INL void avr_cab_init_entries(const AvrTables* tab, int is_intra, int init_idc, int qp, int lane, int nlanes) {
  const int pl = is_intra ? 0 : 1 + init_idc;
  qp = qp < 0 ? 0 : qp > 51 ? 51 : qp;
  for (int i = lane; i < 1024; i += nlanes) {
    int pre = ((tab->ctx_m[pl][i] * qp) >> 4) + tab->ctx_n[pl][i];
    pre = pre < 1 ? 1 : pre > 126 ? 126 : pre;
    const uint32_t st = pre <= 63 ? (uint32_t)((63 - pre) << 1) : (uint32_t)(((pre - 64) << 1) | 1);
    g_dec.cab_ent[i] = avr_cab_entry(st, tab->hot.range_lps);
  }
}
AVR_DEC_INL void avr_cab_push(AvrCabRegs& c, uint32_t b) {
  if (c.n < c.cap) c.out[c.n] = (uint8_t)b; else c.err = AVR_E_OVERFLOW;
  c.n++;
}
AVR_DEC_INL void avr_cab_carry(AvrCabRegs& c) {
  uint32_t i = c.n < c.cap ? c.n : c.cap;
  while (i > 0) { if (++c.out[--i] != 0) break; }
}
AVR_DEC_INL void avr_cab_renorm(AvrCabRegs& c) {
  if (c.range == 0) { c.err = AVR_E_CORRUPT; c.range = 1; }
  while (c.range < 0x8000u) {                             // 16-bit digits, arithmetic_code.h:113-122
    if (c.low >> 31) { avr_cab_carry(c); c.low &= 0x7FFFFFFFu; }
    const uint32_t digit = c.low >> 15;
    avr_cab_push(c, digit >> 8); avr_cab_push(c, digit & 0xFF);
    c.low = (c.low & 0x7FFFu) << 16;
    c.range <<= 16;
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
// One regular or bypass bin with the context entry already in hand (cabac_code.h:33-54).  Returns the entry the
// context holds afterwards (unchanged for bypass).  `ctx` < 1024 or AVR_CTX_BYPASS.
AVR_DEC_INL uint64_t avr_cab_put(AvrCabRegs& c, uint32_t ctx, uint32_t bin, uint64_t ent) {
  const bool regular = ctx < 1024u;
  const uint32_t state = (uint32_t)(ent >> 32), row = (uint32_t)ent;
  const int normalize = 23 - AVR_CLZ32(c.range);             // floor(log2(range / 0x100)), range >= 0x200
  const uint32_t q = (c.range >> (normalize + 6)) & 3u;      // == ((range >> (normalize - 1)) >> 7) & 3
  const uint32_t rlps_tab = ((row >> (q * 8u)) & 0xFFu) << normalize;
  const uint32_t r1 = regular ? rlps_tab : c.range >> 1;     // put_raw's range_of_1
  const uint32_t sym = regular ? (uint32_t)(bin != (state & 1u)) : bin;
  const uint32_t r0 = c.range - r1;
  c.low += sym ? r0 : 0u;
  c.range = sym ? r1 : r0;
  const uint64_t next = regular ? g_dec.cab_trans[((state << 1) | bin) & 255u] : ent;
  if (regular) g_dec.cab_ent[ctx] = next;
  if (c.range < 0x200u) avr_cab_renorm(c);
  return next;
}
// terminate bin (cabac_code.h:57-67); returns 1 when the slice is complete
AVR_DEC_INL int avr_cab_terminate(AvrCabRegs& c, uint32_t bin) {
  const int normalize = 23 - AVR_CLZ32(c.range);
  const uint32_t r1 = 2u << normalize, r0 = c.range - r1;
  if (bin) { c.low += r0; c.range = r1; } else { c.range = r0; }
  if (c.range < 0x200u) avr_cab_renorm(c);
  if (!bin) return 0;
  avr_cab_finish(c);
  return 1;
}
// one record, entry loaded here (host emulation and the unpipelined start of a batch)
AVR_DEC_INL int avr_cab_consume(AvrCabRegs& c, uint32_t rec) {
  const uint32_t ctx = (rec >> 1) & 0x7FF, bin = rec & 1u;
  if (ctx <= AVR_CTX_BYPASS) { avr_cab_put(c, ctx, bin, g_dec.cab_ent[ctx & 1023u]); return 0; }
  if (ctx == AVR_CTX_TERMINATE) return avr_cab_terminate(c, bin);
  return 1;                                                   // AVR_REC_ABORT
}

#if defined(__CUDACC__)
// Encoder thread: consumes the slice's records until end_of_slice / abort.  Records published so far are coded as one
// run: the entry of record i + 1 is requested before record i is coded (and replaced by record i's result when both
// use the same context), the record after that is read two ahead.
static __device__ __noinline__ void avr_cab_encoder_thread(uint8_t* out, uint32_t cap) {
  AvrCabRegs c;
  c.low = 0; c.range = (1u << 22) * 0x1FEu; c.n = 0; c.cap = cap; c.out = out; c.err = 0;   // cabac_code.h:30
  uint32_t tail = g_dec.ring_tail;
  volatile uint16_t* ring = reinterpret_cast<volatile uint16_t*>(g_dec.ring);
  const uint64_t* ent = g_dec.cab_ent;
  for (;;) {
    uint32_t pub = g_dec.ring_pub;
    if (pub == tail) {
      g_dec.ring_tail = tail;
      do { __nanosleep(40); pub = g_dec.ring_pub; } while (pub == tail);
    }
#if defined(AVR_RING_FENCE)
    __threadfence_block();
#endif
    uint32_t rec0 = ring[tail & (AVR_RING - 1)];
    uint32_t rec1 = ring[(tail + 1) & (AVR_RING - 1)];       // may lie beyond `pub`: read, never consumed
    uint64_t e0 = ent[(rec0 >> 1) & 1023u];
    bool finished = false;
    while (tail != pub) {
      const uint32_t ctx0 = (rec0 >> 1) & 0x7FF, bin0 = rec0 & 1u;
      const uint32_t rec2 = ring[(tail + 2) & (AVR_RING - 1)];
      const uint32_t ctx1 = (rec1 >> 1) & 0x7FF;
      uint64_t e1 = ent[ctx1 & 1023u];                        // stale when ctx1 == ctx0: replaced below
      tail++;
      if (ctx0 <= AVR_CTX_BYPASS) {
        const uint64_t ne = avr_cab_put(c, ctx0, bin0, e0);
        if (ctx1 == ctx0) e1 = ne;
      } else if (ctx0 == AVR_CTX_TERMINATE) {
        if (avr_cab_terminate(c, bin0)) { finished = true; break; }
      } else { finished = true; break; }                      // AVR_REC_ABORT
      rec0 = rec1; rec1 = rec2; e0 = e1;
      if ((tail & 63u) == 0) g_dec.ring_tail = tail;
    }
    if (finished) break;
  }
  g_dec.ring_tail = tail;
  g_dec.cab_n = c.n; g_dec.cab_err = c.err;
}
#endif

// ---- ring, producer side ---------------------------------------------------------------------------
// The shared-window address of g_dec, loaded from g_dec itself: an address ptxas cannot rematerialise from
// SR_CgaCtaId inside the loops (it did, one S2UR + ULEA per bin, when the address came from cvta).
AVR_DEC_INL avr_sptr avr_sbase() { return (avr_sptr)g_dec.self; }
#define AVR_S_OFF(member) ((avr_sptr)offsetof(AvrDecShared, member))

// makes records [.., head) visible to the encoder.  Called warp-uniformly after a __syncwarp-ordered commit.
// Ordering: the records and the cursor are volatile shared-memory accesses of one SM, performed in program order
// (__syncwarp orders the winner lane's record stores before lane 0's cursor store); round 1's ring relied on the same.
// -DAVR_RING_FENCE adds fence.cta on both sides (measured: see profiles/).
AVR_DEC_INL void avr_ring_publish(uint32_t head) {
#if defined(__CUDACC__)
#if defined(AVR_RING_FENCE)
  __threadfence_block();
#endif
  if (AVR_LANE0) g_dec.ring_pub = head;
#else
  (void)head;
#endif
}
// room for n more records from `head`; waits for the encoder when the ring is full (everything posted so far is
// published first, or the encoder could never catch up)
AVR_DEC_INL void avr_ring_reserve(uint32_t head, uint32_t& tail_seen, uint32_t n) {
#if defined(__CUDACC__)
  if (head + n - tail_seen > AVR_RING - 2) {
    avr_ring_publish(head);
    uint32_t t;
    do { t = g_dec.ring_tail; } while (head + n - t > AVR_RING - 2);
    tail_seen = t;
  }
#else
  (void)head; (void)tail_seen; (void)n;
#endif
}
// space reserved by the caller; rec = (ctx << 1) | symbol
AVR_DEC_INL void avr_ring_store(avr_sptr s_ring, uint32_t slot, uint32_t rec) {
#if defined(__CUDACC__)
  avr_s_st16(s_ring + ((slot & (AVR_RING - 1)) << 1), rec);
#else
  (void)s_ring; (void)slot;
  avr_cab_consume(g_cab_host, rec);
#endif
}

// ---- range decoder input (arithmetic_code.h:251-288) ---------------------------------------------
// 8 payload bytes at `off` (multiple of 8) as a big-endian number; bytes at or beyond `size` read as zero (:283)
AVR_DEC_INL uint64_t avr_dec_load_be(const uint8_t* in, uint32_t off, uint32_t size) {
  if (off >= size) return 0;
  uint64_t be;
#if defined(__CUDACC__)
  const uint64_t w = __ldg(reinterpret_cast<const unsigned long long*>(in + off));
  be = ((uint64_t)__byte_perm((uint32_t)w, 0, 0x0123) << 32) | __byte_perm((uint32_t)(w >> 32), 0, 0x0123);
#else
  be = 0;
  for (int k = 0; k < 8; k++) be = (be << 8) | in[off + (uint32_t)k];
#endif
  const uint32_t valid = size - off;
  if (valid < 8) be &= ~0ull << (8 * (8 - valid));
  return be;
}
// Decoder registers of the chain warp while it is inside one entry point; loaded from / stored to g_dec at its ends.
struct AvrDecRegs {
  uint64_t low, range, win, w_next;
  uint32_t pos, win_n, carry, head, tail_seen, err;
};
AVR_DEC_INL void avr_regs_load(AvrDecRegs& r) {
  r.low = g_dec.low; r.range = g_dec.range; r.win = g_dec.win; r.w_next = g_dec.w_next;
  r.pos = g_dec.pos; r.win_n = g_dec.win_n; r.carry = g_dec.carry; r.head = g_dec.ring_head; r.tail_seen = g_dec.ring_tail_seen;
  r.err = (uint32_t)g_dec.err;
}
AVR_DEC_INL void avr_regs_store(const AvrDecRegs& r) {
  if (AVR_LANE0) {
    g_dec.low = r.low; g_dec.range = r.range; g_dec.win = r.win; g_dec.w_next = r.w_next;
    g_dec.pos = r.pos; g_dec.win_n = r.win_n; g_dec.carry = r.carry; g_dec.ring_head = r.head; g_dec.ring_tail_seen = r.tail_seen;
    g_dec.err = (int)r.err;
  }
  AVR_WARP_SYNC();
}
// the next 8 digits: the 8 bytes requested ahead, shifted one bit late; requests the 8 bytes after them
AVR_DEC_INL void avr_dec_refill(AvrDecRegs& r) {     // inlined: a reference parameter of a real call would pin r to local memory
  const uint64_t be = r.w_next;
  r.win = ((uint64_t)r.carry << 63) | (be >> 1);
  r.carry = (uint32_t)be & 1u;
  r.win_n = 8;
  r.pos += 8;
  r.w_next = avr_dec_load_be(g_dec.in, r.pos, g_dec.size);
}
// renormalisation (arithmetic_code.h:239-246): digits until range >= 2^55.  Warp-uniform.
AVR_DEC_INL void avr_dec_renorm(AvrDecRegs& r, uint64_t target) {
  if (r.range == 0) { r.err = 1; r.range = 1ull << 63; return; }   // corrupt payload: the slice ends with a status (avr_dec_slice)
  while (r.range < target) {
    if (r.win_n == 0) avr_dec_refill(r);
    r.low = (r.low << 8) + (r.win >> 56);
    r.win <<= 8; r.win_n--;
    r.range <<= 8;
  }
}
AVR_DEC_INL void avr_dec_init_range(const uint8_t* p, uint32_t sz) {   // decoder constructor, arithmetic_code.h:220-230
  AvrDecRegs r;
  g_dec.in = p; g_dec.size = sz;
  const uint64_t be = avr_dec_load_be(p, 0, sz);
  const uint32_t b0 = (uint32_t)(be >> 56);
  r.win = ((uint64_t)(b0 & 1u) << 63) | ((be << 8) >> 1);     // digits 1..7: bytes 1..7 one bit late
  r.win_n = 7; r.carry = (uint32_t)be & 1u; r.pos = 8;
  r.w_next = avr_dec_load_be(p, 8, sz);
  r.low = b0 >> 1; r.range = 128;
  r.head = g_dec.ring_head; r.tail_seen = g_dec.ring_tail_seen; r.err = 0;
  avr_dec_renorm(r, 1ull << 63);                              // 7 digits: range 2^7 -> 2^63 (arithmetic_code.h:226-228)
  avr_regs_store(r);
}

// estimator update on the packed cell (pos | neg << 8), recode.cpp:1037-1054
AVR_DEC_INL uint32_t avr_cell_update(uint32_t cell, uint32_t symbol, uint32_t limit) {
  const uint32_t c = cell + (symbol ? 1u : 0x100u);
  const uint32_t h = ((c + 0x0101u) >> 1) & 0x7F7Fu;
  return (c & 0xFF) + (c >> 8) > limit ? h : c;          // selects only: lanes on different paths stay converged
}
AVR_DEC_INL void avr_s_ld_ms(avr_sptr a, uint64_t& m, uint32_t& sh) { uint64_t t; avr_s_ld_2x64_ro(a, m, t); sh = (uint32_t)t; }
AVR_DEC_INL void avr_ms_fill(const AvrTables* tab, int lane, int nlanes) {
  for (int t = lane; t < 98; t += nlanes) { g_dec.ms[t].m = tab->hot.div.m[t]; g_dec.ms[t].s = tab->hot.div.s[t]; }
}
// one symbol, warp-uniform (all lanes compute the same): recode.cpp:823-827 + arithmetic_code.h:232-248
AVR_DEC_INL uint32_t avr_dec_one(AvrDecRegs& r, avr_sptr s_ms, uint32_t cell) {
  uint64_t m; uint32_t sh;
  avr_s_ld_ms(s_ms + ((cell & 0xFF) + (cell >> 8)) * 16u, m, sh);
  const uint64_t r1 = (AVR_MULHI64(r.range, m) >> sh) * (uint64_t)(cell & 0xFF);
  const uint64_t r0 = r.range - r1;
  const uint32_t symbol = r.low >= r0;
  r.low -= symbol ? r0 : 0ull;
  r.range = symbol ? r1 : r0;
  if (r.range < (1ull << 51)) avr_dec_renorm(r, 1ull << 55);
  return symbol;
}

// ---- bin entry points outside residual blocks (warp-uniform; one copy each in the kernel) ----------
// Only low/range/ring cursor travel through shared memory per bin; the input window is touched by the (rare) renormalisation.
struct AvrLowRange { uint64_t low, range; };
AVR_DEC_FN AvrLowRange avr_dec_single_renorm(uint64_t low, uint64_t range) {     // by value in and out: no local memory
  AvrDecRegs r;
  avr_regs_load(r);
  r.low = low; r.range = range;
  avr_dec_renorm(r, 1ull << 55);
  AVR_WARP_SYNC();
  if (AVR_LANE0) { g_dec.win = r.win; g_dec.w_next = r.w_next; g_dec.pos = r.pos; g_dec.win_n = r.win_n; g_dec.carry = r.carry; g_dec.err = (int)r.err; }
  AvrLowRange o; o.low = r.low; o.range = r.range;
  return o;
}
AVR_DEC_INL int avr_dec_single(uint32_t key) {
  const avr_sptr s_base = avr_sbase();
  const avr_sptr s_ms = s_base + AVR_S_OFF(ms), s_ring = s_base + AVR_S_OFF(ring), a = s_base + AVR_S_OFF(est_def) + 2u * key;
  uint64_t low = g_dec.low, range = g_dec.range;
  uint32_t head = g_dec.ring_head;
  const uint32_t cell = avr_s_ld16(a);
  uint64_t m; uint32_t sh;
  avr_s_ld_ms(s_ms + ((cell & 0xFF) + (cell >> 8)) * 16u, m, sh);
  const uint64_t r1 = (AVR_MULHI64(range, m) >> sh) * (uint64_t)(cell & 0xFF);
  const uint64_t r0 = range - r1;
  const uint32_t sym = low >= r0;
  low -= sym ? r0 : 0ull;
  range = sym ? r1 : r0;
  if (range < (1ull << 51)) { const AvrLowRange o = avr_dec_single_renorm(low, range); low = o.low; range = o.range; }
  if (head + 1u - g_dec.ring_tail_seen > AVR_RING - 2) {         // rare: ring full
    uint32_t ts = g_dec.ring_tail_seen;
    avr_ring_reserve(head, ts, 1);
    AVR_WARP_SYNC();
    if (AVR_LANE0) g_dec.ring_tail_seen = ts;
  }
  AVR_WARP_SYNC();                                               // every lane has read the cell and the registers
  if (AVR_LANE0) {
    avr_s_st16(a, avr_cell_update(cell, sym, 0x60));
    avr_ring_store(s_ring, head, (key << 1) | sym);
    g_dec.low = low; g_dec.range = range;
    g_dec.ring_head = head + 1u;                                 // every bin posts one record: bins of the slice = head delta
  }
  AVR_WARP_SYNC();
  if (key == AVR_CTX_TERMINATE || ((head + 1u) & 15u) == 0) avr_ring_publish(head + 1u);
  return (int)sym;
}
AVR_DEC_FN int avr_dec_get(int ctx) { return avr_dec_single((uint32_t)ctx); }
AVR_DEC_FN int avr_dec_bypass() { return avr_dec_single(AVR_CTX_BYPASS); }
AVR_DEC_FN int avr_dec_terminate() {
  const int sym = avr_dec_single(AVR_CTX_TERMINATE);
  if (sym) g_dec.done = 1;
  return sym;
}

// ---- the speculative core ----------------------------------------------------------------------------
// cell[j] = estimator of step j on this lane's path (already carrying the updates of the path's earlier steps where
// keys repeat), nsteps = steps before the path leaves the phase.  Step j assumes symbol (lane >> j) & 1.  Returns the
// winning lane; low/range = its state after `nout` steps (it stopped early if its range fell below 2^51).
template <int D>
AVR_DEC_INL int avr_spec_chain(uint64_t& low, uint64_t& range, AvrPerLane<uint32_t> (&cell)[D], AvrPerLane<uint32_t>& nsteps,
                               avr_sptr s_ms, uint32_t& nout) {
  AvrPerLane<uint64_t> lo, rg;
  AvrPerLane<uint32_t> ok, cnt;
  AVR_LANE_LOOP(l) {
    uint64_t m[D]; uint32_t sh[D];
#pragma unroll
    for (int j = 0; j < D; j++) { const uint32_t c = cell[j][l]; avr_s_ld_ms(s_ms + ((c & 0xFF) + (c >> 8)) * 16u, m[j], sh[j]); }
    uint64_t L = low, R = range;
    uint32_t good = 1, n = 0;
    bool run = true;
    const uint32_t ns = nsteps[l];
#pragma unroll
    for (int j = 0; j < D; j++) {
      const bool act = run && (uint32_t)j < ns;
      const uint32_t a = ((uint32_t)l >> j) & 1u;
      const uint64_t r1 = (AVR_MULHI64(R, m[j]) >> sh[j]) * (uint64_t)(cell[j][l] & 0xFF);
      const uint64_t r0 = R - r1;
      const uint32_t sym = L >= r0;
      // selects, no branches: the lanes are on different paths and must stay converged
      good = act ? (good & (uint32_t)(sym == a)) : good;
      L = act ? L - (a ? r0 : 0ull) : L;
      R = act ? (a ? r1 : r0) : R;
      n += act ? 1u : 0u;
      run = run && R >= (1ull << 51);
    }
    lo[l] = L; rg[l] = R; ok[l] = good; cnt[l] = n;
  }
  const uint32_t mask = avr_lane_ballot(ok);                     // never empty: the true path is one of the lanes
  const int w = mask ? AVR_FFS(mask) - 1 : 0;
  low = avr_lane_get(lo, w); range = avr_lane_get(rg, w); nout = avr_lane_get(cnt, w);
  return w;
}
// updates of a path's earlier steps carried into later steps that use the same key
template <int D>
AVR_DEC_INL void avr_spec_forward(uint32_t (&cell)[D], const uint32_t (&key)[D], uint32_t lane, uint32_t ns, uint32_t limit) {
  uint32_t u[D];
#pragma unroll
  for (int j = 1; j < D; j++) {
    u[j - 1] = avr_cell_update(cell[j - 1], (lane >> (j - 1)) & 1u, limit);      // cell[j - 1] is final here
#pragma unroll
    for (int i = 0; i < j; i++) cell[j] = ((uint32_t)i < ns && key[i] == key[j]) ? u[i] : cell[j];
  }
}

// significant_coeff_flag / last_significant_coeff_flag ctxIdxInc (ITU-T H.264 Table 9-43, frame coded)
AVR_DEC_INL uint32_t avr_last8x8_ctx(uint32_t i) { return i == 0 ? 0 : i < 16 ? 1 : i < 32 ? 2 : i < 40 ? 3 : i < 48 ? 4 : i < 52 ? 5 : i < 56 ? 6 : i < 60 ? 7 : 8; }

// nnz side channel (recode.cpp:865-891, decompress order :1483-1499): the low nbits of the block's count, LSB first,
// bit i keyed by the bits before it.  One round covers up to 5 levels of the bit tree.
AVR_DEC_INL int avr_dec_nz_bits(AvrDecRegs& r, uint16_t* est, avr_sptr s_ms, const AvrNzCtx& nc, int nbits) {
  int stored = 0;
  for (int i0 = 0; i0 < nbits;) {
    AvrPerLane<uint32_t> cell[AVR_SPEC_D], key[AVR_SPEC_D], ns;
    const uint32_t left = (uint32_t)(nbits - i0);
    uint32_t line[AVR_SPEC_D];                                   // key of cell 0 of each level's line (depends on i only)
#pragma unroll
    for (int j = 0; j < AVR_SPEC_D; j++) line[j] = avr_nz_bit_key(nc, (i0 + j) < 6 ? i0 + j : 5, 0) - (1u << ((i0 + j) < 6 ? i0 + j : 5));
    AVR_LANE_LOOP(l) {
      const uint32_t nsl = left < AVR_SPEC_D ? left : AVR_SPEC_D;
      ns[l] = nsl;
#pragma unroll
      for (int j = 0; j < AVR_SPEC_D; j++) {
        const uint32_t i = (uint32_t)(i0 + j), so_far = (uint32_t)stored | (((uint32_t)l & ((1u << j) - 1u)) << i0);
        key[j][l] = line[j] + (1u << (i < 6 ? i : 5)) + (so_far & ((1u << (i < 6 ? i : 5)) - 1u));
        cell[j][l] = (uint32_t)j < nsl ? avr_g_ld16(est + key[j][l]) : AVR_EST_INIT;
      }
    }
    uint32_t n;
    const int w = avr_spec_chain<AVR_SPEC_D>(r.low, r.range, cell, ns, s_ms, n);
    AVR_LANE_LOOP(l) {
      if (l == w) {
#pragma unroll
        for (int j = 0; j < AVR_SPEC_D; j++)
          if ((uint32_t)j < n) avr_g_st16(est + key[j][l], avr_cell_update(cell[j][l], ((uint32_t)l >> j) & 1u, 0x60));
      }
    }
    AVR_WARP_SYNC();
    stored |= (int)(((uint32_t)w & ((1u << n) - 1u)) << i0);
    i0 += (int)n;
    if (r.range < (1ull << 51)) avr_dec_renorm(r, 1ull << 55);
  }
  return stored;
}

// Significance map (ffmpeg h264_cabac.c:1667-1722; reference order recode.cpp:1442-1456): MAP bins are range-decoded with
// key (cat, stored count, ones so far, position offset); the EOB bin after a one is inferred as stored == ones so far
// (recode.cpp:1444-1445) and only posted to the CABAC encoder.  Returns the ones decoded; *last_out as the walk leaves it.
template <bool IS64>
AVR_DEC_INL int avr_dec_sigmap(AvrDecRegs& r, uint16_t* est, avr_sptr s_ms, avr_sptr s_ring, uint32_t key0, int stored, int max_coeff,
                               int sig_base, int last_base, int* last_out) {
  const uint32_t coefs = (uint32_t)max_coeff - 1u;
  uint32_t last = 0, count = 0;
  for (;;) {
    AvrPerLane<uint32_t> cell[AVR_SPEC_D], key[AVR_SPEC_D], ns;
    avr_ring_reserve(r.head, r.tail_seen, 2 * AVR_SPEC_D);
    AVR_LANE_LOOP(l) {
      uint32_t k[AVR_SPEC_D], c[AVR_SPEC_D], nsl = AVR_SPEC_D;
#pragma unroll
      for (int j = AVR_SPEC_D - 1; j >= 0; j--) {                // the FIRST stopping step wins
        const uint32_t p = last + (uint32_t)j, cnt = count + (uint32_t)AVR_POPC((uint32_t)l & ((1u << j) - 1u)), a = ((uint32_t)l >> j) & 1u;
        const uint32_t pp = p < 63u ? p : 63u;
        k[j] = key0 + 16u * cnt + (IS64 ? (uint32_t)avr_sig8x8_off((int)pp) : (pp & 15u));
        if ((a && cnt + 1u == (uint32_t)stored) || p + 1u >= coefs) nsl = (uint32_t)j + 1u;
      }
#pragma unroll
      for (int j = 0; j < AVR_SPEC_D; j++) c[j] = (uint32_t)j < nsl ? avr_g_ld16(est + k[j]) : AVR_EST_INIT;
      if (IS64) avr_spec_forward<AVR_SPEC_D>(c, k, (uint32_t)l, nsl, 0x50);   // 8x8 offsets repeat: a later position can use a cell written on the way
#pragma unroll
      for (int j = 0; j < AVR_SPEC_D; j++) { key[j][l] = k[j]; cell[j][l] = c[j]; }
      ns[l] = nsl;
    }
    uint32_t n;
    const int w = avr_spec_chain<AVR_SPEC_D>(r.low, r.range, cell, ns, s_ms, n);
    // commit: the winner writes its cells and posts MAP (+ inferred EOB) records in bin order
    AVR_LANE_LOOP(l) {
      if (l == w) {
        uint32_t slot = r.head;
#pragma unroll
        for (int j = 0; j < AVR_SPEC_D; j++) {
          if ((uint32_t)j < n) {
            const uint32_t p = last + (uint32_t)j, cnt = count + (uint32_t)AVR_POPC((uint32_t)l & ((1u << j) - 1u)), a = ((uint32_t)l >> j) & 1u;
            avr_g_st16(est + key[j][l], avr_cell_update(cell[j][l], a, 0x50));
            avr_ring_store(s_ring, slot++, (((uint32_t)sig_base + (IS64 ? (uint32_t)avr_sig8x8_off((int)p) : p)) << 1) | a);
            if (a) avr_ring_store(s_ring, slot++, (((uint32_t)last_base + (IS64 ? avr_last8x8_ctx(p) : p)) << 1) | (uint32_t)(cnt + 1u == (uint32_t)stored));
          }
        }
      }
    }
    AVR_WARP_SYNC();
    const uint32_t bits = (uint32_t)w & ((1u << n) - 1u), ones = (uint32_t)AVR_POPC(bits);
    const uint32_t a_last = ((uint32_t)w >> (n - 1u)) & 1u;
    last += n; count += ones;
    r.head += n + ones;
    avr_ring_publish(r.head);
    if (r.range < (1ull << 51)) avr_dec_renorm(r, 1ull << 55);
    if (a_last && count == (uint32_t)stored) { last = (uint32_t)max_coeff; break; }   // EOB inferred 1: no implied last coefficient
    if (last >= coefs) break;
    if (r.err) break;
  }
  *last_out = (int)last;
  return (int)count;
}

// coeff_abs_level_minus1 + sign, reverse scan order (h264_cabac.c:1727-1768).  Per-lane state machine over the
// assumed symbols: B0 (first bin, ctx by node_ctx) -> GT1 run (one ctx per coefficient) -> SIGN (bypass); the
// Exp-Golomb escape of levels >= 15 is rare and decoded warp-uniformly.
struct AvrLvl { uint32_t ph, nc, ca, g2, cc; };                  // phase, node_ctx, coeff_abs, gt1 ctx offset, coefficients left
AVR_DEC_INL uint32_t avr_lvl_pack(const AvrLvl& s) { return s.ph | (s.nc << 3) | (s.ca << 6) | (s.g2 << 10) | (s.cc << 14); }
AVR_DEC_INL AvrLvl avr_lvl_unpack(uint32_t v) { AvrLvl s; s.ph = v & 7u; s.nc = (v >> 3) & 7u; s.ca = (v >> 6) & 15u; s.g2 = (v >> 10) & 15u; s.cc = v >> 14; return s; }
enum { AVR_LVL_B0 = 0, AVR_LVL_GT1 = 1, AVR_LVL_SIGN = 2, AVR_LVL_ESC = 3, AVR_LVL_DONE = 4 };

AVR_DEC_INL void avr_dec_levels(AvrDecRegs& r, avr_sptr s_ms, avr_sptr s_ring, avr_sptr s_def, int cat, int abs_base, int coeff_count) {
  const uint32_t gt1_max = cat == 3 ? 8u : 9u;
  AvrLvl s; s.ph = AVR_LVL_B0; s.nc = 0; s.ca = 0; s.g2 = 5; s.cc = (uint32_t)coeff_count;
  while (s.ph != AVR_LVL_DONE && !r.err) {
    if (s.ph == AVR_LVL_ESC) {                                   // coeff_abs >= 15: Exp-Golomb k = 0 in bypass bins
      const avr_sptr a_byp = s_def + 2u * AVR_CTX_BYPASS;
      uint32_t cell = avr_s_ld16(a_byp);
      uint32_t j = 0;
      avr_ring_reserve(r.head, r.tail_seen, 64);
      for (;;) {
        const uint32_t b = avr_dec_one(r, s_ms, cell);
        cell = avr_cell_update(cell, b, 0x60);
        if (AVR_LANE0) avr_ring_store(s_ring, r.head, (AVR_CTX_BYPASS << 1) | b);
        r.head++;
        if (!b || j >= 30) break;
        j++;
      }
      while (j--) {
        const uint32_t b = avr_dec_one(r, s_ms, cell);
        cell = avr_cell_update(cell, b, 0x60);
        if (AVR_LANE0) avr_ring_store(s_ring, r.head, (AVR_CTX_BYPASS << 1) | b);
        r.head++;
      }
      AVR_WARP_SYNC();
      if (AVR_LANE0) avr_s_st16(a_byp, cell);
      AVR_WARP_SYNC();
      s.ph = AVR_LVL_SIGN;
      continue;
    }
    AvrPerLane<uint32_t> cell[AVR_SPEC_D], key[AVR_SPEC_D], ns, st[AVR_SPEC_D];
    avr_ring_reserve(r.head, r.tail_seen, AVR_SPEC_D);
    AVR_LANE_LOOP(l) {
      uint32_t k[AVR_SPEC_D], c[AVR_SPEC_D], nsl = AVR_SPEC_D;
      AvrLvl t = s;
      bool open = true;
#pragma unroll
      for (int j = 0; j < AVR_SPEC_D; j++) {
        const uint32_t a = ((uint32_t)l >> j) & 1u;
        const bool b0 = t.ph == AVR_LVL_B0, gt = t.ph == AVR_LVL_GT1, sg = !b0 && !gt;
        const uint32_t i1 = t.nc < 4u ? t.nc + 1u : 0u;
        const uint32_t i2 = t.nc < 4u ? 5u : (t.nc + 2u > gt1_max ? gt1_max : t.nc + 2u);
        k[j] = sg ? (uint32_t)AVR_CTX_BYPASS : (uint32_t)abs_base + (b0 ? i1 : t.g2);
        // next state by selects (lanes are in different phases)
        const uint32_t nc1 = t.nc < 4u ? 4u : (t.nc < 7u ? t.nc + 1u : 7u), nc0 = t.nc < 3u ? t.nc + 1u : t.nc;
        const bool esc = gt && a && t.ca + 1u >= 15u, fin = sg && t.cc == 1u;
        const uint32_t ph_b0 = a ? (uint32_t)AVR_LVL_GT1 : (uint32_t)AVR_LVL_SIGN;
        const uint32_t ph_gt = a ? (esc ? (uint32_t)AVR_LVL_ESC : (uint32_t)AVR_LVL_GT1) : (uint32_t)AVR_LVL_SIGN;
        const uint32_t ph_sg = fin ? (uint32_t)AVR_LVL_DONE : (uint32_t)AVR_LVL_B0;
        t.g2 = (b0 && a) ? i2 : t.g2;
        t.nc = b0 ? (a ? nc1 : nc0) : t.nc;
        t.ca = b0 ? 2u : ((gt && a) ? t.ca + 1u : t.ca);
        t.cc = sg ? t.cc - 1u : t.cc;
        t.ph = b0 ? ph_b0 : gt ? ph_gt : ph_sg;
        st[j][l] = avr_lvl_pack(t);
        const bool stop = esc || fin;
        nsl = (stop && open) ? (uint32_t)j + 1u : nsl;
        open = open && !stop;
        t.ph = open ? t.ph : (uint32_t)AVR_LVL_SIGN;             // beyond the path's end: any state with valid keys
        t.cc = open ? t.cc : 2u;
      }
#pragma unroll
      for (int j = 0; j < AVR_SPEC_D; j++) c[j] = (uint32_t)j < nsl ? avr_s_ld16(s_def + 2u * k[j]) : AVR_EST_INIT;
      avr_spec_forward<AVR_SPEC_D>(c, k, (uint32_t)l, nsl, 0x60);
#pragma unroll
      for (int j = 0; j < AVR_SPEC_D; j++) { key[j][l] = k[j]; cell[j][l] = c[j]; }
      ns[l] = nsl;
    }
    uint32_t n;
    const int w = avr_spec_chain<AVR_SPEC_D>(r.low, r.range, cell, ns, s_ms, n);
    AvrPerLane<uint32_t> fin;
    AVR_LANE_LOOP(l) {
      uint32_t f = st[0][l];
#pragma unroll
      for (int j = 1; j < AVR_SPEC_D; j++) if (n == (uint32_t)j + 1u) f = st[j][l];
      fin[l] = f;
      if (l == w) {
#pragma unroll
        for (int j = 0; j < AVR_SPEC_D; j++) {
          if ((uint32_t)j < n) {
            const uint32_t a = ((uint32_t)l >> j) & 1u;
            avr_s_st16(s_def + 2u * key[j][l], avr_cell_update(cell[j][l], a, 0x60));
            avr_ring_store(s_ring, r.head + (uint32_t)j, (key[j][l] << 1) | a);
          }
        }
      }
    }
    AVR_WARP_SYNC();
    s = avr_lvl_unpack(avr_lane_get(fin, w));
    r.head += n;
    avr_ring_publish(r.head);
    if (r.range < (1ull << 51)) avr_dec_renorm(r, 1ull << 55);
  }
}

// nnz rows of the current macroblock, its left / above neighbours and the co-located macroblock of the previous
// picture (reference FrameBuffer::meta_at, framebuffer.h:75-86; neighbours recode.cpp:426-478), 64 B each, staged
// by the 32 lanes: 8 lanes x 8 bytes per row.  Rows that do not exist read as zero.
AVR_DEC_INL void avr_nb_stage() {
  const uint8_t* cur = g_dec.store[g_dec.cur];
  const uint8_t* prev = g_dec.store[g_dec.cur ^ 1];
  const uint32_t mb = g_dec.mb_addr, mb_w = (uint32_t)g_dec.mb_w;
  AVR_LANE_LOOP(l) {
    const int which = l >> 3, part = l & 7;
    const uint8_t* src = nullptr;
    if (which == 0) src = cur + (size_t)mb * AVR_MB_STRIDE;
    else if (which == 1) { if (g_dec.mb_x > 0) src = cur + (size_t)(mb - 1u) * AVR_MB_STRIDE; }
    else if (which == 2) { if (mb >= mb_w) src = cur + (size_t)(mb - mb_w) * AVR_MB_STRIDE; }
    else src = prev + (size_t)mb * AVR_MB_STRIDE;
    uint64_t v = 0;
    if (src) v = avr_g_ld64(src + part * 8);
    *reinterpret_cast<uint64_t*>(&g_dec.nb[which][part * 8]) = v;
  }
  AVR_WARP_SYNC();
  g_dec.nb_staged = 1;
}
// avr_nz_ctx_x (avr_chain.h) on the staged rows
AVR_DEC_INL void avr_nz_ctx_staged(AvrNzCtx* c, int n, int size, int cat, int is_dc) {
  int cross, nn;
  nn = avr_neighbor_n(n, 0, size, &cross);
  c->has_left = !(cross && g_dec.mb_x == 0);
  c->left_nz = c->has_left ? g_dec.nb[cross ? 1 : 0][nn] : 0;
  nn = avr_neighbor_n(n, 1, size, &cross);
  const int has_above = !(cross && g_dec.mb_addr < (uint32_t)g_dec.mb_w);
  c->above_nz = has_above ? g_dec.nb[cross ? 2 : 0][nn] : 0;
  c->prev_nz = g_dec.nb[3][n];
  const int is8 = g_dec.nb[0][AVR_MB_IS8X8] || size > 32;
  c->third = is8 + 2 * is_dc + 4 * cat;
}

// ---- one residual block (ffmpeg h264_cabac.c:1583-1780 walk; reference order recode.cpp:1483-1499)
// nnz side channel -> significance map (EOB bins inferred) -> levels/signs.  Returns coeff_count.
AVR_DEC_FN int avr_dec_residual_block(int cat, int n, int max_coeff, int is_dc, int sig_base, int last_base, int abs_base) {
  AvrDecRegs r;
  avr_regs_load(r);
  uint16_t* const est = g_dec.est;
  const avr_sptr s_base = avr_sbase();
  const avr_sptr s_ms = s_base + AVR_S_OFF(ms), s_ring = s_base + AVR_S_OFF(ring), s_def = s_base + AVR_S_OFF(est_def);
  if (!g_dec.nb_staged) avr_nb_stage();
  // nnz bits first.  The count is recovered completely (a block that codes as 0 in 2/4 bits is a
  // full one) so that the store holds what the COMPRESSOR held.
  AvrNzCtx nc;
  avr_nz_ctx_staged(&nc, n, max_coeff, cat, is_dc);
  int stored = avr_dec_nz_bits(r, est, s_ms, nc, avr_nz_bits(max_coeff));
  if (stored == 0 && max_coeff <= 16) stored = max_coeff;
  // significance map
  const uint32_t key0 = avr_key_sig(cat, stored, 0, 0);
  int coeff_count = 0, last = 0;
  if (max_coeff > 1 && !r.err) {
    coeff_count = max_coeff == 64 ? avr_dec_sigmap<true>(r, est, s_ms, s_ring, key0, stored, max_coeff, sig_base, last_base, &last)
                                  : avr_dec_sigmap<false>(r, est, s_ms, s_ring, key0, stored, max_coeff, sig_base, last_base, &last);
  }
  if (last == max_coeff - 1) coeff_count++;
  AVR_WARP_SYNC();
  if (AVR_LANE0) {
    uint8_t* mb = g_dec.store[g_dec.cur] + (size_t)g_dec.mb_addr * AVR_MB_STRIDE;
    if (max_coeff > 32) { mb[AVR_MB_IS8X8] = 1; g_dec.nb[0][AVR_MB_IS8X8] = 1; }
    mb[n] = (uint8_t)(coeff_count & 63); g_dec.nb[0][n] = (uint8_t)(coeff_count & 63);
  }
  AVR_WARP_SYNC();
  if (coeff_count > 0 && !r.err) avr_dec_levels(r, s_ms, s_ring, s_def, cat, abs_base, coeff_count);
  avr_regs_store(r);
  return coeff_count;
}

// Bin I/O policy for avr::SliceParser: stateless, forwards to the functions above.
struct AvrDecPolicy {
  static constexpr bool kFastResidual = true;
  AVR_DEC_M int residual_block(int cat, int n, int mc, int dc, int sig, int last, int abs_) { return avr_dec_residual_block(cat, n, mc, dc, sig, last, abs_); }
  AVR_DEC_M int get(int ctx) { return avr_dec_get(ctx); }
  AVR_DEC_M int bypass(int) { return avr_dec_bypass(); }
  AVR_DEC_M int terminate(int) { return avr_dec_terminate(); }
  AVR_DEC_M void mb_begin(int x, int y) {
    AVR_WARP_SYNC();
    if (AVR_LANE0) { g_dec.mb_addr = (uint32_t)(x + y * g_dec.mb_w); g_dec.mb_x = (uint32_t)x; g_dec.nb_staged = 0; }
    AVR_WARP_SYNC();
  }
  AVR_DEC_M void sub_mb_begin(int, int, int, int, int) {}
  AVR_DEC_M void sub_mb_end() {}
  AVR_DEC_M void sigmap_begin() {}
  AVR_DEC_M void sigmap_end() {}
  AVR_DEC_M void hint(int, int) {}
};

// ---- per-file / per-slice driver ---------------------------------------------------------------
AVR_DEC_INL void avr_dec_file_begin(uint16_t* est, uint8_t* s0, uint8_t* s1) {   // one thread
  g_dec.self = (uint64_t)avr_sp(&g_dec);
  g_dec.est = est; g_dec.store[0] = s0; g_dec.store[1] = s1; g_dec.cur = 0; g_dec.mb_w = 1;
  g_dec.ring_head = 0; g_dec.ring_tail_seen = 0; g_dec.ring_tail = 0; g_dec.ring_pub = 0;
  g_dec.nb_staged = 0; g_dec.err = 0;
}
// update_frame_spec flip (recode.cpp:831-850): one thread toggles, then every thread zeroes its share
AVR_DEC_INL void avr_dec_flip_toggle(const AvrDecSlice& ds) { g_dec.cur ^= 1; g_dec.mb_w = ds.sp.mb_w; }
AVR_DEC_INL void avr_dec_flip_zero(const AvrDecSlice& ds, int lane, int nlanes) {
  AvrChain::zero_stores(g_dec.store[g_dec.cur], ds.flip == 1 ? nullptr : g_dec.store[g_dec.cur ^ 1], ds.sp.mb_w, ds.sp.mb_h, lane, nlanes);
}
// One coded slice on the chain WARP (all 32 lanes call it).  g_dec.cab_ent must hold the slice's initial CABAC
// entries and the encoder thread must be consuming the ring.  Returns AVR_OK or an error (after an error the model
// state of the file is meaningless).  The CABAC byte count / overflow flag come from the encoder.
AVR_DEC_FN int avr_dec_slice(const AvrDecSlice& ds, const uint8_t* payloads, avr::MbInfo* row, uint32_t uid, uint32_t* out_bins) {
  if (uid == 1) for (int x = 0; x < ds.sp.mb_w; x++) row[x].uid = 0;
  AVR_WARP_SYNC();
  avr_dec_init_range(payloads + ds.payload_off, ds.payload_len);
  if (AVR_LANE0) { g_dec.done = 0; g_dec.mb_addr = 0; }
  AVR_WARP_SYNC();
  const uint32_t head0 = g_dec.ring_head;                      // every bin posts exactly one record
  AvrDecPolicy io;
  avr::SliceParser<AvrDecPolicy> parser(io, ds.sp, row, uid);
  const int rc = parser.run();
  AVR_WARP_SYNC();
  const uint32_t head1 = g_dec.ring_head;
  if (AVR_LANE0) *out_bins = head1 - head0;
  if (!g_dec.done) {                                            // release the encoder thread
    AvrDecRegs r;
    avr_regs_load(r);
    avr_ring_reserve(r.head, r.tail_seen, 1);
    if (AVR_LANE0) avr_ring_store(avr_sbase() + AVR_S_OFF(ring), r.head, AVR_REC_ABORT << 1);
    r.head++;
    avr_regs_store(r);
  }
  avr_ring_publish(g_dec.ring_head);
  return rc < 0 ? AVR_E_CORRUPT : (!g_dec.done || g_dec.err) ? AVR_E_CORRUPT : g_dec.pos > ds.payload_len + 32 ? AVR_E_CORRUPT : AVR_OK;
}

#if !defined(__CUDACC__)
// Whole file on one host thread (tests/host_emul.cpp); the kernel in avr_kernels.cu runs the same
// steps with the bulk work spread over the block and the CABAC encoder on its own warp.
static void avr_decompress_file_host(const AvrFileDesc& fd, const AvrDecSlice* slices_all, const uint8_t* payloads,
                                     uint16_t* est, uint8_t* store0, uint8_t* store1, avr::MbInfo* row, const AvrTables* tab,
                                     uint8_t* out, uint32_t* out_len, uint32_t* out_bins, int* status) {
  avr_ms_fill(tab, 0, 1);
  avr_cab_fill_trans(tab, 0, 1);
  for (uint32_t i = 0; i < AVR_NUM_CTX_KEYS; i++) g_dec.est_def[i] = AVR_EST_INIT;
  avr_dec_file_begin(est, store0, store1);
  uint32_t uid = 0;
  int failed = 0;
  for (uint32_t s = 0; s < fd.nslices; s++) {
    const AvrDecSlice& ds = slices_all[fd.slice_off + s];
    const uint32_t g = fd.slice_off + s;
    if (failed) { out_len[g] = 0; out_bins[g] = 0; status[g] = AVR_E_CORRUPT; continue; }
    if (ds.flip) { avr_dec_flip_toggle(ds); avr_dec_flip_zero(ds, 0, 1); }
    avr_cab_init_entries(tab, ds.sp.slice_type == avr::SLICE_I, ds.sp.cabac_init_idc, ds.sp.slice_qp, 0, 1);
    g_cab_host.low = 0; g_cab_host.range = (1u << 22) * 0x1FEu; g_cab_host.n = 0; g_cab_host.cap = ds.out_cap;
    g_cab_host.out = out + ds.out_off; g_cab_host.err = 0;
    int rc = avr_dec_slice(ds, payloads, row, ++uid, &out_bins[g]);
    out_len[g] = g_cab_host.n;
    status[g] = rc ? rc : g_cab_host.err;
    if (status[g]) failed = 1;
  }
}
#endif
