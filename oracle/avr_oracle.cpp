// TEST INFRASTRUCTURE -- CPU restatement of the reference's entropy hot path.  Never linked
// into, called from, or shipped with the product (avrecode-ms_b200/); only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline leg may run it.
//
// Parity status: PINNED.  tests/test_oracle.py checks the payloads this program produces
// against the `cabac` fields of the .avrec written by the unmodified reference compressor
// (oracle/_ref/recode, built by oracle/build_ref.sh) on the same clips, and against the
// golden fixtures under tests/golden/ minted from that binary (tests/golden/make_golden.py).
//
// Input: the hook-event trace printed by oracle/ref_harness.cpp (the reference's own
// libavcodec-hooks parse), one event per line:
//   F cur w h | S size off | I <states hex> | M x y | B cat n max_coeff is_dc c422 | [ | ] | c ctx sym | b sym | t sym
// A plain sequential program that follows the reference statement by statement:
//   estimator + thresholds   recode.cpp:823-827, 1037-1054, 1064
//   model key                recode.cpp:683-822 (live part: 685-690, 703-712, 800, 805-816)
//   sig-map state tracking   recode.cpp:958-1033
//   nnz side channel         recode.cpp:851-937, neighbour geometry 426-478 / 270-319
//   frame store              recode.cpp:831-850, framebuffer.h:33-36
//   compress bin order       recode.cpp:1075-1103, 1167-1262
//   decompress bin order     recode.cpp:1442-1527
//   range coder              arithmetic_code.h:106-190 (encoder), 220-288 (decoder)
//   CABAC encoder            cabac_code.h:30-67 over arithmetic_code<uint32_t,uint16_t,0x200>
//
// Usage:
//   avr_oracle encode <trace> > payloads      one line "P <hex>" per coded slice
//   avr_oracle decode <trace> <payloads>      decodes every slice with the trace's CONTEXT
//        sequence only (symbols are used to verify), prints "C <hex>" = re-encoded CABAC bytes
//        per slice and "OK bins=<n>" or "MISMATCH ..." (exit 1)
//   avr_oracle bench <trace> <reps>           times encode and decode, prints Mbins/s
//   avr_oracle streams <digit_bits> <model> <nstreams> <nbins> <bits file> [quiet]
//        BASELINE configs[0]: the coder test of test/arithmetic_code.cpp:93-110 per stream; prints "A <hex>"
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <chrono>
#include <map>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>

typedef uint64_t range_t;

// ---------------------------------------------------------------------------------------
// Generic range coder, arithmetic_code<uint64_t,uint8_t> (recode.cpp:322-323)
// ---------------------------------------------------------------------------------------
struct RangeEncoder64 {
  static constexpr uint64_t fixed_one = 1ull << 63;
  static constexpr uint64_t min_range = (fixed_one / 256) / 16;   // arithmetic_code.h:61-62
  static constexpr uint64_t msd = fixed_one / 256;                // most_significant_digit
  uint64_t low = 0, range = fixed_one;
  std::vector<uint8_t> overflow;
  std::vector<uint8_t>* out;
  explicit RangeEncoder64(std::vector<uint8_t>* o) : out(o) {}
  void put(int symbol, uint64_t range_of_1) {                      // arithmetic_code.h:106-126
    uint64_t range_of_0 = range - range_of_1;
    if (symbol) { low += range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < min_range) {
      if (range == 0) { fprintf(stderr, "zero-probability symbol\n"); exit(2); }
      while (range < fixed_one / 256) renorm();
    }
  }
  void renorm() {                                                  // arithmetic_code.h:148-180
    if (low >= fixed_one) {
      for (int i = (int)overflow.size() - 1; i >= 0; i--) if (++overflow[i] != 0) break;
      low -= fixed_one;
    }
    uint8_t digit = (uint8_t)(low / msd);
    if (digit != (uint8_t)((low + range - 1) / msd)) {
      overflow.push_back(digit);
    } else {
      for (uint8_t d : overflow) out->push_back(d);
      overflow.clear();
      out->push_back(digit);
    }
    low -= digit * msd;
    low *= 256;
    range *= 256;
  }
  void finish() {                                                  // arithmetic_code.h:128-144
    for (uint64_t stop = fixed_one >> 1; stop > 0; stop >>= 1) {
      uint64_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) { range = 1; renorm(); }
    range = 0;
  }
};

struct RangeDecoder64 {
  static constexpr uint64_t fixed_one = 1ull << 63;
  static constexpr uint64_t min_range = (fixed_one / 256) / 16;
  const uint8_t* in; const uint8_t* end;
  uint8_t next_digit;
  uint64_t low, range;
  RangeDecoder64(const uint8_t* b, const uint8_t* e) : in(b), end(e) {  // arithmetic_code.h:220-230
    next_digit = aligned();
    low = next_digit / 2;        // digit_alignment = 2 (arithmetic_code.h:251-252)
    range = 256 / 2;
    while (range < fixed_one) consume();
  }
  uint8_t aligned() { return in != end ? *in++ : 0; }              // arithmetic_code.h:278-287
  void consume() {                                                 // arithmetic_code.h:259-275
    uint8_t d = aligned();
    uint8_t digit = (uint8_t)((next_digit * 128) | (d / 2));
    next_digit = d;
    low = low * 256 + digit;
    range *= 256;
  }
  int get(uint64_t range_of_1) {                                   // arithmetic_code.h:232-248
    uint64_t range_of_0 = range - range_of_1;
    int symbol = low >= range_of_0;
    if (symbol) { low -= range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < min_range) while (range < fixed_one / 256) consume();
    return symbol;
  }
};

// ---------------------------------------------------------------------------------------
// The same coder for any digit width (test/arithmetic_code.cpp:93 uses <uint64_t,uint16_t>): put-time digits are
// Digit wide and leave as bytes MSB first (arithmetic_code.h:185-190); finish() flushes 8-bit digits
// (renormalize_and_emit_digit<OutputDigit>, arithmetic_code.h:139-142); `overflow` holds Digit-wide entries.
// ---------------------------------------------------------------------------------------
template <typename Digit>
struct RangeEncoderT {
  static constexpr uint64_t fixed_one = 1ull << 63;
  static constexpr uint64_t base = 1ull << (8 * sizeof(Digit));
  static constexpr uint64_t min_range = (fixed_one / base) / 16;
  uint64_t low = 0, range = fixed_one;
  std::vector<Digit> overflow;
  std::vector<uint8_t>* out;
  explicit RangeEncoderT(std::vector<uint8_t>* o) : out(o) {}
  void emit(Digit d) { for (int i = (int)sizeof(Digit) - 1; i >= 0; i--) out->push_back((uint8_t)(d >> (8 * i))); }
  void put(int symbol, uint64_t range_of_1) {
    uint64_t range_of_0 = range - range_of_1;
    if (symbol) { low += range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < min_range) {
      if (range == 0) { fprintf(stderr, "zero-probability symbol\n"); exit(2); }
      while (range < fixed_one / base) renorm<Digit>();
    }
  }
  template <typename D>
  void renorm() {                                                  // arithmetic_code.h:148-180 with Digit = D
    const uint64_t b = 1ull << (8 * sizeof(D)), msd = fixed_one / b;
    if (low >= fixed_one) {
      for (int i = (int)overflow.size() - 1; i >= 0; i--) if (++overflow[i] != 0) break;
      low -= fixed_one;
    }
    D digit = (D)(low / msd);
    if (digit != (D)((low + range - 1) / msd)) {
      overflow.push_back(digit);                                   // stored as CompressedDigit
    } else {
      for (Digit d : overflow) emit(d);                            // emit_digit(CompressedDigit)
      overflow.clear();
      for (int i = (int)sizeof(D) - 1; i >= 0; i--) out->push_back((uint8_t)(digit >> (8 * i)));
    }
    low -= digit * msd;
    low *= b;
    range *= b;
  }
  void finish() {
    for (uint64_t stop = fixed_one >> 1; stop > 0; stop >>= 1) {
      uint64_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) { range = 1; renorm<uint8_t>(); }
    range = 0;
  }
};
template <typename Digit>
struct RangeDecoderT {
  static constexpr uint64_t fixed_one = 1ull << 63;
  static constexpr uint64_t base = 1ull << (8 * sizeof(Digit));
  static constexpr uint64_t min_range = (fixed_one / base) / 16;
  const uint8_t* in; const uint8_t* end;
  Digit next_digit;
  uint64_t low, range;
  RangeDecoderT(const uint8_t* b, const uint8_t* e) : in(b), end(e) {
    next_digit = aligned();
    low = next_digit / 2;
    range = base / 2;
    while (range < fixed_one) consume();
  }
  Digit aligned() {                                                // arithmetic_code.h:278-287
    Digit d = 0;
    for (int i = (int)sizeof(Digit) - 1; i >= 0; i--) { d = (Digit)(d * 256); if (in != end) d |= *in++; }
    return d;
  }
  void consume() {
    Digit d = aligned();
    Digit digit = (Digit)((next_digit * (base / 2)) | (d / 2));
    next_digit = d;
    low = low * base + digit;
    range *= base;
  }
  int get(uint64_t range_of_1) {
    uint64_t range_of_0 = range - range_of_1;
    int symbol = low >= range_of_0;
    if (symbol) { low -= range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < min_range) while (range < fixed_one / base) consume();
    return symbol;
  }
};
struct StreamCounts {                                              // recode.cpp:1064, 1037-1054 (limit 0x60), 823-827
  int pos = 1, neg = 1;
  uint64_t p1(uint64_t range) const { return (range / (uint64_t)(pos + neg)) * (uint64_t)pos; }
  void update(int symbol) {
    if (symbol) pos++; else neg++;
    if (pos + neg > 0x60) { pos = (pos + 1) / 2; neg = (neg + 1) / 2; }
  }
};
// test/arithmetic_code.cpp:93-110 for one stream; model 0 = `range/2`, model 1 = adaptive counts
template <typename Digit>
static bool stream_roundtrip(int model, const uint8_t* bits, long nbins, std::vector<uint8_t>* out) {
  {
    RangeEncoderT<Digit> enc(out);
    StreamCounts c;
    for (long i = 0; i < nbins; i++) {
      int bit = bits[i] != 0;
      enc.put(bit, model == 0 ? enc.range / 2 : c.p1(enc.range));
      if (model) c.update(bit);
    }
    enc.finish();
  }
  RangeDecoderT<Digit> dec(out->data(), out->data() + out->size());
  StreamCounts c;
  for (long i = 0; i < nbins; i++) {
    int bit = dec.get(model == 0 ? dec.range / 2 : c.p1(dec.range));
    if (model) c.update(bit);
    if (bit != (bits[i] != 0)) return false;
  }
  return true;
}

// ---------------------------------------------------------------------------------------
// CABAC encoder = arithmetic_code<uint32_t,uint16_t,0x200>::encoder<..., uint8_t>
// (cabac_code.h:16-67).  Put-time renormalisation works in 16-bit digits, finish() in 8-bit
// digits (arithmetic_code.h:113-122 vs 139-142); pending overflow digits are always 16-bit.
// Constant tables: ITU-T H.264 Table 9-44/9-45 (data only; same numbers as
// ffmpeg/libavcodec/cabac.c:36-165 in the standard's layout).
// ---------------------------------------------------------------------------------------
#include "../avrecode-ms_b200/csrc/h264_std_tables.h"

struct CabacEncoderRef {
  static constexpr uint32_t fixed_one = 1u << 31;
  uint32_t low = 0, range = (fixed_one / 0x200) * 0x1FE;           // cabac_code.h:30
  std::vector<uint16_t> overflow;
  std::vector<uint8_t>* out;
  explicit CabacEncoderRef(std::vector<uint8_t>* o) : out(o) {}
  static int ilog2(uint32_t x) { int i = 0; while (x >>= 1) i++; return i; }
  void emit16(uint16_t d) { out->push_back((uint8_t)(d >> 8)); out->push_back((uint8_t)d); }
  void carry() {
    if (low >= fixed_one) {
      for (int i = (int)overflow.size() - 1; i >= 0; i--) if (++overflow[i] != 0) break;
      low -= fixed_one;
    }
  }
  void renorm16() {
    const uint32_t msd = fixed_one >> 16;
    carry();
    uint16_t digit = (uint16_t)(low / msd);
    if (digit != (uint16_t)((low + range - 1) / msd)) {
      overflow.push_back(digit);
    } else {
      for (uint16_t d : overflow) emit16(d);
      overflow.clear();
      emit16(digit);
    }
    low -= digit * msd; low *= 65536u; range *= 65536u;
  }
  void renorm8() {
    const uint32_t msd = fixed_one >> 8;
    carry();
    uint8_t digit = (uint8_t)(low / msd);
    if (digit != (uint8_t)((low + range - 1) / msd)) {
      overflow.push_back(digit);
    } else {
      for (uint16_t d : overflow) emit16(d);
      overflow.clear();
      out->push_back(digit);
    }
    low -= digit * msd; low *= 256u; range *= 256u;
  }
  void put_raw(int symbol, uint32_t range_of_1) {
    uint32_t range_of_0 = range - range_of_1;
    if (symbol) { low += range_of_0; range = range_of_1; } else { range = range_of_0; }
    if (range < 0x200) {
      if (range == 0) { fprintf(stderr, "cabac zero range\n"); exit(2); }
      while (range < fixed_one / 65536u) renorm16();
    }
  }
  void finish() {
    for (uint32_t stop = fixed_one >> 1; stop > 0; stop >>= 1) {
      uint32_t x = (low | stop) & ~(stop - 1);
      if (stop < range && low <= x && x < low + range) { low = x; break; }
    }
    while (low != 0) { range = 1; renorm8(); }
    range = 0;
  }
  void put(int symbol, uint8_t* state) {                           // cabac_code.h:33-49
    uint32_t s = *state, p = s >> 1, mps = s & 1;
    int lps = symbol != (int)mps;
    int normalize = ilog2(range / 0x100);
    uint32_t approx = range >> (normalize - 1);
    uint32_t rlps = (uint32_t)avr_range_tab_lps[(p << 2) | ((approx >> 7) & 3)] << normalize;
    put_raw(lps, rlps);
    if (lps) { if (p == 0) mps ^= 1; p = avr_trans_idx_lps[p]; } else if (p < 62) p++;
    *state = (uint8_t)((p << 1) | mps);
  }
  void put_bypass(int symbol) { put_raw(symbol, range / 2); }      // cabac_code.h:52-54
  void put_terminate(int eos) {                                    // cabac_code.h:57-67
    int normalize = ilog2(range / 0x100);
    put_raw(eos, 2u << normalize);
    if (eos) finish();
  }
};

// ---------------------------------------------------------------------------------------
// h264_model restated (recode.cpp:625-1066)
// ---------------------------------------------------------------------------------------
enum CT { CT_UNKNOWN, CT_UNREACHABLE, CT_SIGMAP, CT_EOB, CT_RESIDUALS, CT_NZ };   // coding_hooks.h:10-22
enum { KEY_CTX = 0, KEY_BYPASS, KEY_TERMINATE, KEY_SIG, KEY_FAKE, KEY_NZBIT0 /* ..+5 */ };  // distinct "addresses"

struct Est { int pos = 1, neg = 1; };                               // recode.cpp:1064
struct MbState {                                                    // live fields of Block/BlockMeta (block.h)
  uint8_t num_nonzeros[51];
  uint8_t is_8x8, coded;
  uint8_t residual[816];
};
struct Frame {
  int w = 0, h = 0;
  std::vector<MbState> mb;
  void init(int w_, int h_) { w = w_; h = h_; mb.assign((size_t)w * h, MbState()); bzero(); }
  void bzero() { if (!mb.empty()) memset(mb.data(), 0, mb.size() * sizeof(MbState)); }
  MbState& at(int x, int y) { return mb[(size_t)x + (size_t)y * w]; }
};

struct Model {
  std::map<std::tuple<int, int, int>, Est> est;                     // recode.cpp:325,1065 (never cleared)
  Frame frames[2];
  int cur = 0;
  int coding_type = CT_UNKNOWN;
  int mb_x = 0, mb_y = 0, n = 0, zz = 0;                            // mb_coord
  int observed = 0, cat = -1, size = -1, is_dc = 0, c422 = 0;

  // The trace gives cur_frame/w/h at each slice; that is the observable effect of
  // update_frame_spec (recode.cpp:831-850): a flip re-inits (size change) or zeroes the new
  // current frame, the other frame keeps the previous picture.
  void frame_observed(int cur_, int w, int h) {
    if (cur_ == cur && frames[cur].w == w && frames[cur].h == h) return;
    cur = cur_;
    if (frames[cur].w != w || frames[cur].h != h) {
      frames[cur].init(w, h);
      if (frames[!cur].w != w || frames[!cur].h != h) frames[!cur].init(w, h);
    } else {
      frames[cur].bzero();
    }
  }
  Est& E(std::tuple<int, int, int> k) { return est[k]; }
  static range_t p1(const Est& e, range_t range) { return (range / (uint64_t)(e.pos + e.neg)) * (uint64_t)e.pos; }  // :823-827

  std::tuple<int, int, int> key(int ctx_kind, int ctx) {            // get_model_key, recode.cpp:683-822
    switch (coding_type) {
      case CT_SIGMAP: {
        static const uint8_t sig8x8[63] = {0, 1, 2, 3, 4, 5, 5, 4, 4, 3, 3, 4, 4, 4, 5, 5, 4, 4, 4, 4, 3, 3, 6, 7, 7, 7, 8, 9, 10, 9, 8, 7,
                                           7, 6, 11, 12, 13, 11, 6, 7, 8, 9, 14, 10, 9, 8, 6, 11, 12, 13, 11, 6, 9, 14, 10, 9, 11, 12, 13, 11, 14, 10, 12};
        static const int cat_lookup[14] = {105, 120, 134, 149, 152, 402, 484, 499, 513, 660, 528, 543, 557, 718};
        static const uint8_t dc422[7] = {0, 0, 1, 1, 2, 2, 2};
        int zo = zz;
        if (is_dc && c422) zo = dc422[zz];
        else if (size > 32) zo = sig8x8[zz];
        int nnz = frames[cur].at(mb_x, mb_y).num_nonzeros[n];
        return std::make_tuple(KEY_SIG, 64 * nnz + observed, is_dc + zo * 2 + 32 * cat_lookup[cat]);
      }
      case CT_EOB: {
        int nnz = frames[cur].at(mb_x, mb_y).num_nonzeros[n];
        return std::make_tuple(KEY_FAKE, nnz == observed, 0);
      }
      default:
        return std::make_tuple(ctx_kind, ctx, 0);
    }
  }
  void begin_sigmap() {                                             // begin_coding_type, :958-981
    coding_type = CT_SIGMAP;
    frames[cur].at(mb_x, mb_y).num_nonzeros[n] = 0;
    observed = 0; zz = 0;
  }
  void end_sigmap() {                                               // end_coding_type, :938-957
    MbState& m = frames[cur].at(mb_x, mb_y);
    uint8_t c = 0;
    for (int i = 0; i < size; i++) if (m.residual[n * 16 + i]) c++;
    m.is_8x8 = m.is_8x8 || size > 32;
    m.coded = 1;
    m.num_nonzeros[n] = c;
    coding_type = CT_UNKNOWN;
  }
  void track(int symbol) {                                          // update_state_tracking, :987-1033
    MbState& m = frames[cur].at(mb_x, mb_y);
    switch (coding_type) {
      case CT_SIGMAP:
        m.residual[n * 16 + zz] = (uint8_t)symbol;
        observed += symbol;
        if (zz + 1 == size) { coding_type = CT_UNREACHABLE; zz = 0; }
        else if (symbol) coding_type = CT_EOB;
        else {
          ++zz;
          if (zz + 1 == size) { m.residual[n * 16 + zz] = 1; ++observed; coding_type = CT_UNREACHABLE; zz = 0; }
        }
        break;
      case CT_EOB:
        if (symbol) { zz = 0; coding_type = CT_UNREACHABLE; }
        else if (zz + 2 == size) { m.residual[n * 16 + zz + 1] = 1; coding_type = CT_UNREACHABLE; }
        else { coding_type = CT_SIGMAP; ++zz; }
        break;
      default: break;
    }
  }
  void update(int symbol, std::tuple<int, int, int> k) {            // update_state_for_model_key, :1037-1054
    Est& e = est[k];
    if (symbol) e.pos++; else e.neg++;
    if ((coding_type != CT_SIGMAP && e.pos + e.neg > 0x60) || (coding_type == CT_SIGMAP && e.pos + e.neg > 0x50)) {
      e.pos = (e.pos + 1) / 2; e.neg = (e.neg + 1) / 2;
    }
    track(symbol);
  }

  // get_neighbor_sub_mb (recode.cpp:426-478) with the scan_8 grid (:270-319) expressed by
  // block coordinates: luma/Cb/Cr planes are 4x4 grids of 4x4 blocks, the left neighbour of
  // column 0 is column 3 of MB (x-1,y), the one above row 0 is row 3 of MB (x,y-1).
  bool neighbor(bool above, int* ox, int* oy, int* on) const {
    *ox = mb_x; *oy = mb_y; *on = n;
    if (n >= 48) {
      if (above) { if (mb_y > 0) { *oy = mb_y - 1; return true; } return false; }
      if (mb_x > 0) { *ox = mb_x - 1; return true; }
      return false;
    }
    int plane = n / 16, i = n % 16;
    int x = (i & 1) | ((i >> 1) & 2), y = ((i >> 1) & 1) | ((i >> 2) & 2);
    if (above) { if (y == 0) { if (mb_y == 0) return false; *oy = mb_y - 1; y = 3; } else y--; }
    else       { if (x == 0) { if (mb_x == 0) return false; *ox = mb_x - 1; x = 3; } else x--; }
    int nn = plane * 16 + ((x & 1) | ((y & 1) << 1) | ((x >> 1) << 2) | ((y >> 1) << 3));
    if (size >= 32) nn = nn / 4 * 4;
    *on = nn;
    return true;
  }

  // finished_queueing (recode.cpp:851-937). put_or_get(key, &bit).
  template <class F>
  void nz_side_channel(const F& put_or_get) {
    int last = coding_type;
    coding_type = CT_NZ;
    MbState& meta = frames[cur].at(mb_x, mb_y);
    int bits[6];
    for (int i = 0; i < 6; i++) bits[i] = (meta.num_nonzeros[n] >> i) & 1;
    const int serialized = size > 16 ? 6 : size > 4 ? 4 : 2;
    int so_far = 0, left_nz = 0, above_nz = 0, x, y, nn;
    bool has_left = neighbor(false, &x, &y, &nn);
    if (has_left) left_nz = frames[cur].at(x, y).num_nonzeros[nn];
    bool has_above = neighbor(true, &x, &y, &nn);
    if (has_above) above_nz = frames[cur].at(x, y).num_nonzeros[nn];
    int prev_nz = frames[!cur].at(mb_x, mb_y).num_nonzeros[n];
    for (int i = 0; i < serialized; i++) {
      int bit = 1 << i;
      int lb = 2, ab = 2;
      if (has_left) lb = left_nz >= bit;
      if (above_nz) ab = above_nz >= bit;                            // sic: recode.cpp:888
      put_or_get(std::make_tuple(KEY_NZBIT0 + i, so_far + 64 * (prev_nz >= bit) + 128 * lb + 384 * ab,
                                 meta.is_8x8 + is_dc * 2 + c422 + cat * 4), &bits[i]);
      if (bits[i]) so_far |= bit;
    }
    meta.num_nonzeros[n] = 0;
    for (int i = 0; i < 6; i++) meta.num_nonzeros[n] |= bits[i] << i;
    coding_type = last;
  }
};

// ---------------------------------------------------------------------------------------
// Trace + drivers
// ---------------------------------------------------------------------------------------
struct Ev { char t; int a, b, c, d, e; };
struct Trace {
  std::vector<Ev> ev;
  std::vector<std::vector<uint8_t>> states;   // one per 'I'
  uint64_t bins = 0;
};
static int hexv(int c) { return c <= '9' ? c - '0' : (c | 32) - 'a' + 10; }
static std::vector<uint8_t> unhex(const char* s) {
  std::vector<uint8_t> v;
  while (s[0] && s[1] && s[0] != '\n') { v.push_back((uint8_t)(hexv(s[0]) * 16 + hexv(s[1]))); s += 2; }
  return v;
}
static bool load_trace(const char* fn, Trace* tr) {
  FILE* f = fopen(fn, "r");
  if (!f) return false;
  static char line[8192];
  while (fgets(line, sizeof line, f)) {
    Ev e = {line[0], 0, 0, 0, 0, 0};
    switch (line[0]) {
      case 'F': sscanf(line + 1, "%d %d %d", &e.a, &e.b, &e.c); break;
      case 'S': { long off = -1; sscanf(line + 1, "%d %ld", &e.a, &off); e.b = off >= 0; } break;
      case 'I': e.a = (int)tr->states.size(); tr->states.push_back(unhex(line + 2)); break;
      case 'M': sscanf(line + 1, "%d %d", &e.a, &e.b); break;
      case 'B': sscanf(line + 1, "%d %d %d %d %d", &e.a, &e.b, &e.c, &e.d, &e.e); break;
      case 'c': sscanf(line + 1, "%d %d", &e.a, &e.b); tr->bins++; break;
      case 'b': case 't': sscanf(line + 1, "%d", &e.a); tr->bins++; break;
      case '[': case ']': break;
      default: continue;
    }
    tr->ev.push_back(e);
  }
  fclose(f);
  return true;
}

struct QSym { int symbol, kind, ctx; };

// compressor::cabac_decoder + h264_symbol::execute (recode.cpp:1075-1103, 1167-1262)
static void encode_trace(const Trace& tr, std::vector<std::vector<uint8_t>>* payloads) {
  Model m;
  std::vector<uint8_t> out;
  RangeEncoder64* enc = nullptr;
  bool coded = false, queueing = false;
  std::vector<QSym> q;
  auto execute = [&](const QSym& s) {
    if (m.coding_type != CT_EOB) enc->put(s.symbol, Model::p1(m.E(m.key(s.kind, s.ctx)), enc->range));
    m.update(s.symbol, m.key(s.kind, s.ctx));
    if (s.kind == KEY_TERMINATE && s.symbol) { enc->finish(); payloads->push_back(out); }
  };
  auto on_bin = [&](int symbol, int kind, int ctx) {
    QSym s = {symbol, kind, ctx};
    if (queueing || !q.empty()) { q.push_back(s); m.track(symbol); } else execute(s);
  };
  for (const Ev& e : tr.ev) {
    switch (e.t) {
      case 'F': m.frame_observed(e.a, e.b, e.c); break;
      case 'S':
        delete enc; enc = nullptr; out.clear(); coded = e.b != 0;
        if (coded) enc = new RangeEncoder64(&out);
        break;
      case 'M': m.mb_x = e.a; m.mb_y = e.b; break;
      case 'B': m.cat = e.a; m.n = e.b; m.size = e.c; m.is_dc = e.d; m.c422 = e.e; break;
      case '[': if (coded) { m.begin_sigmap(); queueing = true; } break;
      case ']':
        if (!coded) break;
        m.end_sigmap();
        queueing = false;
        m.nz_side_channel([&](std::tuple<int, int, int> k, int* bit) {
          enc->put(*bit, Model::p1(m.E(k), enc->range));
          m.update(*bit, k);
        });
        m.zz = 0; m.observed = 0; m.coding_type = CT_SIGMAP;       // reset_mb_significance_state_tracking
        { std::vector<QSym> qq; qq.swap(q); for (const QSym& s : qq) execute(s); }
        m.coding_type = CT_UNKNOWN;
        break;
      case 'c': if (coded) on_bin(e.b, KEY_CTX, e.a); break;
      case 'b': if (coded) on_bin(e.a, KEY_BYPASS, 0); break;
      case 't': if (coded) on_bin(e.a, KEY_TERMINATE, 0); break;
      default: break;
    }
  }
  delete enc;
}

// decompressor::cabac_decoder (recode.cpp:1418-1527) driven by the trace's context sequence.
// ref_quirks=false (default): exact inverse of the compressor -- the count kept after a block
// is what the compressor keeps (count & 63, recode.cpp:928-931).  ref_quirks=true: statement
// faithful to the reference decompressor (keeps the recount of recode.cpp:942-954), which is
// not an inverse for completely full 8x8 blocks.
static bool decode_trace(const Trace& tr, const std::vector<std::vector<uint8_t>>& payloads, bool ref_quirks,
                         std::vector<std::vector<uint8_t>>* cabac_out, uint64_t* mismatches) {
  Model m;
  RangeDecoder64* dec = nullptr;
  std::vector<uint8_t> cab;
  CabacEncoderRef* ce = nullptr;
  uint8_t st[1024];
  bool coded = false;
  size_t slice = 0;
  *mismatches = 0;
  auto get = [&](int kind, int ctx, int expect) {
    int symbol;
    auto k = m.key(kind, ctx);
    if (kind == KEY_CTX && m.coding_type == CT_EOB) symbol = std::get<1>(k);
    else symbol = dec->get(Model::p1(m.E(k), dec->range));
    if (kind == KEY_CTX) ce->put(symbol, &st[ctx]);
    else if (kind == KEY_BYPASS) ce->put_bypass(symbol);
    m.update(symbol, k);
    if (kind == KEY_TERMINATE) {
      ce->put_terminate(symbol);
      if (symbol) {                                                 // finish(), recode.cpp:1508-1515
        if (!cab.empty() && cab.back() == 0x80) cab.pop_back();
        cabac_out->push_back(cab);
      }
    }
    if (symbol != expect) ++*mismatches;
  };
  for (const Ev& e : tr.ev) {
    if (*mismatches) break;   // contexts after a wrong bin are meaningless
    switch (e.t) {
      case 'F': m.frame_observed(e.a, e.b, e.c); break;
      case 'S':
        delete dec; delete ce; dec = nullptr; ce = nullptr; cab.clear(); coded = e.b != 0;
        if (coded) {
          if (slice >= payloads.size()) return false;
          const std::vector<uint8_t>& p = payloads[slice++];
          dec = new RangeDecoder64(p.data(), p.data() + p.size());
          ce = new CabacEncoderRef(&cab);
        }
        break;
      case 'I': memcpy(st, tr.states[e.a].data(), 1024); break;
      case 'M': m.mb_x = e.a; m.mb_y = e.b; break;
      case 'B': m.cat = e.a; m.n = e.b; m.size = e.c; m.is_dc = e.d; m.c422 = e.e; break;
      case '[':
        if (!coded) break;
        m.begin_sigmap();
        // the compressor updates is_8x8 (end_coding_type, :951) BEFORE it codes the nnz bits
        // (:1215-1219); the reference decompressor reads it before (:1485-1487) -- its 8x8 bug
        if (!ref_quirks && m.size > 32) m.frames[m.cur].at(m.mb_x, m.mb_y).is_8x8 = 1;
        m.nz_side_channel([&](std::tuple<int, int, int> k, int* bit) {
          *bit = dec->get(Model::p1(m.E(k), dec->range));
          m.update(*bit, k);
        });
        if (!ref_quirks) {
          uint8_t& v = m.frames[m.cur].at(m.mb_x, m.mb_y).num_nonzeros[m.n];
          if (v == 0 && m.size <= 16) v = (uint8_t)m.size;           // only a full block codes as 0
        }
        break;
      case ']':
        if (!coded) break;
        m.end_sigmap();
        if (!ref_quirks) m.frames[m.cur].at(m.mb_x, m.mb_y).num_nonzeros[m.n] &= 63;
        break;
      case 'c': if (coded) get(KEY_CTX, e.a, e.b); break;
      case 'b': if (coded) get(KEY_BYPASS, 0, e.a); break;
      case 't': if (coded) get(KEY_TERMINATE, 0, e.a); break;
      default: break;
    }
  }
  delete dec; delete ce;
  return *mismatches == 0;
}

static void print_hex(const char* tag, const std::vector<uint8_t>& v) {
  fputs(tag, stdout);
  for (uint8_t b : v) printf("%02x", b);
  fputc('\n', stdout);
}
static bool load_payloads(const char* fn, std::vector<std::vector<uint8_t>>* p) {
  FILE* f = fopen(fn, "r");
  if (!f) return false;
  std::string line; int c;
  while ((c = fgetc(f)) != EOF) {
    if (c != '\n') { line.push_back((char)c); continue; }
    if (line.size() >= 2 && line[0] == 'P') p->push_back(unhex(line.c_str() + 2));
    line.clear();
  }
  fclose(f);
  return true;
}

int main(int argc, char** argv) {
  if (argc >= 7 && std::string(argv[1]) == "streams") {
    // avr_oracle streams <digit_bits> <model> <nstreams> <nbins> <bits file> [quiet]: "A <hex>" per stream
    const int digit_bits = atoi(argv[2]), model = atoi(argv[3]);
    const long nstreams = atol(argv[4]), nbins = atol(argv[5]);
    const bool quiet = argc >= 8;
    std::vector<uint8_t> bits((size_t)(nstreams * nbins));
    FILE* f = fopen(argv[6], "rb");
    if (!f || fread(bits.data(), 1, bits.size(), f) != bits.size()) { fprintf(stderr, "cannot read %s\n", argv[6]); return 2; }
    fclose(f);
    auto t0 = std::chrono::steady_clock::now();
    uint64_t bytes = 0;
    for (long s = 0; s < nstreams; s++) {
      std::vector<uint8_t> out;
      const bool ok = digit_bits == 16 ? stream_roundtrip<uint16_t>(model, bits.data() + s * nbins, nbins, &out)
                                       : stream_roundtrip<uint8_t>(model, bits.data() + s * nbins, nbins, &out);
      if (!ok) { fprintf(stderr, "stream %ld does not round-trip\n", s); return 1; }
      bytes += out.size();
      if (!quiet) print_hex("A ", out);
    }
    double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (quiet) printf("{\"bins\": %lld, \"bytes\": %llu, \"roundtrip_s\": %.6f, \"mbins_s\": %.3f}\n", (long long)(nstreams * nbins),
                      (unsigned long long)bytes, sec, nstreams * nbins / sec / 1e6);
    return 0;
  }
  if (argc < 3) { fprintf(stderr, "usage: %s encode <trace> | decode <trace> <payloads> [--ref-quirks] | bench <trace> <reps>\n", argv[0]); return 2; }
  std::string cmd = argv[1];
  Trace tr;
  if (!load_trace(argv[2], &tr)) { fprintf(stderr, "cannot read %s\n", argv[2]); return 2; }
  if (cmd == "encode") {
    std::vector<std::vector<uint8_t>> payloads;
    encode_trace(tr, &payloads);
    for (auto& p : payloads) print_hex("P ", p);
    return 0;
  }
  if (cmd == "decode" && argc >= 4) {
    std::vector<std::vector<uint8_t>> payloads, cab;
    if (!load_payloads(argv[3], &payloads)) { fprintf(stderr, "cannot read %s\n", argv[3]); return 2; }
    bool quirks = argc >= 5 && std::string(argv[4]) == "--ref-quirks";
    uint64_t mism = 0;
    bool ok = decode_trace(tr, payloads, quirks, &cab, &mism);
    for (auto& c : cab) print_hex("C ", c);
    if (ok) printf("OK bins=%llu\n", (unsigned long long)tr.bins);
    else printf("MISMATCH after %zu slices\n", cab.size());
    return ok ? 0 : 1;
  }
  if (cmd == "bench" && argc >= 4) {
    int reps = atoi(argv[3]);
    std::vector<std::vector<uint8_t>> payloads, cab;
    auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < reps; r++) { payloads.clear(); encode_trace(tr, &payloads); }
    auto t1 = std::chrono::steady_clock::now();
    uint64_t mism = 0;
    for (int r = 0; r < reps; r++) { cab.clear(); decode_trace(tr, payloads, false, &cab, &mism); }
    auto t2 = std::chrono::steady_clock::now();
    double es = std::chrono::duration<double>(t1 - t0).count(), ds = std::chrono::duration<double>(t2 - t1).count();
    printf("{\"bins\": %llu, \"reps\": %d, \"encode_s\": %.6f, \"decode_s\": %.6f, \"encode_mbins_s\": %.3f, \"decode_mbins_s\": %.3f, \"mismatches\": %llu}\n",
           (unsigned long long)tr.bins, reps, es, ds, tr.bins * reps / es / 1e6, tr.bins * reps / ds / 1e6, (unsigned long long)mism);
    return mism ? 1 : 0;
  }
  fprintf(stderr, "bad command\n");
  return 2;
}
