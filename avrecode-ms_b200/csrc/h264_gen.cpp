// Synthetic CABAC H.264 clip generator -- see h264_gen.h.
#include "h264_gen.h"
#include "cabac_std.h"
#include "h264_syntax.h"

namespace avr {
namespace {

struct Rng {
  uint64_t s;
  explicit Rng(uint64_t seed) : s(seed * 0x9E3779B97F4A7C15ull + 0x1234567ull) {}
  uint64_t next() {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  int chance(int p256) { return (int)(next() & 255) < p256; }
  int below(int n) { return (int)(next() % (uint64_t)n); }
};

struct BitWriter {
  std::vector<uint8_t> bytes;
  uint32_t cur = 0; int n = 0;
  void u(int nbits, uint32_t v) {
    for (int i = nbits - 1; i >= 0; i--) {
      cur = (cur << 1) | ((v >> i) & 1);
      if (++n == 8) { bytes.push_back((uint8_t)cur); cur = 0; n = 0; }
    }
  }
  void ue(uint32_t v) {
    uint32_t x = v + 1; int len = 0;
    while ((x >> len) > 1) len++;
    u(len, 0);
    u(len + 1, x);
  }
  void se(int32_t v) { ue(v > 0 ? (uint32_t)(2 * v - 1) : (uint32_t)(-2 * v)); }
  bool aligned() const { return n == 0; }
  void trailing() { u(1, 1); while (n) u(1, 0); }
};

void append_nal(std::vector<uint8_t>* out, int long_sc, int nal_ref_idc, int nal_type, const std::vector<uint8_t>& rbsp) {
  if (long_sc) out->push_back(0);
  out->push_back(0); out->push_back(0); out->push_back(1);
  out->push_back((uint8_t)((nal_ref_idc << 5) | nal_type));
  int zeros = 0;
  for (uint8_t b : rbsp) {
    if (zeros >= 2 && b <= 3) { out->push_back(3); zeros = 0; }
    out->push_back(b);
    zeros = b == 0 ? zeros + 1 : 0;
  }
}

struct GenIO {
  static constexpr bool kFastResidual = false;
  int residual_block(int, int, int, int, int, int, int) { return 0; }
  const GenParams& gp;
  Rng& rng;
  CabacEncoder enc;
  uint8_t st[1024];
  int forced[40]; int nforced = 0, fpos = 0;
  int mbs_left = 0;
  int run = 0;  // current bypass prefix run length

  GenIO(const GenParams& g, Rng& r) : gp(g), rng(r) {}

  static int clampp(int p) { return p < 4 ? 4 : p > 252 ? 252 : p; }
  int prob(int ctx) const {
    int j = ((ctx * 37) & 63) - 32;
    if ((ctx >= 11 && ctx <= 13) || (ctx >= 24 && ctx <= 26)) return gp.p_skip;
    if (ctx >= 27 && ctx <= 29) return 256 - gp.p_direct;     // first mb_type bin of a B slice: 0 = B_Direct_16x16
    if (ctx == 36) return 256 - gp.p_direct;                  // first sub_mb_type bin: 0 = B_Direct_8x8
    if (ctx == 14) return gp.p_intra_in_p;
    if (ctx >= 3 && ctx <= 5) return gp.p_i16;
    if (ctx == 17) return gp.p_i16;
    if ((ctx >= 40 && ctx <= 42) || (ctx >= 47 && ctx <= 49)) return gp.p_mvd_nz;
    if (ctx >= 40 && ctx <= 53) return gp.p_mvd_more;
    if (ctx >= 73 && ctx <= 84) return gp.p_cbp;
    if (ctx >= 85 && ctx <= 104) return clampp(gp.p_cbf + j / 2);
    if (ctx >= 105 && ctx <= 165) return clampp(gp.p_sig + j);
    if (ctx >= 166 && ctx <= 226) return clampp(gp.p_last + j / 2);
    if (ctx >= 227 && ctx <= 275) {
      static const int base[5] = {227, 237, 247, 257, 266};
      int inc = 0;
      for (int c = 4; c >= 0; c--) if (ctx >= base[c]) { inc = ctx - base[c]; break; }
      return inc < 5 ? gp.p_gt1 : gp.p_gt1_more;
    }
    if (ctx >= 399 && ctx <= 401) return gp.p_t8x8;
    if (ctx >= 402 && ctx <= 416) return clampp(gp.p_sig * 3 / 4 + j);
    if (ctx >= 417 && ctx <= 425) return clampp(gp.p_last / 3 + 2);
    if (ctx >= 426 && ctx <= 435) return (ctx - 426) < 5 ? gp.p_gt1 : gp.p_gt1_more;
    return 128;
  }

  int get(int ctx) {
    int b = fpos < nforced ? forced[fpos++] : rng.chance(prob(ctx));
    enc.encode(b, &st[ctx]);
    return b;
  }
  int bypass(int kind) {
    int b;
    if (kind == BP_ESC_PREFIX) { b = run < 8 ? rng.chance(110) : 0; run = b ? run + 1 : 0; }
    else if (kind == BP_MVD_PREFIX) { b = run < 3 ? rng.chance(100) : 0; run = b ? run + 1 : 0; }
    else b = rng.chance(128);
    enc.bypass(b);
    return b;
  }
  int terminate(int kind) {
    int b = 0;
    if (kind == TERM_EOS) b = (--mbs_left == 0);
    enc.terminate(b);
    return b;
  }
  void mb_begin(int, int) {}
  void sub_mb_begin(int, int, int, int, int) {}
  void sub_mb_end() {}
  void sigmap_begin() {}
  void sigmap_end() {}
  void force(int b) { forced[nforced++] = b; }
  void hint(int kind, int arg) {
    nforced = 0; fpos = 0;
    switch (kind) {
      case HINT_I16_PRED: {  // value 0 VERT(top) 1 HOR(left) 2 DC 3 PLANE(both); arg bit0 left, bit1 top
        int cand[4], n = 0;
        cand[n++] = 2;
        if (arg & 2) cand[n++] = 0;
        if (arg & 1) cand[n++] = 1;
        if ((arg & 3) == 3) cand[n++] = 3;
        int v = cand[rng.below(n)];
        force(v >> 1); force(v & 1);
        break;
      }
      case HINT_CHROMA_PRED: {  // 0 DC 1 HOR(left) 2 VERT(top) 3 PLANE(both)
        int cand[4], n = 0;
        cand[n++] = 0;
        if (arg & 1) cand[n++] = 1;
        if (arg & 2) cand[n++] = 2;
        if ((arg & 3) == 3) cand[n++] = 3;
        int m = cand[rng.below(n)];
        if (m == 0) force(0);
        else if (m == 1) { force(1); force(0); }
        else if (m == 2) { force(1); force(1); force(0); }
        else { force(1); force(1); force(1); }
        break;
      }
      case HINT_I4_PRED:
        if (arg) force(1);  // an unavailable neighbour: predicted mode is DC, always legal
        break;
      case HINT_REF_IDX: {
        int r = rng.chance(150) ? 0 : rng.below(arg + 1);
        for (int i = 0; i < r; i++) force(1);
        force(0);
        break;
      }
      case HINT_DQP: {
        if (rng.chance(gp.p_dqp)) {
          int k = 1 + rng.below(6);
          for (int i = 0; i < k; i++) force(1);
          force(0);
        } else force(0);
        break;
      }
    }
  }
};

}  // namespace

std::vector<uint8_t> generate_clip(const GenParams& gp) {
  std::vector<uint8_t> out;
  Rng rng(gp.seed);
  const int log2_max_frame_num = 8;
  const int mbs = gp.mb_w * gp.mb_h;
  const int nslices = gp.slices_per_frame < 1 ? 1 : (gp.slices_per_frame > mbs ? mbs : gp.slices_per_frame);

  {  // SPS
    BitWriter w;
    w.u(8, 100); w.u(8, 0); w.u(8, 51);
    w.ue(0);             // sps id
    w.ue(1);             // chroma_format_idc 4:2:0
    w.ue(0); w.ue(0);    // bit depths
    w.u(1, 0);           // qpprime
    w.u(1, 0);           // scaling matrix
    w.ue(log2_max_frame_num - 4);
    w.ue(2);             // pic_order_cnt_type 2
    w.ue(gp.num_ref < 1 ? 1 : gp.num_ref);
    w.u(1, 0);           // gaps
    w.ue(gp.mb_w - 1); w.ue(gp.mb_h - 1);
    w.u(1, 1);           // frame_mbs_only
    w.u(1, 1);           // direct_8x8_inference
    w.u(1, 0);           // cropping
    w.u(1, 0);           // vui
    w.trailing();
    append_nal(&out, 1, 3, 7, w.bytes);
  }
  {  // PPS
    BitWriter w;
    w.ue(0); w.ue(0);
    w.u(1, 1);           // CABAC
    w.u(1, 0);
    w.ue(0);             // slice groups
    w.ue(0); w.ue(0);    // default num_ref (1): slices override
    w.u(1, 0); w.u(2, 0);
    w.se(gp.qp - 26); w.se(0); w.se(0);
    w.u(1, 1);           // deblocking_filter_control_present
    w.u(1, 0);           // constrained_intra_pred
    w.u(1, 0);           // redundant_pic_cnt_present
    if (gp.transform_8x8) { w.u(1, 1); w.u(1, 0); w.se(0); }
    w.trailing();
    append_nal(&out, 1, 3, 8, w.bytes);
  }

  std::vector<MbInfo> row((size_t)gp.mb_w);
  for (auto& m : row) m.uid = 0;
  uint32_t uid = 0;
  int frame_num = 0, since_idr = 0, prev_was_ref = 1;
  for (int f = 0; f < gp.frames; f++) {
    const bool idr = f == 0 || (gp.idr_period > 0 && f % gp.idr_period == 0);
    bool is_ref = true;
    if (idr) { frame_num = 0; since_idr = 0; }
    else {
      if (prev_was_ref) frame_num = (frame_num + 1) % (1 << log2_max_frame_num);
      since_idr++;
      if (gp.repeat_frame_num && (f & 1)) is_ref = false;
    }
    int avail_refs = since_idr < gp.num_ref ? since_idr : gp.num_ref;
    if (avail_refs < 1) avail_refs = 1;
    for (int s = 0; s < nslices; s++) {
      const int first = (int)((int64_t)mbs * s / nslices), endmb = (int)((int64_t)mbs * (s + 1) / nslices);
      const bool islice = idr || rng.chance(12);
      const bool bslice = !islice && gp.p_bslice > 0 && rng.chance(gp.p_bslice);
      SliceParams sp;
      sp.slice_type = islice ? SLICE_I : bslice ? SLICE_B : SLICE_P;
      sp.first_mb = first; sp.mb_w = gp.mb_w; sp.mb_h = gp.mb_h;
      sp.num_ref[0] = islice ? 1 : avail_refs; sp.num_ref[1] = bslice ? avail_refs : 1;
      sp.cabac_init_idc = gp.cabac_init_idc;
      int qp_delta = rng.below(7) - 3;
      sp.slice_qp = gp.qp + qp_delta;
      sp.transform_8x8_mode = gp.transform_8x8;
      sp.chroma_format_idc = 1;
      sp.direct_8x8_inference = 1;

      BitWriter w;
      w.ue(first);
      w.ue((islice ? 2 : bslice ? 1 : 0) + (rng.chance(128) ? 5 : 0));
      w.ue(0);
      w.u(log2_max_frame_num, frame_num);
      if (idr) w.ue(f & 0xFFFF);  // idr_pic_id
      if (bslice) w.u(1, 1);                 // direct_spatial_mv_pred_flag
      if (!islice) {
        w.u(1, 1); w.ue(sp.num_ref[0] - 1);  // num_ref_idx_active_override
        if (bslice) w.ue(sp.num_ref[1] - 1);
        w.u(1, 0);                           // ref_pic_list_modification_flag_l0
        if (bslice) w.u(1, 0);               // ref_pic_list_modification_flag_l1
      }
      if (is_ref) {
        if (idr) { w.u(1, 0); w.u(1, 0); } else w.u(1, 0);
      }
      if (!islice) w.ue(sp.cabac_init_idc);
      w.se(qp_delta);
      w.ue(1);  // disable_deblocking_filter_idc = 1
      while (!w.aligned()) w.u(1, 1);  // cabac_alignment_one_bit

      GenIO io(gp, rng);
      std::vector<uint8_t> data;
      io.enc.init(&data);
      cabac_init_states(io.st, islice, sp.cabac_init_idc, sp.slice_qp);
      io.mbs_left = endmb - first;
      SliceParser<GenIO> parser(io, sp, row.data(), ++uid);
      int r = parser.run();
      (void)r;
      std::vector<uint8_t> rbsp = w.bytes;
      rbsp.insert(rbsp.end(), data.begin(), data.end());
      append_nal(&out, gp.long_startcodes && s == 0, is_ref ? (idr ? 3 : 1 + rng.below(3)) : 0, idr ? 5 : 1, rbsp);
    }
    prev_was_ref = is_ref;
  }
  return out;
}

}  // namespace avr
