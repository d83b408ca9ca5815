// H.264 CABAC slice_data() syntax state machine (ITU-T H.264 7.3.4/7.3.5, 9.3.2, 9.3.3.1),
// written once, host+device, templated on a bin I/O policy.
//
// What it replaces: in the reference this logic lives inside libavcodec
// (ffmpeg/libavcodec/h264_cabac.c:1288-1580 ctxIdxInc helpers, :1583-1780 residual blocks,
// :1926-2480 ff_h264_decode_mb_cabac, slice loop h264_slice.c:2357-2440) and calls UP through
// AVCodecHooks for every bin (coding_hooks.h:27-55).  Here the same walk decides which
// ctxIdx / bypass / terminate comes next and when residual blocks begin and end, and calls a
// policy object `IO` instead:
//     int  get(int ctxIdx)            -- CABACCodingHooks::get      (state = cabac_state + ctxIdx)
//     int  bypass(int kind)           -- CABACCodingHooks::get_bypass
//     int  terminate(int kind)        -- CABACCodingHooks::get_terminate
//     void mb_begin(x, y)             -- ModelHooks::mb_xy
//     void sub_mb_begin(cat,n,max_coeff,is_dc,c422) / sub_mb_end()  -- begin_sub_mb / end_sub_mb
//     void sigmap_begin() / sigmap_end()   -- begin/end_coding_type(PIP_SIGNIFICANCE_MAP)
//     static constexpr bool kFastResidual; int residual_block(cat,n,max_coeff,is_dc,sig,last,abs)
//                                     -- optional whole-block entry point (see residual_block below)
//     void hint(kind, arg)            -- no reference counterpart; lets the synthetic generator
//                                        keep random bins legal (no-op in the product policies)
// Policies: host CABAC decode + record (compress front end), device range decode + CABAC
// re-encode (decompress kernel), random source + CABAC encode (clip generator).
//
// Supported: I, P and B slices (round 2: B_Skip / B_Direct_16x16 / the 21 two-list partition types / B_8x8 with its 13 sub
// types; h264_cabac.c:1332-1370 skip contexts, :1973-2003 mb_type, :1455-1469 sub_mb_type, :1472-1501 ref_idx with the
// direct_cache mask, :2112-2330 ref/mvd order), frame pictures (no MBAFF/field), chroma_format_idc 0/1,
// transform_8x8_mode 0/1, multiple slices per picture, no I_PCM.  Direct prediction itself (motion vectors of direct blocks)
// is never needed: direct blocks carry no mvd / ref_idx bins and are masked out of their neighbours' contexts.
#pragma once
#include "avr_common.h"

namespace avr {

enum { SLICE_P = 0, SLICE_B = 1, SLICE_I = 2 };
enum { BP_SIGN = 0, BP_ESC_PREFIX, BP_ESC_SUFFIX, BP_MVD_PREFIX, BP_MVD_SUFFIX };
enum { TERM_PCM = 0, TERM_EOS = 1 };
enum { HINT_I16_PRED = 0, HINT_CHROMA_PRED, HINT_I4_PRED, HINT_REF_IDX, HINT_DQP };

enum { SYN_OK = 0, SYN_E_PCM = -10, SYN_E_DQP = -11, SYN_E_REF = -12, SYN_E_SLICE_TYPE = -13,
       SYN_E_MVD = -14, SYN_E_LENGTH = -15 /* set by the callers: the walk did not end where slice_data() ends */ };

struct SliceParams {
  int slice_type;        // SLICE_P / SLICE_I (slice_type % 5 mapped; SP/SI unsupported)
  int first_mb;          // first_mb_in_slice
  int mb_w, mb_h;        // picture size in macroblocks
  int num_ref[2];        // num_ref_idx_active
  int cabac_init_idc;
  int slice_qp;          // 26 + pic_init_qp + slice_qp_delta
  int transform_8x8_mode;
  int chroma_format_idc; // 0 or 1
  int direct_8x8_inference;
};

// Per-macroblock neighbour state kept for the current row (left = row[x-1], top = row[x]).
struct MbInfo {
  uint32_t uid;          // slice uid: availability == same slice (9.3.3.1.1, 6.4.x)
  uint8_t flags;         // MBF_*
  uint8_t chroma_pred;   // intra_chroma_pred_mode (0 for inter)
  uint16_t cbp;          // bits0-3 luma 8x8, 4-5 chroma, 6/7 Cb/Cr DC coded, 8 luma DC coded
  uint8_t nnz[24];       // 16 luma (y*4+x), 4 Cb (y*2+x), 4 Cr : total_coeff per 4x4
  uint8_t refpos;        // bit (list*4 + b8): ref_idx > 0
  uint8_t direct8;       // bit b8: that 8x8 is direct-predicted (B_Skip / B_Direct_16x16: all four; B_8x8: its direct sub-blocks)
  uint8_t pad[2];
  uint8_t mvd[2][16][2]; // [list][y*4+x][comp] = min(|mvd|, 70)
};
enum { MBF_SKIP = 1, MBF_INTRA = 2, MBF_I16 = 4, MBF_8x8DCT = 8, MBF_DIRECT = 16, MBF_PCM = 32 };

// HAS_B = false compiles the B-slice branches out: the device kernels instantiate both and pick per slice, so that I / P
// slices run the smaller code (the B branches cost the chain kernel 7 % on a P-only corpus: profiles/r02_summary.md).
template <class IO, bool HAS_B = true>
struct SliceParser {
  IO& io;
  const SliceParams& sp;
  MbInfo* row;       // [mb_w], caller-owned, persists across slices of a file
  uint32_t uid;      // unique per slice within the file
  int last_dqp_nz;
  MbInfo cur;
  const MbInfo* L;
  const MbInfo* T;
  int mbs_done;

  AVR_HDN SliceParser(IO& io_, const SliceParams& sp_, MbInfo* row_, uint32_t uid_)
      : io(io_), sp(sp_), row(row_), uid(uid_), last_dqp_nz(0), L(nullptr), T(nullptr), mbs_done(0) {}

  // ---- slice loop (h264_slice.c:2370-2440) ------------------------------------------------
  AVR_HDN int run() {
    if (sp.slice_type != SLICE_P && sp.slice_type != SLICE_I && !(HAS_B && sp.slice_type == SLICE_B)) return SYN_E_SLICE_TYPE;
    int x = sp.first_mb % sp.mb_w, y = sp.first_mb / sp.mb_w;
    last_dqp_nz = 0;
    for (;;) {
      io.mb_begin(x, y);
      int r = macroblock(x, y);
      if (r < 0) return r;
      mbs_done++;
      int eos = io.terminate(TERM_EOS);
      if (++x >= sp.mb_w) { x = 0; ++y; }
      if (eos || y >= sp.mb_h) return mbs_done;
    }
  }

 private:
  // ---- neighbour accessors ---------------------------------------------------------------
  AVR_HD int unavail_nnz() const { return (cur.flags & MBF_INTRA) ? 64 : 0; }
  AVR_HD int luma_nnz(int bx, int by) const {
    if (bx < 0) return L ? L->nnz[by * 4 + 3] : unavail_nnz();
    if (by < 0) return T ? T->nnz[12 + bx] : unavail_nnz();
    return cur.nnz[by * 4 + bx];
  }
  AVR_HD int chroma_nnz(int c, int cx, int cy) const {
    if (cx < 0) return L ? L->nnz[16 + 4 * c + cy * 2 + 1] : unavail_nnz();
    if (cy < 0) return T ? T->nnz[16 + 4 * c + 2 + cx] : unavail_nnz();
    return cur.nnz[16 + 4 * c + cy * 2 + cx];
  }
  AVR_HD int nb_cbp(const MbInfo* m) const { return m ? m->cbp : ((cur.flags & MBF_INTRA) ? 0x7CF : 0x00F); }
  AVR_HD int mvd_at(int list, int bx, int by, int comp) const {
    if (bx < 0) return L ? L->mvd[list][by * 4 + 3][comp] : 0;
    if (by < 0) return T ? T->mvd[list][12 + bx][comp] : 0;
    return cur.mvd[list][by * 4 + bx][comp];
  }
  AVR_HD int refpos_at(int list, int bx, int by) const {  // 4x4 coordinates
    if (bx < 0) return L ? (L->refpos >> (list * 4 + (by >> 1) * 2 + 1)) & 1 : 0;
    if (by < 0) return T ? (T->refpos >> (list * 4 + 2 + (bx >> 1))) & 1 : 0;
    return (cur.refpos >> (list * 4 + (by >> 1) * 2 + (bx >> 1))) & 1;
  }
  // sl->direct_cache of ffmpeg (h264_mvpred.h:723-753, h264_cabac.c:2131): is the 8x8 that holds 4x4 (bx,by) direct?
  AVR_HD int direct_at(int bx, int by) const {
    if (bx < 0) return L ? (L->direct8 >> ((by >> 1) * 2 + 1)) & 1 : 0;
    if (by < 0) return T ? (T->direct8 >> (2 + (bx >> 1))) & 1 : 0;
    return (cur.direct8 >> ((by >> 1) * 2 + (bx >> 1))) & 1;
  }
  // luma4x4BlkIdx -> (x,y) in the 4x4 grid (6.4.3)
  static AVR_HD int blk_x(int i) { return (i & 1) | ((i >> 1) & 2); }
  static AVR_HD int blk_y(int i) { return ((i >> 1) & 1) | ((i >> 2) & 2); }

  // ---- mb_type for intra macroblocks (h264_cabac.c:1301-1330) -----------------------------
  AVR_HDN int intra_mb_type(int base, int intra_slice) {
    if (intra_slice) {
      int ctx = 0;
      if (L && (L->flags & (MBF_I16 | MBF_PCM))) ctx++;
      if (T && (T->flags & (MBF_I16 | MBF_PCM))) ctx++;
      if (io.get(base + ctx) == 0) return 0;
      base += 2;
    } else {
      if (io.get(base) == 0) return 0;
    }
    if (io.terminate(TERM_PCM)) return 25;
    int t = 1;
    t += 12 * io.get(base + 1);
    if (io.get(base + 2)) t += 4 + 4 * io.get(base + 2 + intra_slice);
    io.hint(HINT_I16_PRED, (L ? 1 : 0) | (T ? 2 : 0));
    t += 2 * io.get(base + 3 + intra_slice);
    t += 1 * io.get(base + 3 + 2 * intra_slice);
    return t;
  }

  AVR_HDN void intra_pred_mode_bins() {  // h264_cabac.c:1372-1384
    if (io.get(68)) return;
    io.get(69); io.get(69); io.get(69);
  }

  AVR_HDN int b_sub_mb_type() {  // h264_cabac.c:1455-1469
    if (!io.get(36)) return 0;                       // B_Direct_8x8
    if (!io.get(37)) return 1 + io.get(39);          // B_L0_8x8, B_L1_8x8
    int type = 3;
    if (io.get(38)) {
      if (io.get(39)) return 11 + io.get(39);        // B_L1_4x4, B_Bi_4x4
      type += 4;
    }
    type += 2 * io.get(39);
    type += io.get(39);
    return type;
  }

  AVR_HDN int ref_idx(int list, int bx, int by) {  // h264_cabac.c:1472-1501; direct neighbours do not count (B slices)
    int ctx = 0;
    if (refpos_at(list, bx - 1, by) && !(HAS_B && direct_at(bx - 1, by))) ctx++;
    if (refpos_at(list, bx, by - 1) && !(HAS_B && direct_at(bx, by - 1))) ctx += 2;
    int ref = 0;
    io.hint(HINT_REF_IDX, sp.num_ref[list] - 1);
    while (io.get(54 + ctx)) {
      ref++;
      ctx = (ctx >> 2) + 4;
      if (ref >= 32) return SYN_E_REF;
    }
    return ref;
  }

  // one mvd component (h264_cabac.c:1503-1541); returns min(|mvd|,70) or <0
  AVR_HDN int mvd_comp(int ctxbase, int amvd) {
    if (!io.get(ctxbase + (amvd > 2) + (amvd > 32))) return 0;
    int mvd = 1;
    ctxbase += 3;
    while (mvd < 9 && io.get(ctxbase)) {
      if (mvd < 4) ctxbase++;
      mvd++;
    }
    if (mvd >= 9) {
      int k = 3;
      while (io.bypass(BP_MVD_PREFIX)) {
        mvd += 1 << k;
        k++;
        if (k > 24) return SYN_E_MVD;
      }
      while (k--) mvd += io.bypass(BP_MVD_SUFFIX) << k;
    }
    io.bypass(BP_SIGN);
    return mvd < 70 ? mvd : 70;
  }

  // mvd pair for the partition whose first 4x4 is (bx,by), size (w,h) in 4x4 units
  AVR_HDN int mvd_partition(int list, int bx, int by, int w, int h) {
    int a0 = mvd_at(list, bx - 1, by, 0) + mvd_at(list, bx, by - 1, 0);
    int a1 = mvd_at(list, bx - 1, by, 1) + mvd_at(list, bx, by - 1, 1);
    int mx = mvd_comp(40, a0);
    if (mx < 0) return mx;
    int my = mvd_comp(47, a1);
    if (my < 0) return my;
    for (int yy = by; yy < by + h; yy++)
      for (int xx = bx; xx < bx + w; xx++) {
        cur.mvd[list][yy * 4 + xx][0] = (uint8_t)mx;
        cur.mvd[list][yy * 4 + xx][1] = (uint8_t)my;
      }
    return 0;
  }

  AVR_HD void set_ref(int list, int b8, int ref) {
    int bit = list * 4 + b8;
    cur.refpos = (uint8_t)((cur.refpos & ~(1 << bit)) | ((ref > 0) << bit));
  }

  // ---- residual block (h264_cabac.c:1583-1780) --------------------------------------------
  // returns coeff_count
  AVR_HDN int residual_block(int cat, int n, int max_coeff, int is_dc) {
    const int sig_base[6] = {105, 120, 134, 149, 152, 402};
    const int last_base[6] = {166, 181, 195, 210, 213, 417};
    const int abs_base[6] = {227, 237, 247, 257, 266, 426};
    // a policy may take the whole block in one call (the decompress kernel keeps its decoder
    // registers live across the block's bins: avr_decode.h)
    if constexpr (IO::kFastResidual) return io.residual_block(cat, n, max_coeff, is_dc, sig_base[cat], last_base[cat], abs_base[cat]);
    io.sub_mb_begin(cat, n, max_coeff, is_dc, 0);
    io.sigmap_begin();
    int coeff_count = 0;
    int last;
    const int coefs = max_coeff - 1;
    if (max_coeff == 64) {
      for (last = 0; last < coefs; last++) {
        if (io.get(sig_base[cat] + sig8x8_inc(last))) {
          coeff_count++;
          if (io.get(last_base[cat] + last8x8_inc(last))) { last = max_coeff; break; }
        }
      }
    } else {
      for (last = 0; last < coefs; last++) {
        if (io.get(sig_base[cat] + last)) {
          coeff_count++;
          if (io.get(last_base[cat] + last)) { last = max_coeff; break; }
        }
      }
    }
    if (last == max_coeff - 1) coeff_count++;
    io.sigmap_end();

    // coeff_abs_level_minus1 + sign, reverse scan order (h264_cabac.c:1727-1768)
    int node_ctx = 0;
    for (int c = coeff_count; c > 0; c--) {
      // node ctx 0..3: no level>1 seen yet; 4..7: numDecodAbsLevelGt1 + 3
      int ctx1 = abs_base[cat] + (node_ctx < 4 ? node_ctx + 1 : 0);
      if (io.get(ctx1) == 0) {
        node_ctx = node_ctx < 3 ? node_ctx + 1 : (node_ctx == 3 ? 3 : node_ctx);
        io.bypass(BP_SIGN);
      } else {
        int gt1 = node_ctx < 4 ? 5 : (node_ctx + 2 > (cat == 3 ? 8 : 9) ? (cat == 3 ? 8 : 9) : node_ctx + 2);
        int ctx2 = abs_base[cat] + gt1;
        node_ctx = node_ctx < 4 ? 4 : (node_ctx < 7 ? node_ctx + 1 : 7);
        int coeff_abs = 2;
        while (coeff_abs < 15 && io.get(ctx2)) coeff_abs++;
        if (coeff_abs >= 15) {
          int j = 0;
          while (io.bypass(BP_ESC_PREFIX) && j < 30) j++;
          while (j--) io.bypass(BP_ESC_SUFFIX);
        }
        io.bypass(BP_SIGN);
      }
    }
    io.sub_mb_end();
    return coeff_count;
  }

  static AVR_HD int sig8x8_inc(int i) {
    // Table 9-43, frame coded: significant_coeff_flag ctxIdxInc for 8x8 blocks, 4 bits each
    const int q = i >> 4;   // selects, not an indexed stack array
    const uint64_t t = q == 0 ? 0x5544433445543210ull : q == 1 ? 0x789A987776334444ull : q == 2 ? 0xB689AE9876BDCB67ull : 0x0CAEBDCB9AE96BDCull;
    return (int)((t >> ((i & 15) * 4)) & 15);
  }
  static AVR_HD int last8x8_inc(int i) {
    // Table 9-43: last_significant_coeff_flag ctxIdxInc for 8x8 blocks
    return i == 0 ? 0 : i < 16 ? 1 : i < 32 ? 2 : i < 40 ? 3 : i < 48 ? 4 : i < 52 ? 5 : i < 56 ? 6 : i < 60 ? 7 : 8;
  }

  // coded_block_flag ctxIdx (h264_cabac.c:1555-1590)
  AVR_HD int cbf_ctx_dc(int cat, int idx /*0 luma, 0/1 chroma*/) const {
    const int base = cat == 0 ? 85 : 97;
    int nza, nzb;
    if (cat == 3) {
      nza = (nb_cbp(L) >> (6 + idx)) & 1;
      nzb = (nb_cbp(T) >> (6 + idx)) & 1;
    } else {
      nza = nb_cbp(L) & 0x100;
      nzb = nb_cbp(T) & 0x100;
    }
    return base + (nza > 0) + 2 * (nzb > 0);
  }

  AVR_HDN void luma_block4(int cat, int i, int max_coeff) {
    int bx = blk_x(i), by = blk_y(i);
    int ctx = (cat == 1 ? 89 : 93) + (luma_nnz(bx - 1, by) > 0) + 2 * (luma_nnz(bx, by - 1) > 0);
    if (io.get(ctx) == 0) { cur.nnz[by * 4 + bx] = 0; return; }
    cur.nnz[by * 4 + bx] = (uint8_t)residual_block(cat, i, max_coeff, 0);
  }

  AVR_HDN void residuals(int cbp) {
    if (cur.flags & MBF_I16) {
      if (io.get(cbf_ctx_dc(0, 0))) {
        residual_block(0, 48, 16, 1);
        cur.cbp |= 0x100;
      }
      if (cbp & 15)
        for (int i = 0; i < 16; i++) luma_block4(1, i, 15);
    } else {
      for (int i8 = 0; i8 < 4; i8++) {
        if (!(cbp & (1 << i8))) continue;
        if (cur.flags & MBF_8x8DCT) {
          int cc = residual_block(5, 4 * i8, 64, 0);  // no coded_block_flag outside 4:4:4
          for (int k = 0; k < 4; k++) { int i = 4 * i8 + k; cur.nnz[blk_y(i) * 4 + blk_x(i)] = (uint8_t)cc; }
        } else {
          for (int k = 0; k < 4; k++) luma_block4(2, 4 * i8 + k, 16);
        }
      }
    }
    if (sp.chroma_format_idc == 1) {
      if (cbp & 0x30) {
        for (int c = 0; c < 2; c++) {
          if (io.get(cbf_ctx_dc(3, c))) {
            residual_block(3, 49 + c, 4, 1);
            cur.cbp |= (uint16_t)(0x40 << c);
          }
        }
      }
      if (cbp & 0x20) {
        for (int c = 0; c < 2; c++)
          for (int i = 0; i < 4; i++) {
            int cx = i & 1, cy = i >> 1;
            int ctx = 101 + (chroma_nnz(c, cx - 1, cy) > 0) + 2 * (chroma_nnz(c, cx, cy - 1) > 0);
            if (io.get(ctx) == 0) { cur.nnz[16 + 4 * c + i] = 0; continue; }
            cur.nnz[16 + 4 * c + i] = (uint8_t)residual_block(4, 16 + 16 * c + i, 15, 0);
          }
      }
    }
  }

  // ---- one macroblock (h264_cabac.c:1926-2480) --------------------------------------------
  AVR_HDN int macroblock(int x, int y) {
    L = (x > 0 && row[x - 1].uid == uid) ? &row[x - 1] : nullptr;
    T = (y > 0 && row[x].uid == uid) ? &row[x] : nullptr;
    // T aliases row[x], which is overwritten at the end of this function only.
    cur.uid = uid; cur.flags = 0; cur.chroma_pred = 0; cur.cbp = 0; cur.refpos = 0; cur.direct8 = 0;
    for (int i = 0; i < 24; i++) cur.nnz[i] = 0;
    for (int l = 0; l < 2; l++)
      for (int i = 0; i < 16; i++) cur.mvd[l][i][0] = cur.mvd[l][i][1] = 0;

    const int decode_chroma = sp.chroma_format_idc == 1;
    int mbt_intra = -1;   // >=0: intra mb_type 0..25
    int p_type = -1;      // P: 0 16x16, 1 16x8, 2 8x16, 3 8x8
    int b_type = -1;      // B: 0 B_Direct_16x16, 1..21 partition types (Table 7-14), 22 B_8x8
    const int is_b = HAS_B && sp.slice_type == SLICE_B;
    if (sp.slice_type != SLICE_I) {
      int ctx = 0;
      if (L && !(L->flags & MBF_SKIP)) ctx++;
      if (T && !(T->flags & MBF_SKIP)) ctx++;
      if (io.get((is_b ? 24 : 11) + ctx)) {  // mb_skip_flag (h264_cabac.c:1332-1370: +13 in B slices)
        cur.flags = is_b ? (MBF_SKIP | MBF_DIRECT) : MBF_SKIP;   // B_Skip is direct-predicted
        cur.direct8 = is_b ? 15 : 0;
        last_dqp_nz = 0;
        row[x] = cur;
        return 0;
      }
      if (is_b) {                            // h264_cabac.c:1973-2003
        int c = 0;
        if (L && !(L->flags & MBF_DIRECT)) c++;
        if (T && !(T->flags & MBF_DIRECT)) c++;
        if (!io.get(27 + c)) b_type = 0;
        else if (!io.get(27 + 3)) b_type = 1 + io.get(27 + 5);
        else {
          int bits = io.get(27 + 4) << 3;
          bits |= io.get(27 + 5) << 2;
          bits |= io.get(27 + 5) << 1;
          bits |= io.get(27 + 5);
          if (bits < 8) b_type = bits + 3;
          else if (bits == 13) mbt_intra = intra_mb_type(32, 0);
          else if (bits == 14) b_type = 11;
          else if (bits == 15) b_type = 22;
          else { bits = (bits << 1) | io.get(27 + 5); b_type = bits - 4; }
        }
      } else if (io.get(14) == 0) {
        if (io.get(15) == 0) p_type = 3 * io.get(16);
        else p_type = 2 - io.get(17);
      } else {
        mbt_intra = intra_mb_type(17, 0);
      }
    } else {
      mbt_intra = intra_mb_type(3, 1);
    }

    int cbp = 0;
    int dct8x8_allowed = sp.transform_8x8_mode;
    const int nts = ((T && (T->flags & MBF_8x8DCT)) ? 1 : 0) + ((L && (L->flags & MBF_8x8DCT)) ? 1 : 0);

    if (mbt_intra >= 0) {
      if (mbt_intra == 25) return SYN_E_PCM;
      cur.flags |= MBF_INTRA;
      if (mbt_intra == 0) {  // I_NxN
        if (dct8x8_allowed && io.get(399 + nts)) {
          cur.flags |= MBF_8x8DCT;
          for (int i = 0; i < 4; i++) {
            int lft = (i & 1) == 0 && !L, top = (i >> 1) == 0 && !T;
            io.hint(HINT_I4_PRED, lft | top);
            intra_pred_mode_bins();
          }
        } else {
          for (int i = 0; i < 16; i++) {
            int lft = blk_x(i) == 0 && !L, top = blk_y(i) == 0 && !T;
            io.hint(HINT_I4_PRED, lft | top);
            intra_pred_mode_bins();
          }
        }
      } else {
        cur.flags |= MBF_I16;
        int k = mbt_intra - 1;  // 0..23
        cbp = (k >= 12 ? 15 : 0) | (((k >> 2) % 3) << 4);
      }
      if (decode_chroma) {
        io.hint(HINT_CHROMA_PRED, (L ? 1 : 0) | (T ? 2 : 0));
        int ctx = 0;
        if (L && L->chroma_pred != 0) ctx++;
        if (T && T->chroma_pred != 0) ctx++;
        int mode = 0;
        if (io.get(64 + ctx)) {
          mode = 1;
          if (io.get(64 + 3)) { mode = 2; if (io.get(64 + 3)) mode = 3; }
        }
        cur.chroma_pred = (uint8_t)mode;
      }
    } else if (HAS_B && b_type == 0) {  // B_Direct_16x16: no ref_idx, no mvd (h264_cabac.c:2218-2222)
      cur.flags |= MBF_DIRECT;
      cur.direct8 = 15;
      dct8x8_allowed &= sp.direct_8x8_inference;
    } else if (HAS_B && b_type == 22) {  // B_8x8 (h264_cabac.c:2112-2217)
      int shape[4], lists[4];   // shape 0 8x8, 1 8x4, 2 4x8, 3 4x4 (as P sub types); lists bit0 L0, bit1 L1; 0 = direct
      for (int i = 0; i < 4; i++) {
        const int st = b_sub_mb_type();
        // b_sub_mb_type_info (h264data.h:154-168), 2 bits per entry
        shape[i] = (int)((0x3F99900u >> (2 * st)) & 3);
        lists[i] = (int)((0x39FA5E4u >> (2 * st)) & 3);
        if (st == 0) cur.direct8 |= (uint8_t)(1 << i);
      }
      for (int list = 0; list < 2; list++)
        for (int i = 0; i < 4; i++) {
          if (!(lists[i] & (1 << list))) continue;          // direct sub-blocks have lists == 0
          int ref = 0;
          if (sp.num_ref[list] > 1) {
            ref = ref_idx(list, (i & 1) * 2, (i >> 1) * 2);
            if (ref < 0) return ref;
            if (ref >= sp.num_ref[list]) return SYN_E_REF;
          }
          set_ref(list, i, ref);
        }
      for (int i = 0; i < 4; i++)                            // get_dct8x8_allowed (h264.h)
        if (lists[i] == 0 ? !sp.direct_8x8_inference : shape[i] != 0) dct8x8_allowed = 0;
      for (int list = 0; list < 2; list++)
        for (int i = 0; i < 4; i++) {
          if (!(lists[i] & (1 << list))) continue;
          int bx = (i & 1) * 2, by = (i >> 1) * 2, r = 0;
          switch (shape[i]) {
            case 0: r = mvd_partition(list, bx, by, 2, 2); break;
            case 1: r = mvd_partition(list, bx, by, 2, 1); if (r >= 0) r = mvd_partition(list, bx, by + 1, 2, 1); break;
            case 2: r = mvd_partition(list, bx, by, 1, 2); if (r >= 0) r = mvd_partition(list, bx + 1, by, 1, 2); break;
            default:
              for (int j = 0; j < 4 && r >= 0; j++) r = mvd_partition(list, bx + (j & 1), by + (j >> 1), 1, 1);
          }
          if (r < 0) return r;
        }
    } else if (HAS_B && b_type > 0) {  // B 16x16 / 16x8 / 8x16 with one or two lists per partition (h264_cabac.c:2223-2330)
      // b_mb_type_info (h264data.h:128-152): types 1..3 = 16x16 L0 / L1 / Bi; 4..21 alternate 16x8 / 8x16 over nine list pairs
      const int shp = b_type < 4 ? 0 : ((b_type & 1) ? 2 : 1);       // 0 16x16, 1 16x8, 2 8x16
      const int pair = b_type < 4 ? 0 : (b_type - 4) >> 1;
      const int use0 = b_type < 4 ? b_type : (int)((0x3F999u >> (2 * pair)) & 3);     // lists of partition 0: 1 L0, 2 L1, 3 both
      const int use1 = b_type < 4 ? 0 : (int)((0x39F69u >> (2 * pair)) & 3);          // lists of partition 1
      const int nparts = shp == 0 ? 1 : 2;
      for (int list = 0; list < 2; list++)
        for (int i = 0; i < nparts; i++) {
          if (!((i ? use1 : use0) & (1 << list))) continue;
          int bx = shp == 2 ? 2 * i : 0, by = shp == 1 ? 2 * i : 0;
          int ref = 0;
          if (sp.num_ref[list] > 1) {
            ref = ref_idx(list, bx, by);
            if (ref < 0) return ref;
            if (ref >= sp.num_ref[list]) return SYN_E_REF;
          }
          if (shp == 0) { for (int b = 0; b < 4; b++) set_ref(list, b, ref); }
          else if (shp == 1) { set_ref(list, 2 * i, ref); set_ref(list, 2 * i + 1, ref); }
          else { set_ref(list, i, ref); set_ref(list, i + 2, ref); }
        }
      for (int list = 0; list < 2; list++)
        for (int i = 0; i < nparts; i++) {
          if (!((i ? use1 : use0) & (1 << list))) continue;
          int bx = shp == 2 ? 2 * i : 0, by = shp == 1 ? 2 * i : 0;
          int w = shp == 2 ? 2 : 4, h = shp == 1 ? 2 : 4;
          int r = mvd_partition(list, bx, by, w, h);
          if (r < 0) return r;
        }
    } else if (p_type == 3) {  // P_8x8
      int sub[4];
      for (int i = 0; i < 4; i++) {  // h264_cabac.c:1432-1441
        if (io.get(21)) sub[i] = 0;
        else if (!io.get(22)) sub[i] = 1;
        else if (io.get(23)) sub[i] = 2;
        else sub[i] = 3;
      }
      for (int i = 0; i < 4; i++) {
        int ref = 0;
        if (sp.num_ref[0] > 1) {
          ref = ref_idx(0, (i & 1) * 2, (i >> 1) * 2);
          if (ref < 0) return ref;
          if (ref >= sp.num_ref[0]) return SYN_E_REF;
        }
        set_ref(0, i, ref);
      }
      for (int i = 0; i < 4; i++) if (sub[i] != 0) dct8x8_allowed = 0;
      for (int i = 0; i < 4; i++) {
        int bx = (i & 1) * 2, by = (i >> 1) * 2, r = 0;
        switch (sub[i]) {
          case 0: r = mvd_partition(0, bx, by, 2, 2); break;
          case 1: r = mvd_partition(0, bx, by, 2, 1); if (r >= 0) r = mvd_partition(0, bx, by + 1, 2, 1); break;
          case 2: r = mvd_partition(0, bx, by, 1, 2); if (r >= 0) r = mvd_partition(0, bx + 1, by, 1, 2); break;
          default:
            for (int j = 0; j < 4 && r >= 0; j++) r = mvd_partition(0, bx + (j & 1), by + (j >> 1), 1, 1);
        }
        if (r < 0) return r;
      }
    } else {  // 16x16, 16x8, 8x16
      const int nparts = p_type == 0 ? 1 : 2;
      for (int i = 0; i < nparts; i++) {
        int bx = p_type == 2 ? 2 * i : 0, by = p_type == 1 ? 2 * i : 0;
        int ref = 0;
        if (sp.num_ref[0] > 1) {
          ref = ref_idx(0, bx, by);
          if (ref < 0) return ref;
          if (ref >= sp.num_ref[0]) return SYN_E_REF;
        }
        if (p_type == 0) { for (int b = 0; b < 4; b++) set_ref(0, b, ref); }
        else if (p_type == 1) { set_ref(0, 2 * i, ref); set_ref(0, 2 * i + 1, ref); }
        else { set_ref(0, i, ref); set_ref(0, i + 2, ref); }
      }
      for (int i = 0; i < nparts; i++) {
        int bx = p_type == 2 ? 2 * i : 0, by = p_type == 1 ? 2 * i : 0;
        int w = p_type == 2 ? 2 : 4, h = p_type == 1 ? 2 : 4;
        int r = mvd_partition(0, bx, by, w, h);
        if (r < 0) return r;
      }
    }

    if (!(cur.flags & MBF_I16)) {  // coded_block_pattern (h264_cabac.c:1406-1440)
      int cbp_a = nb_cbp(L), cbp_b = nb_cbp(T);
      int c = 0, ctx;
      ctx = !(cbp_a & 0x02) + 2 * !(cbp_b & 0x04);
      c += io.get(73 + ctx);
      ctx = !(c & 0x01) + 2 * !(cbp_b & 0x08);
      c += io.get(73 + ctx) << 1;
      ctx = !(cbp_a & 0x08) + 2 * !(c & 0x01);
      c += io.get(73 + ctx) << 2;
      ctx = !(c & 0x04) + 2 * !(c & 0x02);
      c += io.get(73 + ctx) << 3;
      cbp = c;
      if (decode_chroma) {
        int ca = (cbp_a >> 4) & 3, cb = (cbp_b >> 4) & 3;
        ctx = (ca > 0) + 2 * (cb > 0);
        if (io.get(77 + ctx)) {
          ctx = 4 + (ca == 2) + 2 * (cb == 2);
          cbp |= (1 + io.get(77 + ctx)) << 4;
        }
      }
    }
    cur.cbp = (uint16_t)cbp;

    if (dct8x8_allowed && (cbp & 15) && !(cur.flags & MBF_INTRA)) {
      if (io.get(399 + nts)) cur.flags |= MBF_8x8DCT;
    }

    if (cbp || (cur.flags & MBF_I16)) {
      io.hint(HINT_DQP, 0);
      if (io.get(60 + (last_dqp_nz != 0))) {  // mb_qp_delta (h264_cabac.c:2391-2418)
        int ctx = 2, val = 1;
        while (io.get(60 + ctx)) {
          ctx = 3;
          val++;
          if (val > 2 * 51) return SYN_E_DQP;
        }
        last_dqp_nz = 1;
      } else {
        last_dqp_nz = 0;
      }
      residuals(cbp);
    } else {
      last_dqp_nz = 0;
    }
    row[x] = cur;
    return 0;
  }
};

}  // namespace avr
