// Synthetic CABAC H.264 Annex-B clip generator.
//
// There is no H.264 encoder and no sample clip on the build or GPU boxes, and the reference
// ships no test inputs (its README names data/GOPR4542.MP4, which is not in the repository).
// This generator walks the same slice_data() state machine the product uses (h264_syntax.h)
// with a seeded random bin source and a textbook CABAC encoder (cabac_std.h), so every clip
// is a syntactically legal High-profile stream: SPS/PPS, IDR + P (+ B, p_bslice) pictures (every picture a
// reference, so frame_num changes on every picture), N slices per picture, optional 8x8
// transform.  Legality is judged by the reference's own libavcodec accepting it
// (tests/test_reference_parity.py runs oracle/_ref/recode on generated clips).
#pragma once
#include <stdint.h>
#include <vector>

namespace avr {

struct GenParams {
  uint64_t seed = 1;
  int mb_w = 8, mb_h = 6;
  int frames = 4;
  int slices_per_frame = 1;
  int idr_period = 0;          // 0: only the first picture is IDR
  int num_ref = 1;             // max_num_ref_frames / num_ref_idx_l0_active upper bound
  int transform_8x8 = 0;
  int qp = 26;
  int cabac_init_idc = 0;
  // statistics knobs, probabilities in 1/256 units
  int p_skip = 64;             // P(mb_skip_flag = 1)
  int p_intra_in_p = 24;       // P(intra macroblock in a P slice)
  int p_cbf = 160;             // P(coded_block_flag = 1)
  int p_cbp = 170;             // P(a coded_block_pattern bin = 1)
  int p_sig = 120;             // P(significant_coeff_flag = 1)
  int p_last = 40;             // P(last_significant_coeff_flag = 1)
  int p_gt1 = 80;              // P(coeff_abs_level_minus1 > 0)
  int p_gt1_more = 150;        // P(each further unary level bin = 1)
  int p_mvd_nz = 140;
  int p_mvd_more = 150;
  int p_dqp = 20;
  int p_t8x8 = 128;
  int p_i16 = 96;              // P(intra macroblock is I16x16 rather than I_NxN)
  int long_startcodes = 1;     // 1: 4-byte start code on the first NAL of each picture
  int p_bslice = 0;            // P(a non-IDR slice is a B slice); B pictures are reference pictures too, so frame_num keeps changing
  int p_direct = 70;           // P(B_Direct_16x16) and P(B_Direct_8x8)
  int repeat_frame_num = 0;    // 1: non-reference pictures so consecutive pictures share frame_num (out-of-domain)
};

// Annex-B byte stream.
std::vector<uint8_t> generate_clip(const GenParams& gp);

}  // namespace avr
