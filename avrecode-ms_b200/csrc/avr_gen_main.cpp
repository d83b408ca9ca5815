// avrgen: writes a synthetic CABAC H.264 Annex-B clip (see h264_gen.h).
//   avrgen out.h264 [key=value ...]   keys = GenParams field names
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include "h264_gen.h"

int main(int argc, char** argv) {
  if (argc < 2) { fprintf(stderr, "usage: %s out.h264 [key=value ...]\n", argv[0]); return 1; }
  avr::GenParams gp;
  for (int i = 2; i < argc; i++) {
    char* eq = strchr(argv[i], '=');
    if (!eq) { fprintf(stderr, "bad arg %s\n", argv[i]); return 1; }
    std::string k(argv[i], eq - argv[i]);
    long long v = atoll(eq + 1);
#define K(name) if (k == #name) { gp.name = (decltype(gp.name))v; continue; }
    K(seed) K(mb_w) K(mb_h) K(frames) K(slices_per_frame) K(idr_period) K(num_ref) K(transform_8x8) K(qp)
    K(cabac_init_idc) K(p_skip) K(p_intra_in_p) K(p_cbf) K(p_cbp) K(p_sig) K(p_last) K(p_gt1) K(p_gt1_more)
    K(p_mvd_nz) K(p_mvd_more) K(p_dqp) K(p_t8x8) K(p_i16) K(long_startcodes) K(repeat_frame_num) K(p_bslice) K(p_direct)
#undef K
    fprintf(stderr, "unknown key %s\n", k.c_str()); return 1;
  }
  std::vector<uint8_t> clip = avr::generate_clip(gp);
  FILE* f = fopen(argv[1], "wb");
  if (!f) { perror("fopen"); return 1; }
  fwrite(clip.data(), 1, clip.size(), f);
  fclose(f);
  fprintf(stderr, "wrote %zu bytes\n", clip.size());
  return 0;
}
