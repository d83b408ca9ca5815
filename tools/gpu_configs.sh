#!/bin/bash
# BASELINE configs[1], [2], [4] as single clips (frame counts reduced): compress pipeline and the one decompress chain, verified
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/cfg; mkdir -p $T
$G $T/c2.h264 seed=2 mb_w=120 mb_h=68 frames=8 slices_per_frame=1 2>/dev/null
$G $T/c3.h264 seed=3 mb_w=120 mb_h=68 frames=30 slices_per_frame=8 2>/dev/null
$G $T/c3b.h264 seed=3 mb_w=120 mb_h=68 frames=30 slices_per_frame=8 transform_8x8=1 2>/dev/null
$G $T/c5.h264 seed=5 mb_w=240 mb_h=135 frames=4 slices_per_frame=68 p_sig=150 p_cbf=230 p_cbp=240 transform_8x8=1 2>/dev/null
for c in c2 c3 c3b c5; do echo -n "$c file API: "; $B 2 3 $T/$c.h264; echo -n "$c resident: "; $B 2 0 $T/$c.h264; done
