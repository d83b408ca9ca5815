#!/bin/bash
# per-chain / per-slice rates of the three serial kernels on single clips (avrbench mode 3 = file API)
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/cr; mkdir -p $T
$G $T/a.h264 seed=1 mb_w=40 mb_h=30 frames=4 2>/dev/null
$G $T/b.h264 seed=1 mb_w=40 mb_h=30 frames=4 transform_8x8=1 2>/dev/null
$G $T/e.h264 seed=1 mb_w=80 mb_h=45 frames=2 2>/dev/null
for f in a b e; do echo -n "$f: "; $B 2 3 $T/$f.h264; done
