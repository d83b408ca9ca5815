#!/bin/bash
# decompress chain kernel: per-chain rate against chains resident per SM (identical files, so no imbalance)
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/cont; mkdir -p $T
$G $T/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
for n in 1 148 296 592 1184 2368; do
  args=""; for ((i=0;i<n;i++)); do args="$args $T/a.h264"; done
  echo -n "chains=$n: "; $B 2 2 $args
done
