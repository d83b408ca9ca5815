#!/bin/bash
# round 2: chains per SM vs per-chain rate (N copies of ONE clip: no imbalance); usage: gpu_r2_curve.sh [counts...]
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/r2c; mkdir -p $T
$G $T/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
for n in ${@:-1 148 592 1024 2048}; do
  args=""; for ((i=0;i<n;i++)); do args="$args $T/a.h264"; done
  echo -n "chains=$n: "; timeout 300 $B 2 2 $args | sed 's/"bytes[^}]*"decompress_ms"/"decompress_ms"/'; echo " rc=$?"
done
