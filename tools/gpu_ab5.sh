#!/bin/bash
# single-chain and loaded chain times per build variant (two clips)
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; mkdir -p /tmp/c5
$G /tmp/c5/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
$G /tmp/c5/p.h264 seed=6 mb_w=44 mb_h=30 frames=2 2>/dev/null
same=""; for ((i=0;i<1024;i++)); do same="$same /tmp/c5/a.h264"; done
f='"(decompress_ms)": [0-9.]*'
for rep in 1 2; do for v in "$@"; do b=$R/ab_$v/bin/avrbench
  echo "$v a: $($b 2 2 /tmp/c5/a.h264 | grep -oE "$f") | p: $($b 2 2 /tmp/c5/p.h264 | grep -oE "$f") | same1024: $($b 2 2 $same | grep -oE "$f")"
done; done
