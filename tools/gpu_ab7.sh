#!/bin/bash
# gpu_ab7.sh <variant>...: chain kernel on 1 / 592 / 1024 copies of one clip per build variant, twice
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; mkdir -p /tmp/cont
$G /tmp/cont/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
s592=""; for ((i=0;i<592;i++)); do s592="$s592 /tmp/cont/a.h264"; done
s1024=""; for ((i=0;i<1024;i++)); do s1024="$s1024 /tmp/cont/a.h264"; done
f='"(decompress_ms)": [0-9.]*'
for rep in 1 2; do for v in "$@"; do b=$R/ab_$v/bin/avrbench
  echo "$v single $($b 2 2 /tmp/cont/a.h264 | grep -oE "$f") | same592 $($b 2 2 $s592 | grep -oE "$f") | same1024 $($b 2 2 $s1024 | grep -oE "$f")"
done; done
