#!/bin/bash
# gpu_ab8.sh <variant>...: chain kernel on 1024 / 2048 copies of one clip per build variant, once
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; mkdir -p /tmp/cont
$G /tmp/cont/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
same=""; for ((i=0;i<1024;i++)); do same="$same /tmp/cont/a.h264"; done
f='"(decompress_ms)": [0-9.]*'
for v in "$@"; do b=$R/ab_$v/bin/avrbench
  echo "$v same1024 $($b 2 2 $same | grep -oE "$f") | same2048 $($b 2 2 $same $same | grep -oE "$f")"
done
