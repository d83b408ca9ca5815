#!/bin/bash
# gpu_ab2.sh <variant>...: same-box A/B of the decompress chain kernel (ab_<variant>/ from tools/mk_variant.sh): one chain, 1024 and 2048
# copies of one clip, and the bench corpus; every variant twice, interleaved.  Only numbers of one call compare.
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; T=/tmp/corpus; mkdir -p $T /tmp/cont
cd $R && python tools/gen_corpus.py $T 0 1024 > /dev/null || exit 1
$G /tmp/cont/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
same=""; for ((i=0;i<1024;i++)); do same="$same /tmp/cont/a.h264"; done
f='"(decompress_ms)": [0-9.]*'
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
for rep in 1 2; do
  for v in "$@"; do
    b=$R/ab_$v/bin/avrbench
    echo "$v single $($b 2 2 /tmp/cont/a.h264 | grep -oE "$f") | same1024 $($b 2 2 $same | grep -oE "$f") | same2048 $($b 2 2 $same $same | grep -oE "$f") | corpus $($b 2 2 $T/*.h264 | grep -oE "$f")"
  done
done
