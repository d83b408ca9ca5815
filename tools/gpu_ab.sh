#!/bin/bash
# gpu_ab.sh <variant>...: same-box A/B of builds of the library (ab_<variant>/, made by tools/mk_variant.sh).
# Only numbers taken in one call compare (boxes differ), and each variant is run twice, interleaved.
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; T=/tmp/corpus; mkdir -p $T /tmp/cont
cd $R && python tools/gen_corpus.py $T 0 1024 > /dev/null || exit 1
$G /tmp/cont/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
same=""; for ((i=0;i<1024;i++)); do same="$same /tmp/cont/a.h264"; done
f='"(coder_ms|decompress_ms|parse_ms|compress_kernel_ms|decompress_kernel_ms)": [0-9.]*'
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
for rep in 1 2; do
  for v in "$@"; do
    b=$R/ab_$v/bin/avrbench
    echo "$v corpus:   $($b 2 0 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
    echo "$v same1024: $($b 2 2 $same | grep -oE "$f" | tr '\n' ' ')"
    echo "$v single:   $($b 2 2 /tmp/cont/a.h264 | grep -oE "$f" | tr '\n' ' ')"
    [ -n "$AB_FILEAPI" ] && echo "$v fileapi:  $($b 2 3 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
  done
done
