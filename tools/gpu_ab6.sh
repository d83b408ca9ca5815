#!/bin/bash
# gpu_ab6.sh <variant>...: file-API kernel times on the bench corpus (N clips, default 2048) and single-chain times, per build variant, twice
R=/root/repo; T=/tmp/corpus; G=$R/avrecode-ms_b200/bin/avrgen; mkdir -p $T /tmp/c5
cd $R && python tools/gen_corpus.py $T 0 ${N:-2048} > /dev/null || exit 1
$G /tmp/c5/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
$G /tmp/c5/p.h264 seed=6 mb_w=44 mb_h=30 frames=2 2>/dev/null
f='"(parse_ms|compress_kernel_ms|coder_ms|decompress_kernel_ms)": [0-9.]*'
g='"(parse_ms|decompress_ms)": [0-9.]*'
for rep in 1 2; do for v in "$@"; do b=$R/ab_$v/bin/avrbench
  echo "$v corpus: $($b 2 3 $T/*.h264 | grep -oE "$f" | tr '\n' ' ') | a: $($b 2 2 /tmp/c5/a.h264 | grep -oE "$g" | tr '\n' ' ') | p: $($b 2 2 /tmp/c5/p.h264 | grep -oE "$g" | tr '\n' ' ')"
done; done
