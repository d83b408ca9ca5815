#!/bin/bash
# gpu_ab4.sh <variant>...: file-API kernel times on the bench corpus (2048 clips) per build variant and carve-out setting
R=/root/repo; T=/tmp/corpus; mkdir -p $T
cd $R && python tools/gen_corpus.py $T 0 ${N:-2048} > /dev/null || exit 1
f='"(parse_ms|compress_kernel_ms|coder_ms|decompress_kernel_ms)": [0-9.]*'
for rep in 1 2; do for v in "$@"; do for c in auto 70; do
  if [ $c = auto ]; then unset AVR_DEC_CARVEOUT; else export AVR_DEC_CARVEOUT=$c; fi
  echo "$v carveout=$c: $($R/ab_$v/bin/avrbench 2 3 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
done; done; done
