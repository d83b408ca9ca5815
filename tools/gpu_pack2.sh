#!/bin/bash
R=/root/repo; T=/tmp/corpus; B=$R/avrecode-ms_b200/bin/avrbench; mkdir -p $T
cd $R && python tools/gen_corpus.py $T 0 ${N:-2048} > /dev/null || exit 1
f='"(parse_ms|compress_kernel_ms|compress_wall_ms)": [0-9.]*'
for cfg in "$@"; do set -- $cfg
  echo "pack=$1 long=$2: $(AVR_PARSE_PACK=$1 AVR_PARSE_LONG=$2 $B 3 5 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
done
