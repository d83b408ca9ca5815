#!/bin/bash
# parse kernel: slices per warp (AVR_PARSE_PACK) x share of slices that keep their own warp (AVR_PARSE_LONG), bench corpus, file API
R=/root/repo; T=/tmp/corpus; B=$R/avrecode-ms_b200/bin/avrbench; mkdir -p $T
cd $R && python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -2
python tools/gen_corpus.py $T 0 ${N:-2048} > /dev/null || exit 1
f='"(parse_ms|compress_kernel_ms|compress_wall_ms)": [0-9.]*'
for cfg in "1 45" "4 45" "8 45" "16 45" "32 45" "8 20" "16 20" "8 70" "8 0" "32 0"; do set -- $cfg
  echo "pack=$1 long=$2: $(AVR_PARSE_PACK=$1 AVR_PARSE_LONG=$2 $B 3 5 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
done
