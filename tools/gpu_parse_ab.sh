#!/bin/bash
# device parse: old library (ab_oldparse, one kernel, every slice its own warp) against the two-kernel layout, one lane and two lanes
R=/root/repo; T=/tmp/corpus; mkdir -p $T
cd $R && python -m pytest tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -1
python tools/gen_corpus.py $T 0 2048 > /dev/null || exit 1
f='"(parse_ms|compress_kernel_ms|compress_wall_ms)": [0-9.]*'
first=$(ls $T/*.h264 | head -1024 | tr '\n' ' ')
for cfg in "ab_oldparse 1 35" "avrecode-ms_b200 1 35" "avrecode-ms_b200 4 35" "avrecode-ms_b200 4 50" "avrecode-ms_b200 8 35" "avrecode-ms_b200 8 20" "avrecode-ms_b200 4 20" "avrecode-ms_b200 4 70"; do set -- $cfg; B=$R/$1/bin/avrbench
  echo "$1 pack=$2 long=$3 | one lane, 1024 clips: $(AVR_PARSE_PACK=$2 AVR_PARSE_LONG=$3 AVR_LANES=1 $B 3 5 $first | grep -oE "$f" | tr '\n' ' ') | two lanes, 2048 clips: $(AVR_PARSE_PACK=$2 AVR_PARSE_LONG=$3 $B 3 5 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
done
