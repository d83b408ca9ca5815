#!/bin/bash
# pass D hybrid: slices up to AVR_CODE_LANE_MAX bins share a warp (one per lane); sweep of the threshold on the bench corpus (1024 clips, one batch)
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus1k; mkdir -p $T
cd $R && python tools/gen_corpus.py $T 0 1024 > /dev/null || exit 1
for m in ${@:-0 50000 100000 150000 250000 400000 1000000000}; do
  echo "lane_max=$m: $(AVR_CODE_LANE_MAX=$m $B 3 0 $T/*.h264 | grep -oE '"(compress_ms|coder_ms|bad_slices)": [0-9.]*' | tr '\n' ' ')"
done
