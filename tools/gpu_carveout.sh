#!/bin/bash
# shared-memory carve-out of the chain kernel against its time on the bench corpus (file API, avrbench 1 3)
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus; N=${1:-2048}; mkdir -p $T
cd $R && python tools/gen_corpus.py $T 0 $N > /dev/null || exit 1
for c in auto 50 60 70 80 100; do
  if [ $c = auto ]; then unset AVR_DEC_CARVEOUT; else export AVR_DEC_CARVEOUT=$c; fi
  echo -n "carveout=$c: "; $B 2 3 $T/*.h264 | grep -oE '"decompress_kernel_ms": [0-9.]*'
done
