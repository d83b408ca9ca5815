#!/bin/bash
R=/root/repo; T=/tmp/corpus; mkdir -p $T; B=$R/ab_snake/bin/avrbench
cd $R && python tools/gen_corpus.py $T 0 1024 > /dev/null || exit 1
f='"(coder_ms|decompress_ms|parse_ms|compress_kernel_ms|decompress_kernel_ms)": [0-9.]*'
for rep in 1 2; do
echo "snake:    $($B 2 2 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
echo "no snake: $(AVR_DEC_NO_SNAKE=1 $B 2 2 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"
done
for d in 0 6000 11000 16000 25000 40000; do echo "parse dyn smem $d: $(AVR_PARSE_DYN_SMEM=$d $B 2 3 $T/*.h264 | grep -oE "$f" | tr '\n' ' ')"; done
AVR_TRACE=1 $B 2 3 $T/*.h264 2>&1 | tail -60
