#!/bin/bash
# product chain (avr_decode.h) against the experimental warp-speculative chain (avr_decode_spec.h, AVR_DEC_SPEC=1): same box, same clips
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; O=$R/gpurun_out; mkdir -p /tmp/c6 $O
cd $R && python -m pytest tests/test_gpu_parity.py -q -m gpu -k "speculative or lanes_and_multi or golden" 2>&1 | tail -2
$G /tmp/c6/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
$G /tmp/c6/p.h264 seed=6 mb_w=44 mb_h=30 frames=2 2>/dev/null
same=""; for ((i=0;i<1024;i++)); do same="$same /tmp/c6/a.h264"; done
f='"(decompress_ms|bins)": [0-9.]*'
for rep in 1 2; do
  for v in 0 1; do
    echo "spec=$v a: $(AVR_DEC_SPEC=$v $B 2 2 /tmp/c6/a.h264 | grep -oE "$f" | tr '\n' ' ') | p: $(AVR_DEC_SPEC=$v $B 2 2 /tmp/c6/p.h264 | grep -oE "$f" | tr '\n' ' ') | same1024: $(AVR_DEC_SPEC=$v $B 2 2 $same | grep -oE "$f" | tr '\n' ' ')"
  done
done | tee $O/spec_ab.txt
# instructions and time of the two kernels on ONE chain (after the plain runs above exited 0)
for v in 0 1; do
  AVR_DEC_SPEC=$v ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active \
    --clock-control none -k regex:avr_k_decompress_chain --csv --log-file $O/spec_counters_$v.csv $B 1 1 /tmp/c6/a.h264 > $O/spec_ncu_$v.log 2>&1
done
grep -h "avr_k_decompress" $O/spec_counters_*.csv | cut -d, -f5,13- | head -20
