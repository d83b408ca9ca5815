#!/bin/bash
# avr_k_parse and avr_k_code on the bench corpus (thousands of slices resident): warp states, scheduler, memory, per-line samples
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus; mkdir -p $T $R/gpurun_out
cd $R && python tools/gen_corpus.py $T 0 1024 > /dev/null || exit 1
$B 1 3 $T/*.h264 > $R/gpurun_out/parse_corpus_plain.log 2>&1 || { echo "plain failed"; exit 1; }
cat $R/gpurun_out/parse_corpus_plain.log
for k in avr_k_parse avr_k_code; do
timeout 600 ncu --section WarpStateStats --section SchedulerStats --section MemoryWorkloadAnalysis --section SpeedOfLight --section InstructionStats --section SourceCounters --section Occupancy --section LaunchStats \
  --clock-control none --import-source on -k regex:$k -c 1 -f -o $R/gpurun_out/corpus_$k $B 1 3 $T/*.h264 > $R/gpurun_out/corpus_${k}_ncu.log 2>&1
echo "$k ncu rc=$?"
done
ls -la $R/gpurun_out/*.ncu-rep
