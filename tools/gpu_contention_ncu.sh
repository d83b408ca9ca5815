#!/bin/bash
# decompress chain kernel under 1 / 4 chains per SM: warp states, scheduler, memory (what the slowdown of a chain that shares its SM is made of)
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/cont; mkdir -p $T $R/gpurun_out
$G $T/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
for n in 148 592 1024; do
  args=""; for ((i=0;i<n;i++)); do args="$args $T/a.h264"; done
  $B 1 2 $args > $R/gpurun_out/cont_${n}_plain.log 2>&1 || { echo "plain failed $n"; continue; }
  cat $R/gpurun_out/cont_${n}_plain.log
  timeout 400 ncu --section WarpStateStats --section SchedulerStats --section MemoryWorkloadAnalysis --section SpeedOfLight --section InstructionStats --section SourceCounters \
    --clock-control none --import-source on -k regex:avr_k_decompress_chain -c 1 -f -o $R/gpurun_out/cont_$n $B 1 2 $args > $R/gpurun_out/cont_${n}_ncu.log 2>&1
  echo "ncu rc=$?"
done
ls -la $R/gpurun_out/*.ncu-rep
