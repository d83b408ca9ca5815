#!/bin/bash
# ncu (warp states, scheduler, memory, source counters) of the decompress chain kernel under load: N copies of one clip
# usage: gpu_r2_ncu_load.sh <outname> <copies>
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/r2l; mkdir -p $T $R/gpurun_out
O=$1; N=${2:-2048}
$G $T/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
args=""; for ((i=0;i<N;i++)); do args="$args $T/a.h264"; done
timeout 300 $B 1 2 $args > $R/gpurun_out/${O}_plain.log 2>&1; cat $R/gpurun_out/${O}_plain.log | sed 's/"bytes[^}]*"decompress_ms"/"decompress_ms"/'
timeout 900 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --section InstructionStats --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy \
  --clock-control none --import-source on -k regex:avr_k_decompress_chain -c 1 -f -o $R/gpurun_out/$O $B 1 2 $args > $R/gpurun_out/${O}_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 $R/gpurun_out/${O}_ncu.log
ncu -i $R/gpurun_out/$O.ncu-rep --page details > $R/gpurun_out/${O}_details.txt 2>&1; ls -la $R/gpurun_out/$O.ncu-rep
