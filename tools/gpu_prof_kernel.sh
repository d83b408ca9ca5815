#!/bin/bash
# light ncu capture (source counters + warp states) of ONE kernel on one small clip: gpu_prof_kernel.sh <kernel regex> <outname> [avrgen args]
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/pk; mkdir -p $T $R/gpurun_out
K=$1; O=$2; shift 2
$G $T/c.h264 seed=6 mb_w=44 mb_h=30 frames=2 "$@" 2>/dev/null
$B 1 3 $T/c.h264 > $R/gpurun_out/${O}_plain.log 2>&1 || { echo "plain run failed"; cat $R/gpurun_out/${O}_plain.log; exit 1; }
cat $R/gpurun_out/${O}_plain.log
timeout 280 ncu --section SourceCounters --section WarpStateStats --section InstructionStats --section SpeedOfLight --section MemoryWorkloadAnalysis --section LaunchStats --section Occupancy \
  --clock-control none --import-source on -k regex:$K -c 1 -f -o $R/gpurun_out/$O $B 1 3 $T/c.h264 > $R/gpurun_out/${O}_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 $R/gpurun_out/${O}_ncu.log; ls -la $R/gpurun_out/$O.ncu-rep
