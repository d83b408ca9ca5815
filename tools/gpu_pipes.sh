#!/bin/bash
# which pipe binds the serial kernels under load: ComputeWorkloadAnalysis / InstructionStats of chain, parse and pass D, 2048 copies of one clip
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; O=$R/gpurun_out; mkdir -p /tmp/c7 $O
$G /tmp/c7/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
same=""; for ((i=0;i<2048;i++)); do same="$same /tmp/c7/a.h264"; done
$B 1 1 $same > $O/pipes_plain.json 2>&1 || exit 1
ncu --section ComputeWorkloadAnalysis --section InstructionStats --section SchedulerStats --section WarpStateStats --section SpeedOfLight \
  --clock-control none -k regex:"avr_k_(decompress_chain|parse|code)" -o $O/pipes_same2048 -f $B 1 1 $same > $O/pipes_ncu.log 2>&1
ncu -i $O/pipes_same2048.ncu-rep --page details --print-details all > $O/pipes_same2048_details.txt 2>&1
grep -c . $O/pipes_same2048_details.txt
