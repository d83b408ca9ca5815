#!/bin/bash
# ncu launch list of the whole product path (file API, device parse) on a small batch of clips.
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/ll; mkdir -p $T $R/gpurun_out
for i in 1 2 3 4 5 6 7 8; do $G $T/c$i.h264 seed=$i mb_w=$((20+4*i)) mb_h=$((12+3*i)) frames=3 slices_per_frame=$((1+i%3)) transform_8x8=$((i%2)) 2>/dev/null; done
$B 2 3 $T/c*.h264 > $R/gpurun_out/launchlist_plain.log 2>&1 || { echo "plain run failed"; cat $R/gpurun_out/launchlist_plain.log; exit 1; }
cat $R/gpurun_out/launchlist_plain.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $R/gpurun_out/launches.csv $B 1 3 $T/c*.h264 > $R/gpurun_out/launchlist_ncu.log 2>&1
echo "ncu rc=$?"; wc -l $R/gpurun_out/launches.csv
