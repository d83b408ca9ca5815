#!/bin/bash
# DRAM traffic per launch of the serial kernels on the bench corpus (1024 clips), two-metric ncu pass
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus; mkdir -p $T $R/gpurun_out
cd $R && python tools/gen_corpus.py $T 0 ${1:-1024} > /dev/null || exit 1
ls $T/*.h264 | wc -l
$B 1 3 $T/*.h264 > $R/gpurun_out/traffic_plain.log 2>&1 || { echo "plain run failed"; tail -3 $R/gpurun_out/traffic_plain.log; exit 1; }
cat $R/gpurun_out/traffic_plain.log
timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:'avr_k_decompress_chain|avr_k_parse|avr_k_code' --csv --log-file $R/gpurun_out/traffic.csv $B 1 3 $T/*.h264 > $R/gpurun_out/traffic_ncu.log 2>&1
echo "ncu rc=$?"; cat $R/gpurun_out/traffic.csv | tail -12
