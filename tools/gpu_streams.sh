#!/bin/bash
# config 1 (independent bin streams): rates at three shapes and one full ncu capture of the stream kernels
# (compute-sanitizer is closed on this pool: memory safety is covered by the ASan build of the host-emulated bodies)
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; G=$R/avrecode-ms_b200/bin/avrgen; mkdir -p $R/gpurun_out /tmp/sc
for shape in "8192 16384" "65536 2048" "262144 512"; do for c in "16 0" "8 1"; do echo -n "streams $shape coder $c: "; $B 3 4 $shape $c; done; done
$B 1 4 65536 2048 8 1 > $R/gpurun_out/streams_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:avr_k_stream -c 2 -f -o $R/gpurun_out/streams_full $B 1 4 65536 2048 8 1 > $R/gpurun_out/streams_ncu.log 2>&1
echo "ncu rc=$?"
