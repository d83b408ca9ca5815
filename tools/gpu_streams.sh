#!/bin/bash
# config 1 (independent bin streams): rates at two shapes, one full ncu capture of the encode kernel, and compute-sanitizer memcheck
# of the stream kernels and of the file API on small clips
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; G=$R/avrecode-ms_b200/bin/avrgen; mkdir -p $R/gpurun_out /tmp/sc
for shape in "8192 16384" "65536 2048" "262144 512"; do for c in "16 0" "8 1"; do echo -n "streams $shape coder $c: "; $B 3 4 $shape $c; done; done
$B 1 4 65536 2048 8 1 > $R/gpurun_out/streams_plain.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:avr_k_stream -c 2 -f -o $R/gpurun_out/streams_full $B 1 4 65536 2048 8 1 > $R/gpurun_out/streams_ncu.log 2>&1
echo "ncu rc=$?"
for c in "16 0" "8 1" "8 0" "16 1"; do timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 $B 1 4 96 1001 $c > $R/gpurun_out/memcheck_streams.log 2>&1; echo "memcheck streams $c rc=$? $(grep -c 'ERROR SUMMARY: 0 errors' $R/gpurun_out/memcheck_streams.log)"; done
$G /tmp/sc/a.h264 seed=3 mb_w=8 mb_h=6 frames=3 slices_per_frame=2 2>/dev/null
$G /tmp/sc/b.h264 seed=4 mb_w=8 mb_h=6 frames=2 transform_8x8=1 2>/dev/null
$G /tmp/sc/c.h264 seed=5 mb_w=10 mb_h=5 frames=2 num_ref=3 p_skip=150 2>/dev/null
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 $B 1 3 /tmp/sc/a.h264 /tmp/sc/b.h264 /tmp/sc/c.h264 > $R/gpurun_out/memcheck_fileapi.log 2>&1; echo "memcheck fileapi rc=$?"; tail -3 $R/gpurun_out/memcheck_fileapi.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 $B 1 1 /tmp/sc/a.h264 /tmp/sc/b.h264 > $R/gpurun_out/memcheck_chain.log 2>&1; echo "memcheck chain mode rc=$?"; tail -2 $R/gpurun_out/memcheck_chain.log
