#!/bin/bash
# decompress-kernel probe: isolates per-chain speed factors, then a light ncu capture on a tiny clip
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/probe; mkdir -p $T $R/gpurun_out
$G $T/a.h264 seed=1 mb_w=40 mb_h=30 frames=4 2>/dev/null
$G $T/b.h264 seed=1 mb_w=40 mb_h=30 frames=4 transform_8x8=1 2>/dev/null
$G $T/c.h264 seed=1 mb_w=40 mb_h=30 frames=4 slices_per_frame=4 2>/dev/null
$G $T/d.h264 seed=1 mb_w=40 mb_h=30 frames=4 num_ref=3 2>/dev/null
$G $T/e.h264 seed=1 mb_w=120 mb_h=68 frames=2 2>/dev/null
$G $T/t.h264 seed=1 mb_w=22 mb_h=18 frames=2 transform_8x8=1 2>/dev/null
for f in a b c d e t; do echo -n "$f: "; $B 1 2 $T/$f.h264; done
$B 1 2 $T/t.h264 > $R/gpurun_out/plain_t.log 2>&1 && timeout 200 ncu --section SourceCounters --section WarpStateStats --section InstructionStats --section SpeedOfLight --section MemoryWorkloadAnalysis --clock-control none --import-source on -k regex:avr_k_decompress_chain -c 1 -f -o $R/gpurun_out/prof_dchain_light $B 1 2 $T/t.h264 > $R/gpurun_out/ncu_t.log 2>&1
tail -2 $R/gpurun_out/ncu_t.log; ls -la $R/gpurun_out/*.ncu-rep
