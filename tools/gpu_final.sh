#!/bin/bash
# round-end evidence in one call: gpu tests, bench (both arms), ncu launch list + DRAM traffic on the bench corpus, one full capture of the dominant kernel
R=/root/repo; cd $R; mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"; tail -c 400 gpurun_out/bench_final.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_final.json 2>> gpurun_out/bench_final.err; echo "ref rc=$?"
tools/gpu_launchlist_corpus.sh 2>&1 | tail -4
tools/gpu_traffic.sh 2>&1 | tail -6
G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench
$G /tmp/one.h264 seed=6 mb_w=44 mb_h=30 frames=2 2>/dev/null
$B 1 2 /tmp/one.h264 > gpurun_out/full_plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:avr_k_decompress_chain -c 1 -f -o gpurun_out/dchain_full $B 1 2 /tmp/one.h264 > gpurun_out/full_ncu.log 2>&1
echo "full ncu rc=$?"; cat gpurun_out/full_plain.log
