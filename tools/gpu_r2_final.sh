#!/bin/bash
# round-2 evidence in one call: smoke, bench (both arms), counters + launch list on the bench corpus, one full ncu capture of the chain kernel
R=/root/repo; cd $R; mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/r02_bench_final_1gpu.json 2> gpurun_out/r02_bench_final_1gpu.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench_final_1gpu.err
python bench.py --impl reference --steps 2 --warmup 0 > gpurun_out/r02_bench_final_reference_1gpu.json 2>> gpurun_out/r02_bench_final_1gpu.err; echo "ref rc=$?"
tools/gpu_counters.sh 2048 2>&1 | tail -3 | cut -c1-600
G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench
$G /tmp/one.h264 seed=6 mb_w=44 mb_h=30 frames=2 2>/dev/null
$B 1 2 /tmp/one.h264 > gpurun_out/r02_full_plain.log 2>&1 && timeout 400 ncu --set full --clock-control none --import-source on -k regex:avr_k_decompress_chain -c 1 -f -o gpurun_out/r02_dchain_full $B 1 2 /tmp/one.h264 > gpurun_out/r02_full_ncu.log 2>&1
echo "full ncu rc=$?"; cat gpurun_out/r02_full_plain.log
ncu -i gpurun_out/r02_dchain_full.ncu-rep --page details > gpurun_out/r02_dchain_full_details.txt 2>&1
