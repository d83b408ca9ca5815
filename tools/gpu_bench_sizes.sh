#!/bin/bash
# bench.py at several corpus sizes per GPU (first a small smoke run of every leg)
cd /root/repo; mkdir -p gpurun_out
timeout 600 python bench.py --clips 128 --strong-clips 48 --steps 1 --warmup 1 > gpurun_out/bench_smoke.json 2> gpurun_out/bench_smoke.err; echo "smoke rc=$?"; tail -c 600 gpurun_out/bench_smoke.err; cut -c1-400 gpurun_out/bench_smoke.json
for n in "$@"; do
  timeout 900 python bench.py --clips $n --strong-clips 0 --no-single --steps 3 --warmup 1 > gpurun_out/bench_c$n.json 2> gpurun_out/bench_c$n.err; echo "clips=$n rc=$?"; tail -c 300 gpurun_out/bench_c$n.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/bench_c$n.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "compress_gbins_s", "decompress_gbins_s", "bit_exact")}, "e2e", d["e2e"]["value"], d["e2e"]["rank0_parts"])
except Exception as e:
    print("no line", e)
PY
done
