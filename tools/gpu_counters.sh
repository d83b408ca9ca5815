#!/bin/bash
# Round 2: counters of the product path on the bench corpus for the committed build, in one call:
#  (1) ncu launch list (gpu__time_duration) of avr_compress_files + avr_decompress_files,
#  (2) per-launch DRAM bytes, executed warp instructions, issue activity and L1 hit rate of the three serial kernels
# -> gpurun_out/r02_launches_fileapi.csv, r02_counters.csv, r02_chain_counters.json (bench.py reads profiles/r02_chain_counters.json)
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus; N=${1:-2048}; mkdir -p $T $R/gpurun_out
cd $R && python tools/gen_corpus.py $T 0 $N > /dev/null || exit 1
$B 1 3 $T/*.h264 > $R/gpurun_out/r02_counters_plain.json 2>&1 || { echo "plain run failed"; tail -3 $R/gpurun_out/r02_counters_plain.json; exit 1; }
cat $R/gpurun_out/r02_counters_plain.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $R/gpurun_out/r02_launches_fileapi.csv $B 1 3 $T/*.h264 > $R/gpurun_out/r02_launches_ncu.log 2>&1
echo "launch list rc=$?"; wc -l $R/gpurun_out/r02_launches_fileapi.csv
timeout 1200 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active \
  --clock-control none -k regex:'avr_k_decompress_chain|avr_k_parse|avr_k_code' --csv --log-file $R/gpurun_out/r02_counters.csv $B 1 3 $T/*.h264 > $R/gpurun_out/r02_counters_ncu.log 2>&1
echo "counters rc=$?"
python - <<PY
import csv, json
plain = json.load(open("$R/gpurun_out/r02_counters_plain.json"))
rows = [r for r in csv.reader(open("$R/gpurun_out/r02_counters.csv")) if len(r) > 10]
hdr = rows[0]; iK = hdr.index("Kernel Name"); iM = hdr.index("Metric Name"); iV = hdr.index("Metric Value"); iU = hdr.index("Metric Unit")
agg = {}
for r in rows[1:]:
    k = r[iK].split("(")[0].replace("void ", "").split("<")[0].strip(); v = float(r[iV].replace(",", "")); u = r[iU]
    if r[iM].startswith("gpu__time_duration"): v = v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v
    if r[iM].startswith("dram__bytes"): v = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    agg.setdefault(k, {}).setdefault(r[iM], []).append(v)
out = {"source": "tools/gpu_counters.sh: ncu --clock-control none on avrbench 1 3 (file API) over the %d-clip bench corpus; sums over the launches of one compress + decompress call" % $N,
       "clips": $N, "bins": plain["bins"], "kernels": {}}
for k, m in agg.items():
    out["kernels"][k] = {"launches": len(m["gpu__time_duration.sum"]), "ms_under_ncu": sum(m["gpu__time_duration.sum"]),
                         "dram_read": sum(m["dram__bytes_read.sum"]), "dram_write": sum(m["dram__bytes_write.sum"]),
                         "inst_executed": sum(m["smsp__inst_executed.sum"]),
                         "issue_active_pct": sum(m["smsp__issue_active.avg.pct_of_peak_sustained_active"]) / len(m["gpu__time_duration.sum"]),
                         "l1_hit_pct": sum(m["l1tex__t_sector_hit_rate.pct"]) / len(m["gpu__time_duration.sum"]),
                         "warps_active_pct": sum(m["sm__warps_active.avg.pct_of_peak_sustained_active"]) / len(m["gpu__time_duration.sum"])}
ch = out["kernels"].get("avr_k_decompress_chain")
if ch:
    out["inst_executed_per_bin"] = ch["inst_executed"] / plain["bins"]
    out["dram_bytes_per_launch"] = (ch["dram_read"] + ch["dram_write"]) / ch["launches"]
    out["issue_active_pct"] = ch["issue_active_pct"]; out["l1_hit_pct"] = ch["l1_hit_pct"]
json.dump(out, open("$R/gpurun_out/r02_chain_counters.json", "w"), indent=1)
print(json.dumps(out)[:1500])
PY
