#!/bin/bash
# SIMT efficiency of the packed device parse: warp and thread instructions of avr_k_parse per AVR_PARSE_PACK, bench corpus (N clips)
R=/root/repo; T=/tmp/corpus; B=$R/avrecode-ms_b200/bin/avrbench; O=$R/gpurun_out; mkdir -p $T $O
cd $R && python tools/gen_corpus.py $T 0 ${N:-1024} > /dev/null || exit 1
for p in 1 4 8; do
  AVR_PARSE_PACK=$p AVR_LANES=1 $B 1 5 $T/*.h264 > $O/parse_simt_plain_$p.json 2>&1 || exit 1
  AVR_PARSE_PACK=$p AVR_LANES=1 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum \
    --clock-control none -k regex:avr_k_parse -c 1 --csv --log-file $O/parse_simt_$p.csv $B 1 5 $T/*.h264 > /dev/null 2>&1
  echo "pack=$p $(cat $O/parse_simt_plain_$p.json | grep -oE '"(bins|parse_ms)": [0-9.]*' | tr '\n' ' ') $(grep avr_k_parse $O/parse_simt_$p.csv | awk -F'","' '{print $(NF-2), $NF}' | tr -d '"' | tr '\n' ' ')"
done
