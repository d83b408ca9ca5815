#!/bin/bash
# old (one static AvrParShared per block) against new (per-lane instances) device parse: time and instructions, N clips, one lane
R=/root/repo; T=/tmp/corpus; O=$R/gpurun_out; mkdir -p $T $O
cd $R && python tools/gen_corpus.py $T 0 ${N:-1024} > /dev/null || exit 1
for cfg in "ab_oldparse 1" "avrecode-ms_b200 1" "avrecode-ms_b200 4"; do set -- $cfg; B=$R/$1/bin/avrbench
  plain=$(AVR_PARSE_PACK=$2 AVR_LANES=1 $B 2 5 $T/*.h264 | grep -oE '"(bins|parse_ms|compress_wall_ms)": [0-9.]*' | tr '\n' ' ')
  AVR_PARSE_PACK=$2 AVR_LANES=1 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -k regex:avr_k_parse -c 1 --csv --log-file $O/po.csv $B 1 5 $T/*.h264 > /dev/null 2>&1
  echo "$1 pack=$2 $plain $(grep avr_k_parse $O/po.csv | awk -F'","' '{print $(NF-2), $NF}' | tr -d '"' | tr '\n' ' ')"
done
