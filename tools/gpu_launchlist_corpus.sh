#!/bin/bash
# ncu launch list (gpu__time_duration only) of the whole product path on the bench corpus
R=/root/repo; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus; mkdir -p $T $R/gpurun_out
cd $R && python tools/gen_corpus.py $T 0 ${1:-1024} > /dev/null || exit 1
$B 1 3 $T/*.h264 > $R/gpurun_out/launchlist_corpus_plain.log 2>&1 || { echo "plain run failed"; tail -3 $R/gpurun_out/launchlist_corpus_plain.log; exit 1; }
cat $R/gpurun_out/launchlist_corpus_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $R/gpurun_out/launches_corpus.csv $B 1 3 $T/*.h264 > $R/gpurun_out/launchlist_corpus_ncu.log 2>&1
echo "ncu rc=$?"; wc -l $R/gpurun_out/launches_corpus.csv
