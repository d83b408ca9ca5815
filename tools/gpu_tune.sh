#!/bin/bash
# serial kernels on the bench corpus and on 1024 copies of one clip (no imbalance): avrbench mode 2 (decompress timed), mode 0 (pipeline)
R=/root/repo; G=$R/avrecode-ms_b200/bin/avrgen; B=$R/avrecode-ms_b200/bin/avrbench; T=/tmp/corpus; mkdir -p $T /tmp/cont
cd $R && python tools/gen_corpus.py $T 0 1024 > /dev/null || exit 1
$G /tmp/cont/a.h264 seed=1 mb_w=40 mb_h=30 frames=3 2>/dev/null
$G /tmp/cont/b.h264 seed=1 mb_w=40 mb_h=30 frames=3 transform_8x8=1 2>/dev/null
same=""; for ((i=0;i<1024;i++)); do same="$same /tmp/cont/a.h264"; done
same8=""; for ((i=0;i<1024;i++)); do same8="$same8 /tmp/cont/b.h264"; done
echo -n "corpus: "; $B 2 0 $T/*.h264
echo -n "same1024: "; $B 2 2 $same
echo -n "same1024 8x8: "; $B 2 2 $same8
echo -n "single: "; $B 2 2 /tmp/cont/a.h264
echo -n "single 8x8: "; $B 2 2 /tmp/cont/b.h264
