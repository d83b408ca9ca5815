#!/bin/bash
# TEST INFRASTRUCTURE: builds the UNMODIFIED reference (pbluc/avrecode-ms) from the sources
# where they lie under /root/reference into oracle/_ref/ (git-ignored, travels to the GPU box).
# Nothing under oracle/_ref is linked into or called from the product path; it is the
# parity checker (tests/) and the `--impl reference` / cpu_baseline arm of bench.py.
#
# Recipe = SURVEY.md Appendix A, but out-of-tree (reference tree is read-only):
#   1. ffmpeg (dropbox/libavcodec-hooks) configured in a scratch dir with the minimal
#      component set, --disable-inline-asm (mandatory on x86, h264_cabac.c:1683), then
#      HAVE_ARC4RANDOM 0 in the generated config.h.
#   2. recode.cpp + test.cpp compiled -O2 -DNDEBUG (no ASan) against oracle/shim/recode.pb.h.
#   3. test/arithmetic_code.cpp as-is.
#   4. oracle/ref_harness.cpp (includes reference recode.cpp; drives its own classes).
#   5. oracle/gpu_driver.cpp (includes reference recode.cpp; its hooks feed include/avrecode_b200.h, links the product .so).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${AVR_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
SCRATCH="${AVR_REF_SCRATCH:-/tmp/avr_ref}"
JOBS="${JOBS:-$(nproc)}"
mkdir -p "$OUT"
# the CPU restatement (needs no reference sources)
${CXX:-g++} -std=c++17 -O2 -o "$OUT/avr_oracle" "$HERE/avr_oracle.cpp"
[ -d "$REF/ffmpeg/libavcodec" ] || { echo "reference not present at $REF; keeping prebuilt $OUT" >&2; exit 0; }
mkdir -p "$OUT" "$SCRATCH/ffbuild"

if [ ! -f "$SCRATCH/ffbuild/libavcodec/libavcodec.a" ]; then
  ( cd "$SCRATCH/ffbuild"
    "$REF/ffmpeg/configure" --disable-yasm --disable-programs --disable-doc --disable-inline-asm \
      --disable-everything --enable-decoder=h264 --enable-parser=h264 \
      --enable-demuxer=mov,h264,matroska,mpegts --enable-protocol=file \
      --disable-avdevice --disable-postproc --disable-network --disable-avfilter --disable-swscale \
      > configure.log 2>&1
    sed -i 's/#define HAVE_ARC4RANDOM 1/#define HAVE_ARC4RANDOM 0/' config.h
    make -j"$JOBS" > make.log 2>&1 )
fi
FB="$SCRATCH/ffbuild"
INC="-I$HERE/shim -I$REF -I$REF/ffmpeg -I$FB"
LIBS="-L$FB/libavformat -lavformat -L$FB/libavcodec -lavcodec -L$FB/libswresample -lswresample -L$FB/libavutil -lavutil -lm -lpthread -lz"
CXX="${CXX:-g++}"
FLAGS="-std=c++17 -O2 -DNDEBUG -w"

$CXX $FLAGS $INC -c "$REF/recode.cpp" -o "$SCRATCH/recode.o"
$CXX $FLAGS $INC -c "$REF/test.cpp" -o "$SCRATCH/test.o"
$CXX "$SCRATCH/recode.o" "$SCRATCH/test.o" -o "$OUT/recode" $LIBS
$CXX -std=c++17 -O2 -w $INC "$REF/test/arithmetic_code.cpp" -o "$OUT/arith_test" $LIBS
if [ -f "$HERE/ref_harness.cpp" ]; then
  $CXX $FLAGS $INC -DAVR_REF_DIR="\"$REF\"" "$HERE/ref_harness.cpp" "$SCRATCH/test.o" -o "$OUT/ref_harness" $LIBS
fi
# oracle/gpu_driver.cpp: the reference's av_decoder<> + hooks driving the C ABI's recorder sinks and the GPU (lower-boundary test)
PKGDIR="$HERE/../avrecode-ms_b200"
if [ -f "$HERE/gpu_driver.cpp" ] && [ -f "$PKGDIR/libavrecode_b200.so" ]; then
  $CXX $FLAGS $INC "$HERE/gpu_driver.cpp" "$SCRATCH/test.o" -o "$OUT/gpu_driver" $LIBS -L"$PKGDIR" -lavrecode_b200 \
    -Wl,-rpath,'$ORIGIN/../../avrecode-ms_b200'
fi
echo "built: $(ls "$OUT")"
