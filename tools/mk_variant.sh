#!/bin/bash
# mk_variant.sh <name> [EXTRA nvcc flags]: builds the working tree's library + avrbench into ab_<name>/ (git-ignored; travels to the GPU box) for tools/gpu_ab.sh
set -e
R=/root/repo; N=$1; shift
touch $R/avrecode-ms_b200/csrc/avr_kernels.cu
make -j8 -C $R/avrecode-ms_b200/csrc EXTRA="$*" > /dev/null
mkdir -p $R/ab_$N/bin
cp $R/avrecode-ms_b200/libavrecode_b200.so $R/ab_$N/
cp $R/avrecode-ms_b200/bin/avrbench $R/ab_$N/bin/
echo "ab_$N ready"
