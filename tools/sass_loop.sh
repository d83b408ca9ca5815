#!/bin/bash
# compile avr_kernels.cu, dump SASS of one kernel with encodings, run the static issue model on a source-line range
# usage: tools/sass_loop.sh <kernel mangled-name regex> <file> <first> <last>
cd /root/repo/avrecode-ms_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xptxas -v -cubin -o /tmp/k.cubin avr_kernels.cu 2>&1 | grep -A2 "$1" | grep "Used\|stack"
nvdisasm -g -hex -c /tmp/k.cubin > /tmp/k.sass 2>/dev/null
awk -v pat="^\\.text\\..*$1" '$0 ~ /^\.text\./ {f = ($0 ~ pat)} f' /tmp/k.sass > /tmp/one.sass
python /root/repo/tools/sass_model.py /tmp/one.sass $2 $3 $4 2>/dev/null | cut -c1-118
