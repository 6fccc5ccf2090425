#!/bin/bash
# ncu capture of one star-kernel configuration: bash profiles/lab/run_one.sh <bx> <by> <chunk> <out-name>
set -e
cd "$(dirname "$0")/../.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Iinclude \
     -Ifinite-difference_b200/csrc -Iprofiles/lab -o /tmp/star_one profiles/lab/star_one.cu
/tmp/star_one $1 $2 $3 && ncu --set full --clock-control none --import-source on -s 5 -c 1 -o gpurun_out/$4 /tmp/star_one $1 $2 $3 > /dev/null
