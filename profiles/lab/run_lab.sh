#!/bin/bash
# Build and run the star-kernel tuning lab on the GPU box: bash profiles/lab/run_lab.sh
set -e
cd "$(dirname "$0")/../.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Iinclude \
     -Ifinite-difference_b200/csrc -o /tmp/star_lab profiles/lab/star_lab.cu
/tmp/star_lab
