#!/bin/bash
set -e
cd "$(dirname "$0")/../.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Iinclude \
     -Ifinite-difference_b200/csrc -o /tmp/bulk_lab profiles/lab/bulk_lab.cu
timeout 120 /tmp/bulk_lab > /dev/null && timeout 300 ncu --set full --clock-control none --import-source on -k regex:bulk_kernel -s 12 -c 1 -o gpurun_out/prof_bulk /tmp/bulk_lab > /dev/null
