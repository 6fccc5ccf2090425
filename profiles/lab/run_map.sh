#!/bin/bash
set -e
cd "$(dirname "$0")/../.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Iinclude \
     -Ifinite-difference_b200/csrc -o /tmp/map_lab profiles/lab/map_lab.cu
timeout 120 /tmp/map_lab
