#!/bin/bash
# Build and run the TMA-bulk staging lab on the GPU box: bash profiles/lab/run_bulk.sh
set -e
cd "$(dirname "$0")/../.."
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Iinclude \
     -Ifinite-difference_b200/csrc -o /tmp/bulk_lab profiles/lab/bulk_lab.cu
timeout 120 /tmp/bulk_lab
