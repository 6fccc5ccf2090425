#!/bin/bash
# Build and run the pair-kernel lab on the GPU box: bash profiles/lab/run_pair.sh
set -e
cd "$(dirname "$0")/../.."
LIBDIR=$PWD/finite-difference_b200/lib
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iinclude -o /tmp/pair_lab profiles/lab/pair_lab.cu \
     -L$LIBDIR -lfd_b200 -Xlinker -rpath=$LIBDIR
timeout 200 /tmp/pair_lab 256 2048 2048 f32
timeout 100 /tmp/pair_lab 64 258 1000 f32 0.1
timeout 100 /tmp/pair_lab 130 512 512 f64
timeout 100 /tmp/pair_lab 37 63 516 f64 0.5
timeout 100 /tmp/pair_lab 66 256 512 f32 1 slab
timeout 100 /tmp/pair_lab 66 256 512 f64 0.5 slab
