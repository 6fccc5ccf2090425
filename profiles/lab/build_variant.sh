#!/bin/bash
# Build a variant of libfd_b200.so with extra nvcc flags into profiles/lab/_ab/<name>/ (A/B runs on one box:
# FD_B200_LIB=profiles/lab/_ab/<name>/libfd_b200.so python bench.py ...).  usage: build_variant.sh <name> <flags...>
set -e
name=$1; shift
root=$(cd "$(dirname "$0")/../.." && pwd)
out=$root/profiles/lab/_ab/$name
mkdir -p "$out"
pids=()
for src in "$root"/finite-difference_b200/csrc/*.cu; do
  obj=$out/$(basename "${src%.cu}").o
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC \
    -I"$root/include" -I"$root/finite-difference_b200/csrc" "$@" -c "$src" -o "$obj" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
/usr/local/cuda/bin/nvcc -shared -o "$out/libfd_b200.so" "$out"/*.o -cudart static
rm -f "$out"/*.o
echo "built $out/libfd_b200.so"
