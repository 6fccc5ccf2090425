#!/bin/bash
# A/B of the lean star kernel with per-value NaN/inf tests (lib built with
# -DFD_LEAN_TWO_STAGE=0, copied to lib/libfd_b200_exact.so) against the two-stage test.
set -e
cd "$(dirname "$0")/../.."
L=finite-difference_b200/lib
cp $L/libfd_b200.so /tmp/new.so; cp $L/libfd_b200_exact.so /tmp/exact.so
run() { for w in c3 c2 c4; do python bench.py --workload $w --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$1', '$w', round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['clocks'])"; done; }
for i in 1 2; do
cp /tmp/exact.so $L/libfd_b200.so; run exact
cp /tmp/new.so $L/libfd_b200.so; run twostage
done
