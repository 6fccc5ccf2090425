#!/bin/bash
# A/B on one box: a variant build (here: an integer-pipe finite guard in front of the exact inf/NaN tests, tried and rejected) against the plain
# FP64 compares (profiles/lab/_ab/plain_checks), configs 3, 2, 4; then the parity tests that plant NaN / inf.
mkdir -p gpurun_out
show='import json,sys; d=json.loads(sys.stdin.read()); print(sys.argv[1], "%.4f ms" % d["ms_per_step"], "%.4f" % d["roofline"]["frac"])'
for rep in 1 2; do
for w in c3 c2 c4; do
  B="python bench.py --workload $w --steps 100 --warmup 5 --no-workloads --no-cpu-baseline --no-history-legs"
  $B 2>/dev/null | python -c "$show" "$w bits "
  FD_B200_LIB=$PWD/profiles/lab/_ab/plain_checks/libfd_b200.so $B 2>/dev/null | python -c "$show" "$w plain"
done
done | tee gpurun_out/r02_bits_ab.txt
python -m pytest tests/test_gpu_parity.py tests/test_gpu_random.py tests/test_gpu_bench_geometry.py -m gpu -x -q 2>&1 | tail -3
