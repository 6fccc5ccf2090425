#!/bin/bash
# round 2: row kernel (radius-3/4, 2-D) -- parity tests, config 4 timing, march-length sweep, A/B against the
# general star kernel, one ncu --set full capture.  Run from the repository root on a B200 box.
mkdir -p gpurun_out
(time python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_geometry.py -m gpu -x -q -k "wide or wave or config4 or scenario") > gpurun_out/r02_row_tests.log 2>&1
tail -5 gpurun_out/r02_row_tests.log
B="python bench.py --workload c4 --steps 50 --warmup 5 --no-workloads --no-cpu-baseline --no-history-legs"
$B > gpurun_out/r02_c4_row.json 2> gpurun_out/r02_c4_row.err
cut -c1-400 gpurun_out/r02_c4_row.json
show='import json,sys; d=json.loads(sys.stdin.read()); print(sys.argv[1], d["ms_per_step"], d["roofline"]["frac"])'
for c in 16 32 128 256; do FD_B200_STAR_CHUNK=$c $B 2>/dev/null | python -c "$show" "chunk=$c"; done | tee gpurun_out/r02_c4_row_chunks.txt
FD_B200_NO_ROW=1 $B 2>/dev/null | python -c "$show" general | tee -a gpurun_out/r02_c4_row_chunks.txt
ncu --set full --clock-control none --import-source on -k regex:star_row -c 1 -f -o gpurun_out/r02_row_c4 \
    python bench.py --workload c4 --steps 2 --warmup 3 --no-workloads --no-cpu-baseline --no-history-legs > gpurun_out/ncu_row.log 2>&1
ls -la gpurun_out/*.ncu-rep
