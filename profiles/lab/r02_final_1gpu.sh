#!/bin/bash
# round 2, one GPU: the whole -m gpu suite, the driver's bench command, the launch list of the same command under
# ncu (shares only -- a time taken under ncu is never a bench value), a short run of the reference arm, smoke().
mkdir -p gpurun_out
(time python -m pytest tests -m gpu -x -q) > gpurun_out/r02_gputest_final.log 2>&1; tail -3 gpurun_out/r02_gputest_final.log
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r02_bench_k20.json 2> gpurun_out/r02_bench_k20.err; cut -c1-300 gpurun_out/r02_bench_k20.json
python bench.py --gpus 1 --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_k100.json 2> gpurun_out/r02_bench_k100.err; cut -c1-300 gpurun_out/r02_bench_k100.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_bench_c3.csv \
  python bench.py --steps 20 --warmup 5 --no-workloads --no-history-legs --no-cpu-baseline > gpurun_out/ncu_bench.log 2>&1
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r02_reference_arm_short.json 2>&1; cut -c1-400 gpurun_out/r02_reference_arm_short.json
python -c "import __graft_entry__ as g; g.smoke()"
python profiles/host_overhead.py
