#!/bin/bash
# 8 GPUs, after the per-step short cut for native slab steps: multi-GPU parity tests, strong scaling at 8 and 4 GPUs,
# weak K=20 at 8.
mkdir -p gpurun_out
(time python -m pytest tests/test_gpu_dist.py -m gpu -x -q) > gpurun_out/r02_gpu_dist_tests_final.log 2>&1
grep -E "passed|failed" gpurun_out/r02_gpu_dist_tests_final.log
run() {
  local N=$1 tag=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 \
    bench.py --gpus $N "$@" > gpurun_out/r02_fast_${tag}.json 2> gpurun_out/r02_fast_${tag}.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/r02_fast_{tag}.json").read().strip().splitlines()[-1])
    print(tag, "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "host %.3f" % d["host_enqueue_ms_per_step"],
          "launches/step", d["gpu_launches"] / d["steps"], "parity", d.get("parity_check", {}).get("bit_exact"), "clocks", d["clocks"]["samples"], d["clocks"]["reasons"])
except Exception as exc:
    print(tag, "FAILED", exc)
    print(open(f"gpurun_out/r02_fast_{tag}.err").read()[-1500:])
PY
}
run 8 strong_n8_k100 --steps 100 --warmup 5 --scaling strong --no-workloads
run 4 strong_n4_k100 --steps 100 --warmup 5 --scaling strong --no-workloads
run 8 n8_k20 --steps 20 --warmup 5 --no-workloads
