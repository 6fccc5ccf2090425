#!/bin/bash
# round 2: native slab-step sequencer + one-launch exchange on 4 GPUs: multi-GPU parity tests, weak K=20,
# strong K=100, A/B against the Python-issued sequence (FD_B200_NO_SEQ=1).
mkdir -p gpurun_out
(time python -m pytest tests/test_gpu_dist.py -m gpu -x -q) > gpurun_out/r02_gpu_dist_tests_seq.log 2>&1
tail -3 gpurun_out/r02_gpu_dist_tests_seq.log
run() {  # run N tag args...
  local N=$1 tag=$2; shift 2
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N "$@" > gpurun_out/r02_seq_${tag}.json 2> gpurun_out/r02_seq_${tag}.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/r02_seq_{tag}.json").read().strip().splitlines()[-1])
    print(tag, "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "e2e_s %.4f" % d["e2e"]["seconds"],
          "host %.3f" % d["host_enqueue_ms_per_step"], "launches/step", d["gpu_launches"] / d["steps"], "parity", d.get("parity_check", {}).get("bit_exact"),
          d["config"]["engine_stats"])
except Exception as exc:
    print(tag, "FAILED", exc)
    print(open(f"gpurun_out/r02_seq_{tag}.err").read()[-1500:])
PY
}
run 4 n4_k20 --steps 20 --warmup 5 --no-workloads
FD_B200_NO_SEQ=1 run 4 n4_k20_noseq --steps 20 --warmup 5 --no-workloads
run 4 strong_n4_k100 --steps 100 --warmup 5 --scaling strong --no-workloads
FD_B200_NO_SEQ=1 run 4 strong_n4_k100_noseq --steps 100 --warmup 5 --scaling strong --no-workloads
run 2 strong_n2_k100 --steps 100 --warmup 5 --scaling strong --no-workloads
run 4 c5_n4_k20 --steps 20 --warmup 5 --workload c5 --no-workloads
