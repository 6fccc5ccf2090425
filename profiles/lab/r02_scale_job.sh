#!/bin/bash
# round 2: the driver's scaling command (K=20, W=5) at N = 1, 2, 4, 8 on one 8-GPU box, with per-step timelines;
# N=8 again at K=100; strong scaling of the one 512^3 grid; N=8 with NCCL send/recv instead of peer memory.
mkdir -p gpurun_out
run() {  # run N tag args...
  local N=$1 tag=$2; shift 2
  if [ "$N" = 1 ]; then
    python bench.py --gpus 1 "$@" > gpurun_out/r02_scale_${tag}.json 2> gpurun_out/r02_scale_${tag}.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --gpus $N "$@" > gpurun_out/r02_scale_${tag}.json 2> gpurun_out/r02_scale_${tag}.err
  fi
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/r02_scale_{tag}.json").read().strip().splitlines()[-1])
    w = d.get("workloads", {}).get("c5", {})
    print(tag, "ms/step %.4f" % d["ms_per_step"], "value %.4g" % d["value"], "e2e_s %.4f" % d["e2e"]["seconds"],
          "host %.3f" % d["host_enqueue_ms_per_step"], "launches", d["gpu_launches"], "parity", d.get("parity_check", {}).get("bit_exact"),
          "c5 %.4f" % w.get("ms_per_step", float("nan")), "same-run", d.get("single_gpu_same_run"), "clocks", d["clocks"]["samples"], d["clocks"]["reasons"])
except Exception as exc:
    print(tag, "FAILED", exc)
    print(open(f"gpurun_out/r02_scale_{tag}.err").read()[-1500:])
PY
}
for N in 1 2 4 8; do
  run $N n${N}_k20 --steps 20 --warmup 5 --no-cpu-baseline --no-history-legs --timeline gpurun_out/r02_timeline_n${N}_k20.json
done
run 8 n8_k100 --steps 100 --warmup 5 --timeline gpurun_out/r02_timeline_n8_k100.json
for N in 2 4 8; do
  run $N strong_n${N}_k100 --steps 100 --warmup 5 --scaling strong --no-workloads --timeline gpurun_out/r02_timeline_strong_n${N}_k100.json
done
FD_B200_NO_P2P=1 run 8 n8_k20_nccl --steps 20 --warmup 5 --no-workloads
FD_B200_NO_SEQ=1 run 8 n8_k20_noseq --steps 20 --warmup 5 --no-workloads
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 profiles/pcie_probe.py 2>/dev/null | tee gpurun_out/r02_pcie_probe_n8.txt
