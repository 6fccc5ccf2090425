#!/bin/bash
# Weak-scaling runs on one multi-GPU box: bash profiles/run_scaling.sh "2 4 8"  (writes gpurun_out/scale_*.json)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in c3 c5; do
  python bench.py --workload $w --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/scale_${w}_n1.json 2>/dev/null
  for n in ${1:-2 4 8}; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n \
      bench.py --gpus $n --workload $w --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_${w}_n$n.json
  done
done
if [ -n "$2" ]; then   # same run with NCCL isend/irecv instead of peer memory
FD_B200_NO_P2P=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 29611 \
  bench.py --gpus $2 --workload c3 --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | tail -1 > gpurun_out/scale_c3_n$2_nccl.json
fi
for f in gpurun_out/scale_*.json; do python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print('$f'.split('/')[-1], d['n_gpus'], round(d['ms_per_step'],4), '%.4g'%d['value'], round(d['roofline']['frac'],4), 'e2e %.3g'%d['e2e']['value'], d['config'].get('engine_stats'))
"; done
