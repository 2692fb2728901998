#!/bin/bash
# Expert-parallel bench at N GPUs: peer-memory exchange (default) and the NCCL exchange, each under its own timeout.
set -u
N=${1:-2}; TAG=${2:-ep$N}
O=gpurun_out; mkdir -p $O
export NCCL_DEBUG=WARN
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3"
for B in auto nccl; do
  timeout 420 $RUN --ep-backend $B > $O/${TAG}_$B.json 2> $O/${TAG}_$B.err; echo "$B rc=$?"
  tail -n 4 $O/${TAG}_$B.err | cut -c1-300
  python - <<PY
import json
try:
    d=json.loads(open('$O/${TAG}_$B.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus','gpu_launches','ep_matches_single_gpu')}, d.get('exchange'))
    print('e2e', d['e2e'].get('value'))
    for k in ('strong_256','skewed_routing','prefill_4096'): print(k, d.get(k))
except Exception as e: print('parse failed', e)
PY
done
