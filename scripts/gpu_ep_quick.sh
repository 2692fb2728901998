#!/bin/bash
# One expert-parallel bench run at N GPUs (peer-memory exchange) + the virtual-rank EP tests on rank 0's GPU.
set -u
N=${1:-2}; TAG=${2:-epq$N}; shift; shift
O=gpurun_out; mkdir -p $O
export NCCL_DEBUG=WARN
timeout 300 python -m pytest tests/test_gpu_ep.py -m gpu -x -q 2>&1 | tail -2
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 "$@" > $O/${TAG}.json 2> $O/${TAG}.err; echo "rc=$?"
grep -v "Warning\|warn\|OMP_NUM\|^\*\*\*\|^$\|symm_mem" $O/${TAG}.err | tail -n 4 | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('$O/${TAG}.json').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus','gpu_launches','ep_matches_single_gpu')}, d['config']['per_layer_us'], d['roofline']['avg_launch_us'])
    print('e2e', d['e2e'].get('value'))
    for k in ('strong_256','skewed_routing','prefill_4096'):
        v=d.get(k); print(k, {kk:vv for kk,vv in v.items() if kk not in ('note','routing')} if isinstance(v,dict) else v)
except Exception as e: print('parse failed', e)
PY
