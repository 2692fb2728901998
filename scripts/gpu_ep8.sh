#!/bin/bash
set -u
N=${1:-8}; TAG=${2:-ep$N}
O=gpurun_out; mkdir -p $O
export NCCL_DEBUG=WARN
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3"
show() { python - <<PY
import json
try:
    d=json.loads(open('$1').read().strip().splitlines()[-1])
    print({k:d.get(k) for k in ('value','ms_per_step','n_gpus','gpu_launches','ep_matches_single_gpu')}, d.get('config',{}).get('workload'))
    print('e2e', d['e2e'].get('value'), 'roofline', {k:d['roofline'].get(k) for k in ('achieved','frac')})
    for k in ('strong_256','skewed_routing','prefill_4096'):
        v=d.get(k); print(k, {kk:vv for kk,vv in v.items() if kk!='note'} if isinstance(v,dict) else v)
except Exception as e: print('parse failed', e)
PY
}
timeout 420 $RUN > $O/${TAG}_auto.json 2> $O/${TAG}_auto.err; echo "auto rc=$?"; grep -v "Warning\|warn\|OMP_NUM\|^\*\*\*\|^$" $O/${TAG}_auto.err | tail -n 4 | cut -c1-300; show $O/${TAG}_auto.json
timeout 420 $RUN --ep-backend nccl --no-sweep > $O/${TAG}_nccl.json 2> $O/${TAG}_nccl.err; echo "nccl rc=$?"; show $O/${TAG}_nccl.json
timeout 600 $RUN --shape 120b > $O/${TAG}_120b.json 2> $O/${TAG}_120b.err; echo "120b rc=$?"; grep -v "Warning\|warn\|OMP_NUM\|^\*\*\*\|^$" $O/${TAG}_120b.err | tail -n 4 | cut -c1-300; show $O/${TAG}_120b.json
