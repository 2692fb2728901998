#!/bin/bash
# BASELINE.json configs[3] (20b shape, decode batch 256 global, EP over N GPUs) and configs[4] decode leg (120b shape: 128 experts,
# 36 layers, decode 256 global) on N GPUs of one box.  Usage: bash scripts/ep_configs.sh N
set -u
N=${1:-8}
O=gpurun_out; mkdir -p $O
TOK=$((256 / N))
run() {  # name, extra args
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $2 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline ${@:3} > $O/$1.json 2> $O/$1.err; echo "$1 rc=$?"
  tail -n 1 $O/$1.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); c=d['config']; print(d['n_gpus'], c['experts'], c['layers'], c['tokens_per_rank'], round(d['value']), round(d['ms_per_step'],3), c['regime'], c['capacity_rows_per_peer'], round(d['e2e']['value']))"
  grep -v "OMP_NUM\|\*\*\*\*" $O/$1.err | tail -n 3
}
run r1c_cfg3_ep${N}_decode256 29571 --tokens $TOK
run r1c_cfg4_ep${N}_120b_decode256 29572 --tokens $TOK --experts 128 --layers 36
