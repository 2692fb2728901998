set -u
O=gpurun_out; mkdir -p $O
for N in 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 > $O/r1c_bench_ep$N.json 2> $O/r1c_bench_ep$N.err; echo "ep$N rc=$?"
tail -n 1 $O/r1c_bench_ep$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value']), d['ms_per_step'], d['config']['regime'], d['config']['capacity_rows_per_peer'], round(d['e2e']['value']), d['roofline']['avg_launch_us'])"
done
