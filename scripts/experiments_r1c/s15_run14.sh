set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/s15_pytest14.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/s15_pytest14.log
for N in 2 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$N bench.py --gpus $N --steps 20 --warmup 5 > $O/r1c_bench_ep$N.json 2> $O/r1c_bench_ep$N.err; echo "ep$N rc=$?"
tail -n 1 $O/r1c_bench_ep$N.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value']), d['ms_per_step'], d['config']['regime'], d['config']['capacity_rows_per_peer'], round(d['e2e']['value']), round(d['e2e']['eager']['value']), d['roofline']['avg_launch_us'])"
grep -v "OMP_NUM\|\*\*\*\*" $O/r1c_bench_ep$N.err | tail -n 3
done
