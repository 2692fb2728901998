set -u
O=gpurun_out; mkdir -p $O
for C in 0 1; do
NF4MOE_EP_COALESCE=$C timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2953$C bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline > $O/s15_ep2_c$C.json 2> $O/s15_ep2_c$C.err; echo "coalesce=$C rc=$?"
tail -n 1 $O/s15_ep2_c$C.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['n_gpus'], round(d['value']), d['ms_per_step'], d['config']['cuda_graph'], round(d['e2e']['value']))"
tail -n 3 $O/s15_ep2_c$C.err
done
