set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > $O/r1c_bench_ep2.json 2> $O/r1c_bench_ep2.err; echo "ep2 rc=$?"
tail -c 2500 $O/r1c_bench_ep2.json; tail -n 5 $O/r1c_bench_ep2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > $O/r1c_bench_ep2_ref.json 2>&1; echo "ref rc=$?"; tail -c 300 $O/r1c_bench_ep2_ref.json
