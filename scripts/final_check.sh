#!/bin/bash
# What the driver runs at round end, in one gpurun call (use --gpus 2): GPU tests, smoke(), both bench arms at N=1 and N=2.
set -u
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/final_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/final_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1
timeout 300 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > $O/final_ref1.json 2>$O/final_ref1.err; echo "ref1 rc=$? lines=$(wc -l < $O/final_ref1.json)"
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/final_bench1.json 2>$O/final_bench1.err; echo "bench1 rc=$? lines=$(wc -l < $O/final_bench1.json)"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29591 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > $O/final_ref2.json 2>$O/final_ref2.err; echo "ref2 rc=$? lines=$(wc -l < $O/final_ref2.json)"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29592 bench.py --gpus 2 --steps 20 --warmup 5 > $O/final_bench2.json 2>$O/final_bench2.err; echo "bench2 rc=$? lines=$(wc -l < $O/final_bench2.json)"
python - <<'PY'
import json
for f in ("final_ref1", "final_bench1", "final_ref2", "final_bench2"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, d.get("impl", "ours"), d["n_gpus"], round(d["value"], 3), round(d["e2e"]["value"], 3), d.get("gpu_launches"), (d.get("roofline") or {}).get("frac"))
PY
