#!/bin/bash
# Round 2, first GPU call: GPU tests, then CUDA-event kernel timings of the decode tcgen05 kernel in the shipped dequant-pool
# shape and in the variant builds under build/ (python gpt-oss-vllm_b200/build.py -DNF4MOE_DEC_GROUPS=g -DNF4MOE_DEC_SPLIT=s --out=...).
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2a}
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
timeout 1700 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/${TAG}_pytest.log
python -c "import bitsandbytes" > $O/${TAG}_bnb_probe.txt 2>&1; echo "bnb import rc=$?" >> $O/${TAG}_bnb_probe.txt; tail -n 2 $O/${TAG}_bnb_probe.txt
timeout 400 python scripts/kernel_times.py --tokens 16 64 128 256 --absmax nested flat --regime gemm_fast > $O/${TAG}_kt_fast.txt 2>&1; cat $O/${TAG}_kt_fast.txt | grep -v "^#"
timeout 300 python scripts/kernel_times.py --tokens 64 --absmax nested flat --regime gemm > $O/${TAG}_kt_exact.txt 2>&1; cat $O/${TAG}_kt_exact.txt | grep -v "^#"
for V in g3s1 g2s2 g2s1 g5s1; do
  timeout 300 python scripts/kernel_times.py --tokens 16 64 --absmax nested --regime gemm_fast --lib build/libnf4moe_$V.so --tag "$V " > $O/${TAG}_kt_$V.txt 2>&1; grep -v "^#" $O/${TAG}_kt_$V.txt
done
timeout 600 python bench.py --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}); print('e2e', d['e2e']['value'], d['e2e']['eager']['value']); r=d['roofline']; print({k:r[k] for k in ('achieved','frac','avg_launch_us')}, r['w2_kernel'], r['whole_step']); print(d['prefill']['exact'], d['prefill']['fast'])
except Exception as e: print('bench parse failed', e)
PY
