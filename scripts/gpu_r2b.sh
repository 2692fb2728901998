#!/bin/bash
# Round 2 GPU pass: a 90-second sanity launch of the decode tcgen05 kernel first (a hang there ends the call at once), then
# CUDA-event kernel timings (shipped build + variant builds under build/), the GPU tests (per-test timeout), the bench line.
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2b}
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader
KT="python scripts/kernel_times.py"
timeout 120 $KT --tokens 16 --layers 1 --reps 2 --absmax nested --regime gemm_fast > $O/${TAG}_sanity.txt 2>&1; rc=$?
cat $O/${TAG}_sanity.txt | tail -n 3
if [ $rc -ne 0 ]; then echo "sanity rc=$rc: decode kernel broken, stopping"; exit 1; fi
timeout 300 $KT --tokens 16 64 128 256 --absmax nested flat --regime gemm_fast > $O/${TAG}_kt_fast.txt 2>&1; echo "kt_fast rc=$?"; grep -v "^#" $O/${TAG}_kt_fast.txt
timeout 200 $KT --tokens 64 --absmax nested flat --regime gemm > $O/${TAG}_kt_exact.txt 2>&1; echo "kt_exact rc=$?"; grep -v "^#" $O/${TAG}_kt_exact.txt
for V in g3s1 g2s2 g2s1 g5s1; do
  timeout 200 $KT --tokens 16 64 --absmax nested --regime gemm_fast --lib build/libnf4moe_$V.so --tag "$V " > $O/${TAG}_kt_$V.txt 2>&1; echo "$V rc=$?"; grep -v "^#" $O/${TAG}_kt_$V.txt
done
# prefill: the shipped cluster size and the variants (fingerprints in the # lines must agree)
timeout 200 $KT --tokens 1024 8192 --layers 2 --reps 3 --absmax nested --regime gemm_fast > $O/${TAG}_kt_prefill.txt 2>&1; echo "prefill rc=$?"; grep -v "^#" $O/${TAG}_kt_prefill.txt
for V in c2 c4; do
  timeout 200 $KT --tokens 1024 8192 --layers 2 --reps 3 --absmax nested --regime gemm_fast --lib build/libnf4moe_$V.so --tag "$V " > $O/${TAG}_kt_$V.txt 2>&1; echo "$V rc=$?"; grep -v "^#" $O/${TAG}_kt_$V.txt
done
timeout 150 $KT --tokens 8192 --layers 2 --reps 3 --absmax nested --regime gemm > $O/${TAG}_kt_prefill_exact.txt 2>&1; grep -v "^#" $O/${TAG}_kt_prefill_exact.txt
grep -h "fingerprint" $O/${TAG}_kt_*.txt | sort | uniq -c | sort -k3 > $O/${TAG}_fingerprints.txt
python -c "import bitsandbytes" > $O/${TAG}_bnb_probe.txt 2>&1; echo "bnb import rc=$?" >> $O/${TAG}_bnb_probe.txt
timeout 1300 python -m pytest tests -m gpu -x -q --timeout=400 --timeout-method=thread --durations=15 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 25 $O/${TAG}_pytest.log
timeout 800 python bench.py --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -n 5 $O/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}); print('e2e', d['e2e']['value']); r=d['roofline']; print({k:r.get(k) for k in ('achieved','frac','avg_launch_us')})
    for k in ('decode_sweep','exact_operand_decode','flat_absmax_decode','strong_256','parity_of_benched_output','prefill'): print(k, d.get(k))
except Exception as e: print('bench parse failed', e)
PY
