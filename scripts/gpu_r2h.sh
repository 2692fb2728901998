#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2h}
KT="python scripts/kernel_times.py"
timeout 100 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; echo "debug rc=$? $(grep BAD $O/${TAG}_debug.txt)"
timeout 300 $KT --tokens 16 64 128 256 --absmax nested flat --regime gemm_fast > $O/${TAG}_kt_fast.txt 2>&1; echo "kt_fast rc=$?"; grep -v "^#" $O/${TAG}_kt_fast.txt
timeout 200 $KT --tokens 64 --absmax nested flat --regime gemm > $O/${TAG}_kt_exact.txt 2>&1; echo "kt_exact rc=$?"; grep -v "^#" $O/${TAG}_kt_exact.txt
for V in g4s1 g2s2; do
  timeout 200 $KT --tokens 16 64 256 --absmax nested flat --regime gemm_fast --lib build/libnf4moe_$V.so --tag "$V " > $O/${TAG}_kt_$V.txt 2>&1; echo "$V rc=$?"; grep -v "^#" $O/${TAG}_kt_$V.txt
done
for R in 3 6; do for B in 3 6; do
  NF4MOE_DEC_MAXRAW=$R NF4MOE_DEC_MAXB=$B timeout 200 $KT --tokens 64 256 --absmax nested --regime gemm_fast --lib build/libnf4moe_exp.so --tag "raw$R b$B " > $O/${TAG}_kt_r${R}b${B}.txt 2>&1; grep -v "^#" $O/${TAG}_kt_r${R}b${B}.txt
done; done
grep -h "fingerprint" $O/${TAG}_kt_*.txt | sed 's/^# [a-z0-9 ]*T=/T=/' | sort | uniq -c > $O/${TAG}_fingerprints.txt; cat $O/${TAG}_fingerprints.txt
timeout 1300 python -m pytest tests -m gpu -x -q --timeout=400 --timeout-method=thread --durations=8 > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 14 $O/${TAG}_pytest.log
timeout 800 python bench.py --no-cpu-baseline > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -n 5 $O/${TAG}_bench.err
python - <<PY
import json
try:
    d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
    print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}); print('e2e', d['e2e']['value']); r=d['roofline']; print({k:r.get(k) for k in ('achieved','frac','avg_launch_us')})
    for s in d['decode_sweep']: print({k:(round(v,2) if isinstance(v,float) else v) for k,v in s.items() if k in ('tokens','regime','per_layer_us','tokens_per_s','w13_us','w2_us','w13_frac')})
    for k in ('exact_operand_decode','flat_absmax_decode','strong_256'): print(k, {kk:(round(v,2) if isinstance(v,float) else v) for kk,v in d[k].items() if kk in ('per_layer_us','tokens_per_s','w13_us','w2_us','w13_frac')})
    print(d['prefill']['exact']['layer_ms'], d['prefill']['fast']['layer_ms'])
except Exception as e: print('bench parse failed', e)
PY
