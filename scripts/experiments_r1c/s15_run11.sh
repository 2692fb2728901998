set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/s15_pytest11.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/s15_pytest11.log
for P in 1 0; do for T in 1 8; do NF4MOE_PDL=$P timeout 300 python bench.py --tokens $T --steps 20 --warmup 5 --no-cpu-baseline --no-prefill 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PDL=$P', $T, d['config']['regime_resolved'], round(d['value'],1), round(d['config']['per_layer_us'],1), round(d['e2e']['value'],1), round(d['roofline']['whole_step']['achieved'],1))"; done; done
