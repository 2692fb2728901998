set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/s15_pytest10.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/s15_pytest10.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-prefill > $O/s15_bench10.json 2> $O/s15_bench10.err; echo "bench rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/s15_bench10.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','step_ms_min_max','launches_per_layer')}, d['e2e']['value'], d['roofline']['avg_launch_us'], d['roofline']['w2_kernel']['avg_launch_us'])
PY
for T in 1 8 16 32; do timeout 300 python bench.py --tokens $T --steps 20 --warmup 5 --no-cpu-baseline --no-prefill 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print($T, d['config']['regime_resolved'], round(d['value'],1), round(d['config']['per_layer_us'],1), round(d['e2e']['value'],1), round(d['roofline']['whole_step']['achieved'],1))"; done
