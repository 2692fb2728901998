set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/s15_pytest9.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/s15_pytest9.log
for P in 1 0 1 0; do
echo "== PDL=$P"
NF4MOE_PDL=$P timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-prefill > $O/s15_bench_pdl$P.json 2> $O/s15_bench_pdl$P.err; echo "bench rc=$?"; python - <<PY
import json
d=json.loads(open('gpurun_out/s15_bench_pdl$P.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','step_ms_min_max')}, d['e2e']['value'], d['roofline']['avg_launch_us'], d['roofline']['w2_kernel']['avg_launch_us'])
PY
done
