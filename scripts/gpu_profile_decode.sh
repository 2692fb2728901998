#!/bin/bash
# Round evidence for the decode path in one gpurun call: the bench line, the reference arm, kernel timings, the ncu launch list of
# the bench command and one `ncu --set full` capture of the decode tcgen05 kernel (merged w13 + w2 launch).  Every ncu command
# runs only after the same command line exited 0 without ncu.  Usage: bash scripts/gpu_profile_decode.sh [tag]
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2}
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -n 3 $O/${TAG}_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2>&1; echo "reference rc=$?"
timeout 300 python scripts/kernel_times.py --tokens 16 32 64 128 256 --absmax nested flat --regime gemm_fast > $O/${TAG}_kernel_times_dec.txt 2>&1
timeout 300 python scripts/kernel_times.py --tokens 1 2 4 8 --absmax nested flat --regime gemv > $O/${TAG}_kernel_times_gemv.txt 2>&1
timeout 300 python scripts/layer_times.py --tokens 1 8 16 64 128 256 > $O/${TAG}_layer_times.txt 2>&1; cat $O/${TAG}_layer_times.txt
timeout 200 python scripts/dec_roles.py --tokens 64 > $O/${TAG}_dec_roles.txt 2>&1
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-prefill --no-sweep"
$BENCH > $O/${TAG}_bench_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/${TAG}_launches.csv $BENCH > $O/${TAG}_ncu_bench.log 2>&1
DEC="python scripts/prof_decode.py --tokens 64 --layers 1 --iters 2 --regime auto --absmax nested"
$DEC > $O/${TAG}_dec_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dec_kernel -s 1 -c 1 -f -o $O/${TAG}_dec $DEC > $O/${TAG}_ncu_dec.log 2>&1
tail -n 2 $O/${TAG}_dec_plain.log $O/${TAG}_ncu_dec.log
python - <<PY
import json
d=json.loads(open('$O/${TAG}_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','step_ms_min_max','gpu_launches')}); print(d['e2e']['value'], d['e2e']['eager']['value']); r=d['roofline']; print({k:r[k] for k in ('kernel','achieved','frac','avg_launch_us','traffic')}, r.get('split_launches'), r['whole_step']); print(d['prefill']['exact']['gemm_tflops'], d['prefill']['fast']['gemm_tflops'], d['cpu_baseline'])
PY
