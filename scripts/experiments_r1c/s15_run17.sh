set -u
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/s15_pytest17.log 2>&1; echo "pytest rc=$?"; tail -n 6 $O/s15_pytest17.log
for C in 1 0; do echo "== BUCKET_CLUSTER=$C"; NF4MOE_BUCKET_CLUSTER=$C timeout 200 python scripts/prof_prefill.py --tokens 1024 8192 --absmax nested --regime gemm 2>&1 | tail -n 2; done
