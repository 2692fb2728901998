#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2i}
KT="python scripts/kernel_times.py"
timeout 100 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; echo "debug rc=$? $(grep BAD $O/${TAG}_debug.txt)"
timeout 300 $KT --tokens 16 64 128 256 --absmax nested flat --regime gemm_fast > $O/${TAG}_kt_fast.txt 2>&1; echo "kt_fast rc=$?"; grep -v "^#" $O/${TAG}_kt_fast.txt
timeout 200 $KT --tokens 64 --absmax nested flat --regime gemm > $O/${TAG}_kt_exact.txt 2>&1; echo "kt_exact rc=$?"; grep -v "^#" $O/${TAG}_kt_exact.txt
timeout 200 python scripts/dec_roles.py --tokens 64 > $O/${TAG}_roles.txt 2>&1; echo "roles rc=$?"; cat $O/${TAG}_roles.txt
timeout 200 python scripts/dec_roles.py --tokens 64 --absmax flat > $O/${TAG}_roles_flat.txt 2>&1; cat $O/${TAG}_roles_flat.txt
grep -h "fingerprint" $O/${TAG}_kt_*.txt | sed 's/^# [a-z0-9 ]*T=/T=/' | sort | uniq -c > $O/${TAG}_fingerprints.txt; cat $O/${TAG}_fingerprints.txt
