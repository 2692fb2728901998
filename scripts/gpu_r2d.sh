#!/bin/bash
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2d}
run() { name=$1; shift; timeout 200 "$@" > $O/${TAG}_$name.txt 2>&1; echo "$name rc=$? $(grep BAD $O/${TAG}_$name.txt)"; }
run default python scripts/debug_small.py
run default2 python scripts/debug_small.py
run exp_pdl1 python scripts/debug_small.py build/libnf4moe_exp.so
NF4MOE_PDL=0 run exp_pdl0 python scripts/debug_small.py build/libnf4moe_exp.so
run g3s1 python scripts/debug_small.py build/libnf4moe_g3s1.so
run g2s2 python scripts/debug_small.py build/libnf4moe_g2s2.so
CUDA_LAUNCH_BLOCKING=1 run blocking python scripts/debug_small.py
grep -h "E=" $O/${TAG}_exp_pdl0.txt | awk '{print $1,$2,$3,$4,$5,$6,$9,$10}' | head -40
