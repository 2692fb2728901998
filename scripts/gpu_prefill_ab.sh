#!/bin/bash
# Prefill kernel variants (build/lib_<name>.so) against the shipped library: T=8192 full-size layer, exact + fast operand, with the
# sampled-token check against the decode kernels.  Each run under its own timeout (cluster launches).
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-pf}; shift
PP="python scripts/prof_prefill.py --tokens 8192 --check --absmax nested"
for r in gemm gemm_fast; do timeout 150 $PP --regime $r > $O/${TAG}_shipped_$r.txt 2>&1; echo "shipped $r rc=$? $(grep 'T=' $O/${TAG}_shipped_$r.txt | cut -c1-260)"; done
for n in "$@"; do
  for r in gemm gemm_fast; do timeout 150 $PP --regime $r --lib build/lib_$n.so > $O/${TAG}_${n}_$r.txt 2>&1; echo "$n $r rc=$? $(grep 'T=\|check\|Error' $O/${TAG}_${n}_$r.txt | cut -c1-260 | tr '\n' ' ')"; done
done
