#!/bin/bash
# Same-box A/B of decode-kernel build variants (build/lib_<name>.so, made with gpt-oss-vllm_b200/build.py -D... --out=...):
# whole-layer times at the given token counts; the output fingerprint (fp) must not change between FAST-operand variants.
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-ab}; shift
TOK=${TOKENS:-"64"}
LT="python scripts/layer_times.py --tokens $TOK --reps 8"
timeout 200 $LT --tag "shipped " > $O/${TAG}_shipped.txt 2>&1; cat $O/${TAG}_shipped.txt
for n in "$@"; do
  timeout 100 python scripts/debug_small.py build/lib_$n.so > $O/${TAG}_${n}_debug.txt 2>&1; echo "$n debug rc=$? $(grep BAD $O/${TAG}_${n}_debug.txt | head -2)"
  timeout 200 $LT --lib build/lib_$n.so --tag "$n " > $O/${TAG}_$n.txt 2>&1; cat $O/${TAG}_$n.txt
done
