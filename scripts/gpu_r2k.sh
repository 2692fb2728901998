#!/bin/bash
# Round-2 checkpoint: bitsandbytes probe, GPU test suite, smoke, then the decode evidence bundle (bench line, reference arm,
# kernel/layer timings, ncu launch list + one full capture).  Usage: bash scripts/gpu_r2k.sh [tag]
set -u
O=gpurun_out; mkdir -p $O
TAG=${1:-r2k}
python -c "import bitsandbytes as b; print('bitsandbytes', b.__version__)" > $O/${TAG}_bnb_probe.txt 2>&1; tail -n 1 $O/${TAG}_bnb_probe.txt
timeout 100 python scripts/debug_small.py > $O/${TAG}_debug.txt 2>&1; rc=$?; echo "debug rc=$rc $(grep BAD $O/${TAG}_debug.txt | head -3)"
timeout 1500 python -m pytest tests -m gpu -x -q --timeout=400 --timeout-method=thread > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -n 4 $O/${TAG}_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.txt 2>&1; echo "smoke rc=$?"; tail -n 1 $O/${TAG}_smoke.txt
bash scripts/gpu_profile_decode.sh $TAG
