#!/usr/bin/env python3
"""Small-shape check of the tcgen05 regimes against the GEMV regime (scratch diagnostic)."""
import importlib, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
pkg = importlib.import_module("gpt-oss-vllm_b200")
if len(sys.argv) > 1:
    pkg.load_library(path=os.path.abspath(sys.argv[1]))
from test_gpu_parity import make_layer
dev = torch.device("cuda:0")
bad = 0
for (E, H, I, T) in [(32, 192, 128, 40), (32, 192, 128, 8), (4, 192, 128, 40), (32, 256, 256, 40), (32, 576, 320, 40), (32, 2880, 128, 40), (8, 128, 64, 9)]:
    st13, st2, w13, w2 = make_layer(pkg, dev, 7, E, H, I)
    rng = np.random.default_rng(5)
    x = torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(dev)
    lg = torch.from_numpy(rng.standard_normal((T, E)).astype(np.float16)).to(dev)
    ref = pkg.ops.moe_forward(x, w13, w2, 4, router_logits=lg, regime=pkg.REGIME_GEMV).float()
    for name, reg in (("gemm", pkg.REGIME_GEMM), ("gemm_fast", pkg.REGIME_GEMM_FAST)):
        for nested in (True, False):
            ws = pkg.ops.new_workspace(dev, T, E, H, I, 4)
            out = pkg.ops.moe_forward(x, w13, w2, 4, router_logits=lg, regime=reg, nested=nested, workspace=ws).float()
            torch.cuda.synchronize()
            bad += int((out - ref).abs().max().item() > 0.05)
            print(f"E={E} H={H} I={I} T={T} {name} nested={nested}: status={pkg.ops.workspace_status(ws)} |out|max={out.abs().max().item():.4f} "
                  f"|ref|max={ref.abs().max().item():.4f} maxdiff={(out - ref).abs().max().item():.5f} zero_rows={(out.abs().sum(1) == 0).sum().item()}", flush=True)
print("BAD", bad, sys.argv[1:], os.environ.get("NF4MOE_PDL"))
