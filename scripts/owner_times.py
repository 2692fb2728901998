#!/usr/bin/env python3
"""Owner-side chain of the expert-parallel layer (nf4moe_expert_rows: bucket -> permute -> decode kernel) on ONE GPU with a
synthetic receive buffer: `world` source chunks of `cap` slots, `rows` valid rows in total over `e_local` experts.  Compares the
worst-case capacity layout (mostly empty slots) with a compact one."""
import argparse, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc
p = argparse.ArgumentParser(); p.add_argument("--world", type=int, default=8); p.add_argument("--tokens", type=int, default=64)
p.add_argument("--experts", type=int, default=32); a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
dev = torch.device("cuda:0"); H = I = 2880; k, L = 4, 12
W, T = a.world, a.tokens
El = a.experts // W
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
stack = [(pkg.NF4ExpertWeights.quantize((torch.randn(El, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2),
          pkg.NF4ExpertWeights.quantize((torch.randn(El, H, I, generator=g, device=dev) * 0.02).half(), code2)) for _ in range(L)]
rows_total = T * k            # what arrives at one rank on average: W ranks x T*k/W pairs each
for cap, label in ((T * k, "worst-case capacity"), ((rows_total // W + 15) // 16 * 16, "compact")):
    n = W * cap
    rows = torch.randn(n, H, generator=g, device=dev).half()
    ids = torch.full((n,), -1, dtype=torch.int32, device=dev)
    per = rows_total // W
    for q in range(W):
        ids[q * cap:q * cap + per] = torch.randint(0, El, (per,), generator=g, device=dev, dtype=torch.int32)
    y = torch.empty(n, H, dtype=torch.float16, device=dev)
    def run():
        for w13, w2 in stack:
            pkg.ops.expert_rows(rows, ids, w13, w2, y, nested=True, rows_hint=rows_total)
    run(); torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph(); side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(gr, stream=side):
            run()
    torch.cuda.synchronize()
    ts = []
    for rep in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
        if rep: ts.append(e0.elapsed_time(e1) * 1e3 / L)
    print(f"world {W}, {El} local experts, {rows_total} rows in {n} slots ({label}): {min(ts):6.1f} us per layer, launches {pkg.ops.last_launch_count()}", flush=True)
