#!/usr/bin/env python3
"""Marginal cost of the small kernels of one MoE layer in the programmatic-dependent-launch chain: CUDA graphs of 24 repetitions of
growing prefixes (route | + bucket | + permute | + combine), CUDA events.  The expert kernels are left out on purpose."""
import argparse, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
p = argparse.ArgumentParser(); p.add_argument("--tokens", type=int, nargs="+", default=[1, 8, 64, 256]); p.add_argument("--lib", default=None); p.add_argument("--tag", default=""); a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
if a.lib:
    pkg.load_library(path=os.path.abspath(a.lib))
dev = torch.device("cuda:0"); H = 2880; E, k, L = 32, 4, 24
g = torch.Generator(device=dev).manual_seed(1)
for T in a.tokens:
    x = torch.randn(T, H, generator=g, device=dev).half()
    lgs = [torch.randn(T, E, generator=g, device=dev).half() for _ in range(L)]
    y = torch.randn(T * k, H, generator=g, device=dev).half()
    use_tiles = T > 8

    def chain(n):
        for l in range(L):
            ids, w = pkg.ops.route_topk(lgs[l], k)
            if n >= 2:
                b = pkg.ops.bucket(ids, E, want_tiles=use_tiles)
            if n >= 3 and use_tiles:
                pkg.ops.permute_rows(x, b, k)
            if n >= 4:
                pkg.ops.combine(y, w, ids, E)
    res = []
    for n in (1, 2, 3, 4):
        chain(n); torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph(); side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            with torch.cuda.graph(gr, stream=side):
                chain(n)
        torch.cuda.synchronize()
        ts = []
        for rep in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
            if rep: ts.append(e0.elapsed_time(e1) * 1e3 / L)
        res.append(min(ts))
    print(f"{a.tag}T={T:4d}: route {res[0]:5.2f} | +bucket {res[1]:5.2f} | +permute {res[2]:5.2f} | +combine {res[3]:5.2f}  us per layer (cumulative)", flush=True)
