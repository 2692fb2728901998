#!/usr/bin/env python3
"""Small driver for ncu: a few decode-regime MoE layer forwards at the BASELINE shapes (used to capture profiles/)."""
import argparse
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc  # code2 table (data) only

p = argparse.ArgumentParser()
p.add_argument("--tokens", type=int, default=64)
p.add_argument("--layers", type=int, default=2)
p.add_argument("--iters", type=int, default=3)
p.add_argument("--experts", type=int, default=32)
p.add_argument("--absmax", default="nested")
p.add_argument("--regime", default="gemv")
a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
dev = torch.device("cuda:0")
H = I = 2880
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
stack = []
for _ in range(a.layers):
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(a.experts, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(a.experts, H, I, generator=g, device=dev) * 0.02).half(), code2)
    stack.append((w13, w2))
x = torch.randn(a.tokens, H, generator=g, device=dev).half()
lg = torch.randn(a.tokens, a.experts, generator=g, device=dev).half()
regime = {"gemv": pkg.REGIME_GEMV, "gemm": pkg.REGIME_GEMM, "gemm_fast": pkg.REGIME_GEMM_FAST, "auto": pkg.REGIME_AUTO}[a.regime]
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(a.iters):
    if it == a.iters - 1:
        ev0.record()
    for w13, w2 in stack:
        out = pkg.ops.moe_forward(x, w13, w2, 4, router_logits=lg, nested=(a.absmax == "nested"), regime=regime)
    if it == a.iters - 1:
        ev1.record()
torch.cuda.synchronize()
print(f"T={a.tokens} layers={a.layers} absmax={a.absmax}: {ev0.elapsed_time(ev1) / a.layers * 1e3:.1f} us/layer, finite={bool(torch.isfinite(out).all())}")
