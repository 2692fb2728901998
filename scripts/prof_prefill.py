#!/usr/bin/env python3
"""Prefill-regime timing of one full-size gpt-oss MoE layer (tcgen05 path): per-kernel CUDA-event times, tensor-roofline
fraction, and a cross-check of sampled tokens against the decode-regime kernels."""
import argparse
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc  # code2 table (data) only

p = argparse.ArgumentParser()
p.add_argument("--tokens", type=int, nargs="+", default=[8192])
p.add_argument("--experts", type=int, default=32)
p.add_argument("--reps", type=int, default=5)
p.add_argument("--absmax", default="flat")
p.add_argument("--check", action="store_true")
p.add_argument("--regime", default="gemm", choices=["gemm", "gemm_fast"])
p.add_argument("--lib", default=None)
a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
if a.lib:
    pkg.load_library(path=os.path.abspath(a.lib))
dev = torch.device("cuda:0")
H = I = 2880
E, k = a.experts, 4
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=dev) * 0.02).half(), code2)
nested = a.absmax == "nested"
REG = pkg.REGIME_GEMM_FAST if a.regime == "gemm_fast" else pkg.REGIME_GEMM
for T in a.tokens:
    x = torch.randn(T, H, generator=g, device=dev).half()
    lg = torch.randn(T, E, generator=g, device=dev).half()
    names = ["route", "bucket", "permute", "w13", "w2", "combine"]
    acc = {n: [] for n in names}
    tot = []
    for rep in range(a.reps + 1):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(7)]
        ev[0].record()
        ids, wts = pkg.ops.route_topk(lg, k); ev[1].record()
        b = pkg.ops.bucket(ids, E); ev[2].record()
        xs = pkg.ops.permute_rows(x, b, k); ev[3].record()
        h = pkg.ops.expert_w13(xs, w13, b, T, k, nested=nested, regime=REG); ev[4].record()
        y = pkg.ops.expert_w2(h, w2, b, T, k, nested=nested, regime=REG); ev[5].record()
        out = pkg.ops.combine(y, wts, ids, E); ev[6].record()
        torch.cuda.synchronize()
        if rep:
            for i, n in enumerate(names):
                acc[n].append(ev[i].elapsed_time(ev[i + 1]) * 1e3)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out2 = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, regime=REG, nested=nested)
        e1.record()
        torch.cuda.synchronize()
        if rep:
            tot.append(e0.elapsed_time(e1) * 1e3)
    flops = 2.0 * T * k * (2 * I * H + H * I)
    t13, t2 = min(acc["w13"]), min(acc["w2"])
    layer = min(tot)
    print(f"T={T} {a.regime} {a.absmax}: " + " ".join(f"{n}={min(v):.1f}us" for n, v in acc.items()) + f" | layer {layer:.1f} us "
          f"{T / layer:.3f} Mtok/s | gemms {flops / (t13 + t2) / 1e6:.1f} TFLOP/s ({flops / (t13 + t2) / 1e6 / 2250 * 100:.1f}% of 2250) "
          f"| layer {flops / layer / 1e6:.1f} TFLOP/s", flush=True)
    assert torch.equal(out, out2)
    if a.check:
        idx = torch.randperm(T, generator=torch.Generator().manual_seed(3))[:48].to(dev)
        ref = pkg.ops.moe_forward(x[idx].contiguous(), w13, w2, k, router_logits=lg[idx].contiguous(), regime=pkg.REGIME_GEMV,
                                  nested=nested)
        d = (out[idx].float() - ref.float()).abs()
        rel = (out[idx].float() - ref.float()).norm() / ref.float().norm()
        print(f"   sampled-token check vs decode kernels: max-abs {d.max().item():.3e} rel {rel.item():.3e} finite={bool(torch.isfinite(out).all())}")
        assert d.max().item() <= 1e-2 and rel.item() <= 2e-3
