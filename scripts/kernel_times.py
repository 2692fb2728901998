#!/usr/bin/env python3
"""CUDA-event timing of the two expert kernels alone (w13+SwiGLU, w2) in one regime, rotating over several layers so the
126 MB L2 never holds the weights.  --lib PATH times a differently built libnf4moe (build.py --out=...; e.g. another dequant
pool shape -DNF4MOE_DEC_GROUPS=3, or an -DNF4MOE_EXPERIMENTS build, where NF4MOE_GEMV_CFG = 19 / 13 / 23 picks the GEMV shape)."""
import argparse
import importlib
import os
import sys

p = argparse.ArgumentParser()
p.add_argument("--tokens", type=int, nargs="+", default=[1, 8, 16, 32, 64])
p.add_argument("--layers", type=int, default=4)
p.add_argument("--reps", type=int, default=5)
p.add_argument("--experts", type=int, default=32)
p.add_argument("--absmax", nargs="+", default=["nested", "flat"])
p.add_argument("--cfg", type=int, nargs="+", default=[0])
p.add_argument("--regime", default="gemv", choices=["gemv", "gemm", "gemm_fast"])
p.add_argument("--lib", default=None)
p.add_argument("--tag", default="")
a = p.parse_args()
if len(a.cfg) > 1:   # the library reads the override once per process: fan out
    import subprocess
    for c in a.cfg:
        argv = [sys.executable, __file__, "--cfg", str(c), "--layers", str(a.layers), "--reps", str(a.reps), "--experts",
                str(a.experts), "--regime", a.regime, "--tokens", *map(str, a.tokens), "--absmax", *a.absmax]
        if a.lib:
            argv += ["--lib", a.lib]
        subprocess.run(argv, check=True)
    sys.exit(0)
os.environ["NF4MOE_GEMV_CFG"] = str(a.cfg[0])
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc  # code2 table (data) only

pkg = importlib.import_module("gpt-oss-vllm_b200")
if a.lib:
    pkg.load_library(path=os.path.abspath(a.lib))
dev = torch.device("cuda:0")
H = I = 2880
E = a.experts
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
stack = []
for _ in range(a.layers):
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=dev) * 0.02).half(), code2)
    stack.append((w13, w2))
for T in a.tokens:
    x = torch.randn(T, H, generator=g, device=dev).half()
    lg = torch.randn(T, E, generator=g, device=dev).half()
    ids, wts = pkg.ops.route_topk(lg, 4)
    gemm = a.regime != "gemv"
    regime = {"gemv": pkg.REGIME_GEMV, "gemm": pkg.REGIME_GEMM, "gemm_fast": pkg.REGIME_GEMM_FAST}[a.regime]
    b = pkg.ops.bucket(ids, E, want_tiles=gemm)
    xin = pkg.ops.permute_rows(x, b, 4) if gemm else x
    ea = int(b["meta"][0].item())
    print(f"# T={T}: active experts {ea}, largest per-expert count {int(b['meta'][3].item())}")
    for mode in a.absmax:
        nested = mode == "nested"
        # CUDA-graph replay: a Python/ctypes launch costs more CPU time than these kernels take on the GPU
        def run13():
            return [pkg.ops.expert_w13(xin, w13, b, T, 4, nested=nested, regime=regime) for w13, _ in stack]

        def run2(hs):
            for h, (_, w2) in zip(hs, stack):
                pkg.ops.expert_w2(h, w2, b, T, 4, nested=nested, regime=regime)
        hs = run13(); run2(hs); torch.cuda.synchronize()
        # order-independent fingerprints of layer 0's outputs: pool shapes / cluster sizes must not change a single bit
        y0 = pkg.ops.expert_w2(hs[0], stack[0][1], b, T, 4, nested=nested, regime=regime)
        fp = [int(t.contiguous().view(torch.int16).to(torch.int64).square().sum().item()) & 0xFFFFFFFFFFFF for t in (hs[0], y0)]
        print(f"# {a.tag}T={T} {mode}: fingerprint h {fp[0]:012x} y {fp[1]:012x}", flush=True)
        g13, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g13, stream=side):
                hs = run13()
            with torch.cuda.graph(g2, stream=side):
                run2(hs)
        torch.cuda.synchronize()
        t13, t2 = [], []
        for rep in range(a.reps + 1):
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record(); g13.replay(); e1.record(); g2.replay(); e2.record()
            torch.cuda.synchronize()
            if rep:
                t13.append(e0.elapsed_time(e1) * 1e3 / len(stack))
                t2.append(e1.elapsed_time(e2) * 1e3 / len(stack))
        b13, b2 = ea * stack[0][0].nbytes(nested), ea * stack[0][1].nbytes(nested)
        m13, m2 = min(t13), min(t2)
        a13, a2 = sum(t13) / len(t13), sum(t2) / len(t2)
        print(f"{a.tag}{a.regime} cfg={a.cfg[0]:2d} T={T:3d} E_act={ea:2d} {mode:6s} w13 {a13:7.1f} us (min {m13:6.1f}) {b13 / a13 / 1e3:7.1f} GB/s | "
              f"w2 {a2:7.1f} us (min {m2:6.1f}) {b2 / a2 / 1e3:7.1f} GB/s | both {(b13 + b2) / (a13 + a2) / 1e3:7.1f} GB/s", flush=True)
