#!/usr/bin/env python3
"""CUDA-event time of whole MoE layers (route -> bucket -> permute -> experts -> combine) replayed as one CUDA graph over a stack
of distinct layers (weights never in L2).  --lib PATH times another build (e.g. build/libnf4moe_exp.so, where
NF4MOE_DEC_MERGED=0 keeps w13 and w2 in separate launches)."""
import argparse, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import nf4_oracle as orc  # code2 table (data) only
p = argparse.ArgumentParser()
p.add_argument("--tokens", type=int, nargs="+", default=[64])
p.add_argument("--layers", type=int, default=12)
p.add_argument("--reps", type=int, default=5)
p.add_argument("--absmax", nargs="+", default=["nested", "flat"])
p.add_argument("--regime", default="auto", choices=["auto", "gemm", "gemm_fast"])
p.add_argument("--lib", default=None)
p.add_argument("--tag", default="")
a = p.parse_args()
pkg = importlib.import_module("gpt-oss-vllm_b200")
if a.lib:
    pkg.load_library(path=os.path.abspath(a.lib))
dev = torch.device("cuda:0")
H = I = 2880; E, k = 32, 4
g = torch.Generator(device=dev).manual_seed(1)
code2 = torch.from_numpy(orc.create_dynamic_map())
stack = [(pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2),
          pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=dev) * 0.02).half(), code2)) for _ in range(a.layers)]
regime = {"auto": pkg.REGIME_AUTO, "gemm": pkg.REGIME_GEMM, "gemm_fast": pkg.REGIME_GEMM_FAST}[a.regime]
for T in a.tokens:
    x = torch.randn(T, H, generator=g, device=dev).half()
    lg = torch.randn(T, E, generator=g, device=dev).half()
    for mode in a.absmax:
        nested = mode == "nested"
        wss = [pkg.ops.new_workspace(dev, T, E, H, I, k) for _ in stack]
        outs = [torch.empty(T, H, dtype=torch.float16, device=dev) for _ in stack]
        def run():
            for (w13, w2), ws, o in zip(stack, wss, outs):
                pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, regime=regime, nested=nested, workspace=ws, out=o)
        run(); torch.cuda.synchronize()
        ref = outs[0].clone()
        gr = torch.cuda.CUDAGraph(); side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            with torch.cuda.graph(gr, stream=side):
                run()
        torch.cuda.synchronize()
        ts = []
        for rep in range(a.reps + 1):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); gr.replay(); e1.record(); torch.cuda.synchronize()
            if rep: ts.append(e0.elapsed_time(e1) * 1e3 / len(stack))
        st = max(pkg.ops.workspace_status(w) for w in wss)
        fp = int(outs[0].view(torch.int16).to(torch.int64).square().sum().item()) & 0xFFFFFFFFFFFF
        print(f"{a.tag}T={T:4d} {mode:6s} {a.regime}: {sum(ts) / len(ts):7.1f} us/layer (min {min(ts):6.1f})  {T / (sum(ts) / len(ts) * 24) * 1e6:8.0f} tok/s at 24 layers  "
              f"status {st} same_as_eager {torch.equal(ref, outs[0])} fp {fp:012x}", flush=True)
