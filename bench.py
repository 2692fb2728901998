#!/usr/bin/env python3
"""bench.py -- NF4 fused-MoE expert path on B200.

    python bench.py --gpus 1 --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W     # the reference's CPU path (oracle port) on host cores
    torchrun ... bench.py --gpus N ...                        # expert-parallel, one rank per GPU

Workload at N=1 (BASELINE.json configs[1]): the 24 MoE layers of gpt-oss-20b (H=I=2880, 32 experts, top-4, NF4 bs64,
double-quant absmax) in the decode regime, T=64 tokens; one "step" = one pass of all 24 layers over the batch.
`value` = T / step-time with everything resident in HBM (CUDA-graph replay, CUDA events); `e2e` = the same through the
public `B200NF4MoEMethod.apply_legacy` call with host buffers (H2D of x + logits and D2H of the output inside the timed
region).  9.6 GB of weights are streamed per step, far more than the 126 MB L2, so no explicit L2 flush is needed.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

H = I = 2880
TOP_K = 4
SEED = 20250805
METRIC = "NF4 MoE-layer tokens/s (24-layer gpt-oss-20b MoE stack, decode)"


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=20)
    p.add_argument("--warmup", type=int, default=5)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--tokens", type=int, default=64, help="tokens per rank (decode batch)")
    p.add_argument("--layers", type=int, default=24)
    p.add_argument("--experts", type=int, default=32)
    p.add_argument("--absmax", default="nested", choices=["nested", "flat"])
    p.add_argument("--regime", default="auto", choices=["auto", "gemv", "gemm", "gemm_fast"],
                   help="auto = what the drop-in apply() uses: GEMV up to 8 tokens, decode tcgen05 kernel above")
    p.add_argument("--ep-backend", default="auto", choices=["auto", "p2p", "nccl"],
                   help="N>1 exchange: p2p = NVLink peer-memory kernels (drop-free, worst-case capacity), nccl = fixed-capacity "
                        "all-to-all, auto = p2p when torch symmetric memory can be set up")
    p.add_argument("--ep-capacity-factor", type=float, default=2.0,
                   help="N>1, nccl back end only: rows reserved per peer = factor x mean load (an overflow is detected and fatal); "
                        "0 = worst case.  The p2p back end always uses the worst case (only rows that exist cross the link)")
    p.add_argument("--no-sweep", action="store_true", help="skip the decode sweep / exact-operand / flat-absmax / T=256 extras")
    p.add_argument("--shape", default="20b", choices=["20b", "120b"],
                   help="120b = BASELINE.json configs[4]: 128 experts, 36 layers (expert-parallel runs; decode 256 + prefill 4096 legs)")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-prefill", action="store_true", help="skip the extra 8192-token prefill measurement")
    p.add_argument("--prefill-tokens", type=int, default=8192)
    p.add_argument("--cpu-tokens", type=int, default=None, help="tokens of the CPU arm's layer sample (default: --tokens, the GPU arm's batch)")
    return p.parse_args()


# ------------------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [l.split(", ") for ts, l in self.lines if t0 - 0.05 <= ts <= t1 + 0.15] or [l.split(", ") for _, l in self.lines]
        sm = [float(r[1]) for r in rows if len(r) >= 9]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            if len(r) >= 9:
                for n, v in zip(names, r[5:9]):
                    if v.strip().lower() == "active":
                        reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": float(rows[0][2]) if rows and len(rows[0]) >= 3 else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks() -> tuple[float, str]:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------------- CPU arm
def cpu_states(rng, E, rows, cols):
    """Synthetic bnb quant states made directly (random codes, plausible double-quant absmax): same bytes and arithmetic
    per weight as a quantised N(0, 0.02^2) matrix, without minutes of CPU quantisation."""
    import numpy as np
    from oracle import nf4_oracle as orc
    nb = rows * cols // 64
    n2 = -(-nb // 256)
    code2 = orc.create_dynamic_map()
    return [{"packed": rng.integers(0, 256, size=rows * cols // 2, dtype=np.uint8),
             "absmax_u8": rng.integers(0, 256, size=nb, dtype=np.uint8),
             "absmax2": (rng.random(n2, dtype=np.float32) * np.float32(0.01) + np.float32(0.005)), "code2": code2,
             "offset": np.float32(0.05), "shape": (rows, cols), "blocksize": 64} for _ in range(E)]


def cpu_layer_seconds(st13, st2, T, E, reps, warm):
    """The reference path on CPU (oracle): bnb-semantics dequant of every expert, per-expert fp16 matmul (fp32 acc),
    clamped SwiGLU, top-4 weighted sum -- one MoE layer, T tokens."""
    import numpy as np
    import torch
    from oracle import nf4_oracle as orc
    torch.set_num_threads(os.cpu_count() or 1)
    rng = np.random.default_rng(SEED)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    times = []
    for i in range(warm + reps):
        t0 = time.perf_counter()
        orc.moe_forward_from_states(x, lg, st13, st2, k=TOP_K, nested=True)
        dt = time.perf_counter() - t0
        if i >= warm:
            times.append(dt)
    return times


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The reference's files cannot be imported on
    a box without bitsandbytes, so this is the oracle port; each step is ONE layer of the 24-layer workload on
    `--cpu-tokens` tokens (bounded sample), scaled to the stack metric."""
    import numpy as np
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    E = args.experts
    rng = np.random.default_rng(SEED)
    st13, st2 = cpu_states(rng, E, 2 * I, H), cpu_states(rng, E, H, I)
    steps, warm = max(1, min(args.steps, 6)), max(0, min(args.warmup, 1))
    T = args.cpu_tokens or args.tokens           # the same batch as the GPU arm
    times = cpu_layer_seconds(st13, st2, T, E, steps, warm)
    sec = sum(times) / len(times)
    val = T / (sec * args.layers)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": steps,
            "warmup": warm, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f16 storage / f32 accumulate (CPU)", "data": "synthetic (random NF4 codes + double-quant absmax made in numpy)",
            "config": {"workload": f"gpt-oss-20b NF4 MoE stack decode, CPU oracle port: 1 of {args.layers} layers x {T} tokens per step, "
                                   f"tokens/s scaled by 1/{args.layers}", "hidden": H, "intermediate": I, "experts": E, "top_k": TOP_K,
                       "layers": args.layers, "tokens": T},
            "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": os.cpu_count(), "kind": "port",
                             "sample": f"one MoE layer, T={T}, all {E} experts dequantised (as V/.../bitsandbytes.py:588-603 does); "
                                       f"mean of {steps} after {warm} warm-up; steps capped at 6 to bound CPU time"},
            "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------- GPU arm
def make_stack(pkg, dev, layers, E, code2):
    import torch
    g = torch.Generator(device=dev).manual_seed(SEED)
    stack = []
    for _ in range(layers):
        w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
        w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=dev) * 0.02).half(), code2)
        stack.append((w13, w2))
    torch.cuda.synchronize(dev)
    return stack


def prefill_section(pkg, dev, layer, E, nested, T=8192, reps=5):
    """BASELINE.json configs[2] beside the headline: ONE gpt-oss-20b MoE layer, 8192 tokens, tcgen05 grouped-GEMM regime, both
    operand modes; CUDA events around the whole layer call and around the two GEMM kernels alone."""
    import torch
    w13, w2 = layer
    g = torch.Generator(device=dev).manual_seed(SEED + 2)
    x = torch.randn(T, H, generator=g, device=dev).half()
    lg = torch.randn(T, E, generator=g, device=dev).half()
    flops = 2.0 * T * TOP_K * (2 * I * H + H * I)
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            mp = json.load(f)
        burst, sust = float(mp["bf16_tflops"]), float(mp.get("bf16_tflops_sustained", mp["bf16_tflops"]))
    except Exception:
        burst, sust = 1590.0, 1400.0
    out = {"workload": f"one gpt-oss-20b NF4 MoE layer, T={T} tokens prefill (BASELINE.json configs[2])", "flops_per_layer": flops,
           "peak_tflops": {"nominal_dense_fp16": 2250.0, "measured_bf16_burst": burst, "measured_bf16_sustained": sust}}
    ws = pkg.ops.get_workspace(dev, T, E, H, I, TOP_K)
    for name, regime in (("exact", pkg.REGIME_GEMM), ("fast", pkg.REGIME_GEMM_FAST)):
        o = torch.empty(T, H, dtype=torch.float16, device=dev)
        tl, tg = [], []
        for rep in range(reps + 2):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            pkg.ops.moe_forward(x, w13, w2, TOP_K, router_logits=lg, out=o, workspace=ws, nested=nested, regime=regime)
            e1.record()
            torch.cuda.synchronize()
            ids, wts = pkg.ops.route_topk(lg, TOP_K)
            b = pkg.ops.bucket(ids, E)
            xs = pkg.ops.permute_rows(x, b, TOP_K)
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            hbuf = pkg.ops.expert_w13(xs, w13, b, T, TOP_K, nested=nested, regime=regime)
            pkg.ops.expert_w2(hbuf, w2, b, T, TOP_K, nested=nested, regime=regime)
            a1.record()
            torch.cuda.synchronize()
            if rep >= 2:
                tl.append(e0.elapsed_time(e1))
                tg.append(a0.elapsed_time(a1))
        ml, mg = statistics.mean(tl), statistics.mean(tg)
        out[name] = {"layer_ms": ml, "tokens_per_s": T / (ml * 1e-3), "layer_tflops": flops / (ml * 1e-3) / 1e12,
                     "gemm_kernels_ms": mg, "gemm_tflops": flops / (mg * 1e-3) / 1e12,
                     "gemm_frac_of_nominal": flops / (mg * 1e-3) / 1e12 / 2250.0, "gemm_frac_of_measured_sustained": flops / (mg * 1e-3) / 1e12 / sust}
    out["operand_modes"] = "exact = fp16_rn(fp32 code x fp32 absmax), bit-identical to bnb dequantize_4bit; fast = half2 arithmetic"
    return out


def state_to_cpu(w, e):
    return {"packed": w.packed[e].cpu().numpy().reshape(-1), "absmax_u8": w.absmax_u8[e].cpu().numpy(),
            "absmax2": w.absmax2[e].cpu().numpy(), "code2": w.code2.cpu().numpy(), "offset": w.offset[e].item(),
            "shape": (w.rows, w.cols), "blocksize": 64}


# ------------------------------------------------------------------------------------------------------- expert parallel
def _graph_of(fn, warm=1):
    """Capture fn() (kernels + collectives) into a CUDA graph on a side stream; None if capture is not possible."""
    import torch
    try:
        g, side = torch.cuda.CUDAGraph(), torch.cuda.Stream()
        with torch.cuda.stream(side):
            for _ in range(warm):
                fn()
            torch.cuda.synchronize()
            with torch.cuda.graph(g, stream=side):
                fn()
        torch.cuda.synchronize()
        return g
    except Exception as exc:   # noqa: BLE001
        print(f"note: CUDA-graph capture failed ({type(exc).__name__}: {exc}); timing eager launches", file=sys.stderr, flush=True)
        return None


def _timed_max_over_ranks(step, K, dev, dist):
    """K steps between two CUDA events, barrier + synchronize on both sides, MAX over ranks (ms per step)."""
    import torch
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / K], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def run_ep(args, pkg, dev, rank, world, code2):
    """N > 1: the decode stack with the experts sharded over `world` GPUs (BASELINE.json configs[3] / configs[4]).
    `value`: --tokens tokens per rank (weak scaling in tokens; each rank streams only its E/world experts).  Beside it:
    `strong_256` (256 tokens GLOBAL, the configs[3] shape as written), a skewed-routing run, a numerical check of the
    expert-parallel output against the single-GPU layer on the same weights, and (120b shape) the prefill-4096 leg."""
    import torch
    import torch.distributed as dist
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    K_TOP = TOP_K
    big = args.shape == "120b"
    E = 128 if big else args.experts
    L = 36 if big else args.layers
    T, K, Wm = args.tokens, args.steps, max(3, args.warmup)
    nested = args.absmax == "nested"
    plan = ep.EPPlan(rank, world, E)
    El = plan.experts_per_rank
    Ts = max(1, 256 // world)                                     # tokens per rank of the 256-global strong-scaling shape
    Tp = (4096 // world) if big else 0                            # tokens per rank of the configs[4] prefill leg
    Tmax = max(T, Ts, Tp)
    cap_factor = (args.ep_capacity_factor or None) if args.ep_backend == "nccl" else None
    exchange, xnote = ep.make_exchange(plan, K_TOP, H, Tmax, dev, backend=args.ep_backend, capacity_factor=cap_factor)
    if exchange.kind == "nccl" and args.ep_backend == "auto":
        cap_factor = args.ep_capacity_factor or None
        exchange = ep.NcclExchange(plan, plan.capacity(Tmax, K_TOP, cap_factor), H, dev)
    g = torch.Generator(device=dev).manual_seed(SEED + 17 * rank)
    gc = torch.Generator(device=dev).manual_seed(SEED)            # common seed: layer 0 is the same full expert set on every rank
    s = slice(plan.local_experts.start, plan.local_experts.stop)
    full13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=gc, device=dev) * 0.02).half(), code2)
    full2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=gc, device=dev) * 0.02).half(), code2)
    weights = [(ep.shard_experts(full13, plan), ep.shard_experts(full2, plan))]
    for _ in range(L - 1):
        weights.append((pkg.NF4ExpertWeights.quantize((torch.randn(El, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2),
                        pkg.NF4ExpertWeights.quantize((torch.randn(El, H, I, generator=g, device=dev) * 0.02).half(), code2)))

    def make_layers(tokens):
        return [ep.EPMoELayer(w13, w2, plan, K_TOP, H, tokens, dev, device_ops=ep.CudaEPOps(nested=nested), exchange=exchange)
                for w13, w2 in weights]

    def inputs(tokens, skew=False, seed=0):
        gi = torch.Generator(device=dev).manual_seed(SEED + 1000 * seed + rank)
        x = torch.randn(tokens, H, generator=gi, device=dev).half()
        lgs = []
        for _ in range(L):
            lg = torch.randn(tokens, E, generator=gi, device=dev).float()
            if skew:   # logits += 2 * onehot(t mod 4): the first experts (rank 0's) get most of the rows
                lg[torch.arange(tokens, device=dev), torch.arange(tokens, device=dev) % 4] += 2.0
            lgs.append(lg.half())
        return x, lgs

    def timed_stack(tokens, skew=False, seed=0, steps=K):
        layers = make_layers(tokens)
        x, lgs = inputs(tokens, skew, seed)
        outs = [torch.empty(tokens, H, dtype=torch.float16, device=dev) for _ in range(L)]

        def fwd():
            for l, layer in enumerate(layers):
                layer.forward(x, router_logits=lgs[l], out=outs[l])
        fwd()
        torch.cuda.synchronize()
        dist.barrier()
        graph = _graph_of(fwd)
        flag = torch.tensor([1 if graph is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        step = graph.replay if int(flag.item()) else fwd
        for _ in range(Wm):
            step()
        ms = _timed_max_over_ranks(step, steps, dev, dist)
        for layer in layers:
            layer.check_overflow()
        return ms, layers, x, lgs, outs, bool(int(flag.item()))

    # ---- numerical check first: EP output of layer 0 == the single-GPU layer on the same full weights, this rank's tokens ----
    chk_layers = make_layers(T)
    cx, clg = inputs(T, False, 7)
    ep_out = chk_layers[0].forward(cx, router_logits=clg[0])
    single = pkg.ops.moe_forward(cx, full13, full2, K_TOP, router_logits=clg[0], nested=nested,
                                 regime=chk_layers[0].regime_resolved)
    torch.cuda.synchronize()
    same = torch.tensor([1 if torch.equal(ep_out, single) else 0], device=dev)
    maxdiff = (ep_out.float() - single.float()).abs().max().reshape(1)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    dist.all_reduce(maxdiff, op=dist.ReduceOp.MAX)
    ep_matches, ep_maxdiff = bool(int(same.item())), float(maxdiff.item())
    del chk_layers
    if world > 1:
        del full13, full2   # layer 0's shards are views of them; the views keep the storage alive

    # ---- value: T tokens per rank ----
    sampler = ClockSampler(dev.index or 0)
    sampler.start()
    time.sleep(0.25)
    t0 = time.time()
    ms_per_step, layers, x, lgs, outs, graphed_ok = timed_stack(T)
    clocks = sampler.stop(t0, time.time())

    # ---- e2e: host buffers, H2D of x + logits and D2H of the output inside the timed region ----
    hx = x.cpu().pin_memory()
    hlg = torch.stack([l.cpu() for l in lgs]).pin_memory()
    hout = torch.empty(T, H, dtype=torch.float16).pin_memory()

    def stack_apply(dx, dlg):
        o = None
        for l, layer in enumerate(layers):
            o = layer.forward(dx, router_logits=dlg[l])
        return o

    e2e_api = "EPMoELayer.forward(x, router_logits) x layers launched eagerly, pinned host buffers, per rank"
    if graphed_ok:
        graphed = pkg.GraphedForward(stack_apply, [x, torch.stack(lgs)])
        step_e2e = lambda: graphed(hx, hlg, out=hout)
        e2e_api = ("GraphedForward(EPMoELayer.forward(x, router_logits) x layers)(host x, host logits) -> host out per rank: pinned H2D, "
                   "one CUDA-graph replay (exchange inside), D2H, stream sync; host wall clock, max over ranks")
    else:
        def step_e2e():
            o = stack_apply(hx.to(dev, non_blocking=True), hlg.to(dev, non_blocking=True))
            hout.copy_(o, non_blocking=True)
            torch.cuda.current_stream().synchronize()
    for _ in range(Wm):
        step_e2e()
    torch.cuda.synchronize()
    dist.barrier()
    tw = time.perf_counter()
    for _ in range(K):
        step_e2e()
    ms3 = torch.tensor([(time.perf_counter() - tw) * 1e3 / K], device=dev, dtype=torch.float64)
    dist.all_reduce(ms3, op=dist.ReduceOp.MAX)
    e2e_ms = float(ms3.item())

    # ---- per-rank compute-only time of the expert kernels (roofline of this rank's weight stream) ----
    def compute_all():
        for layer in layers:
            layer.compute()
    cg = _graph_of(compute_all)
    tk = []
    for rep in range(4):
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        (cg.replay if cg is not None else compute_all)()
        a1.record()
        torch.cuda.synchronize()
        if rep:
            tk.append(a0.elapsed_time(a1) / L)
    peak, peak_src = measured_peaks()
    wbytes = El * (weights[1][0].nbytes(nested) + weights[1][1].nbytes(nested))
    d = statistics.mean(tk) * 1e-3

    # ---- beside it: the 256-global strong-scaling shape, skewed routing, (120b) the prefill leg ----
    extras = {}
    if not args.no_sweep:
        ms_s, *_ = timed_stack(Ts, False, 11, steps=max(5, K // 2))
        extras["strong_256"] = {"global_tokens": Ts * world, "tokens_per_rank": Ts, "ms_per_step": ms_s, "per_layer_us": ms_s * 1e3 / L,
                                "tokens_per_s": Ts * world / (ms_s * 1e-3),
                                "note": "BASELINE.json configs[3] as written: 256 tokens GLOBAL over N GPUs; the N=1 anchor is `strong_256` of "
                                        "the N=1 line (efficiency = this / (N x anchor) is the driver's to compute)"}
        ms_k, *_ = timed_stack(T, True, 13, steps=max(5, K // 2))
        extras["skewed_routing"] = {"tokens_per_rank": T, "ms_per_step": ms_k, "per_layer_us": ms_k * 1e3 / L, "tokens_per_s": T * world / (ms_k * 1e-3),
                                    "routing": "logits += 2 * onehot(t mod 4) (SURVEY 8(d)): rank 0's experts take most rows",
                                    "dropped_pairs": 0 if exchange.kind == "p2p" else "checked after the run (overflow would have raised)"}
        if big:
            ms_p, *_ = timed_stack(Tp, False, 17, steps=3)
            flops = 2.0 * Tp * world * K_TOP * (2 * I * H + H * I) * L
            extras["prefill_4096"] = {"global_tokens": Tp * world, "tokens_per_rank": Tp, "ms_per_step": ms_p, "per_layer_us": ms_p * 1e3 / L,
                                      "tokens_per_s": Tp * world / (ms_p * 1e-3), "tflops_all_ranks": flops / (ms_p * 1e-3) / 1e12,
                                      "tensor_frac_of_nominal_per_gpu": flops / (ms_p * 1e-3) / 1e12 / world / 2250.0}
    wire = layers[0].wire_bytes()
    # this library's kernels per layer on this rank, by phase (EPMoELayer.forward): counted, not estimated -- the owner chain is
    # what nf4moe_expert_rows reported for the last layer it ran; NCCL's own kernels are not ours and are not counted
    p2p = exchange.kind == "p2p"
    ep_launches = {"route": 1, "bucket": 1, "dispatch_p2p" if p2p else "pack": 1,
                   "wait": 2 if (p2p and not getattr(exchange, "fused_wait", False)) else 0,
                   "expert_rows": pkg.ops.last_launch_count(), "return_p2p": 1 if p2p else 0, "combine_indexed": 1}
    if rank == 0:
        T_all = T * world
        shape_name = "gpt-oss-120b-shape" if big else "gpt-oss-20b"
        line = {"metric": METRIC, "value": T_all / (ms_per_step * 1e-3), "unit": "tokens/s", "n_gpus": world, "steps": K,
                "warmup": Wm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "nf4 weights -> f16 mma operands, f32 accumulate, f16 activations",
                "data": "synthetic (seeded N(0,0.02^2) weights quantised on device, N(0,1) tokens and logits)",
                "config": {"workload": f"{shape_name} {L}-layer NF4 MoE stack decode, {T} tokens per rank ({T_all} global), "
                                       f"{E} experts expert-parallel over {world} GPUs ({El} per GPU), dispatch/combine over {xnote} "
                                       f"(BASELINE.json configs[{4 if big else 3}] shape)", "hidden": H, "intermediate": I, "experts": E,
                           "top_k": K_TOP, "layers": L, "tokens_per_rank": T, "absmax": args.absmax, "parallelism": f"ep{world}",
                           "exchange": exchange.kind, "exchange_note": xnote,
                           "regime": {pkg.REGIME_GEMV: "gemv", pkg.REGIME_GEMM: "gemm_exact",
                                      pkg.REGIME_GEMM_FAST: "gemm_fast (decode tcgen05 kernel)"}.get(layers[0].regime_resolved, "auto"),
                           "capacity_rows_per_peer": layers[0].cap,
                           "capacity": ("worst case (T*k rows per peer): nothing can be dropped" if cap_factor is None else
                                        f"{cap_factor} x mean load per peer, overflow checked after the run (none)"),
                           "cuda_graph": graphed_ok,
                           "l2": f"{wbytes * L / 1e9:.2f} GB of weights streamed per rank per step >> 126 MB L2",
                           "per_layer_us": ms_per_step * 1e3 / L},
                "clocks": clocks,
                "ep_matches_single_gpu": ep_matches, "ep_vs_single_gpu_max_abs_diff": ep_maxdiff,
                "e2e": {"value": T_all / (e2e_ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": hx.numel() * 2 + hlg.numel() * 2,
                        "d2h_bytes_per_step": hout.numel() * 2, "ms_per_step": e2e_ms, "api": e2e_api},
                "gpu_launches": sum(ep_launches.values()) * L * K, "launches_per_layer": ep_launches,
                "collectives_per_layer": exchange.collectives_per_layer, "wire_bytes_per_rank_per_layer": wire,
                "roofline": {"bound": "hbm", "kernel": "expert_rows chain (bucket + permute + w13 + w2) on this rank's experts",
                             "achieved": wbytes / d / 1e9, "peak": peak, "unit": "GB/s", "frac": wbytes / d / 1e9 / peak,
                             "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": wbytes,
                             "avg_launch_us": d * 1e6}}
        line.update(extras)
        print(json.dumps(line), flush=True)
    # Leave without NCCL teardown: destroying a communicator that a live CUDA graph captured can block forever, and a
    # benchmark process has nothing left to release.  Every rank has finished its collectives at this point.
    torch.cuda.synchronize()
    dist.barrier()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


def stack_and_kernel_times(pkg, dev, stack, T, E, regime, nested, seed, reps=6):
    """One configuration beside the headline: the whole 24-layer stack (CUDA-graph replay, CUDA events) and the w13 / w2 kernels
    alone (one graph each, all layers back to back), for T tokens in `regime` with the given absmax form."""
    import torch
    L = len(stack)
    g = torch.Generator(device=dev).manual_seed(SEED + seed)
    x = torch.randn(T, H, generator=g, device=dev).half()
    logits = [torch.randn(T, E, generator=g, device=dev).half() for _ in range(L)]
    outs = [torch.empty(T, H, dtype=torch.float16, device=dev) for _ in range(L)]
    ws = pkg.ops.new_workspace(dev, T, E, H, I, TOP_K)
    eff = pkg.ops.auto_regime(T, TOP_K, E, nested) if regime == pkg.REGIME_AUTO else regime
    use_tc = eff != pkg.REGIME_GEMV

    def fwd():
        for l, (w13, w2) in enumerate(stack):
            pkg.ops.moe_forward(x, w13, w2, TOP_K, router_logits=logits[l], out=outs[l], workspace=ws, nested=nested, regime=regime)
    graph = _graph_of(fwd)
    prep = []
    for l in range(L):
        ids, _ = pkg.ops.route_topk(logits[l], TOP_K)
        b = pkg.ops.bucket(ids, E, want_tiles=use_tc)
        prep.append((b, pkg.ops.permute_rows(x, b, TOP_K) if use_tc else x))
    eact = sum(int(b["meta"][0].item()) for b, _ in prep) / L
    hs = []

    def run13():
        hs.clear()
        hs.extend(pkg.ops.expert_w13(xin, w13, b, T, TOP_K, nested=nested, regime=eff) for (w13, _), (b, xin) in zip(stack, prep))

    def run2():
        for hbuf, (_, w2), (b, _) in zip(hs, stack, prep):
            pkg.ops.expert_w2(hbuf, w2, b, T, TOP_K, nested=nested, regime=eff)
    side = torch.cuda.Stream()
    run13(); run2(); torch.cuda.synchronize()
    g13, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g13, stream=side):
            run13()
        with torch.cuda.graph(g2, stream=side):
            run2()
    torch.cuda.synchronize()
    ts, t13, t2 = [], [], []
    for rep in range(reps + 2):
        a = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        a[0].record(); graph.replay(); a[1].record(); g13.replay(); a[2].record(); g2.replay(); a[3].record()
        torch.cuda.synchronize()
        if rep >= 2:
            ts.append(a[0].elapsed_time(a[1])); t13.append(a[1].elapsed_time(a[2]) / L); t2.append(a[2].elapsed_time(a[3]) / L)
    pkg.ops.check_status(ws)
    peak, _ = measured_peaks()
    ms = statistics.mean(ts)
    # algorithmic bytes: weights + absmax of the active experts, x in, out out; the h / y intermediates do not count
    b13 = eact * stack[0][0].nbytes(nested) + T * H * 2
    b2 = eact * stack[0][1].nbytes(nested) + T * H * 2
    d13, d2 = statistics.mean(t13) * 1e-3, statistics.mean(t2) * 1e-3
    name = {pkg.REGIME_GEMV: "gemv", pkg.REGIME_GEMM: "tcgen05 exact operand", pkg.REGIME_GEMM_FAST: "tcgen05 fast operand"}[eff]
    return {"tokens": T, "regime": name, "absmax": "nested" if nested else "flat", "experts_active_mean": eact,
            "ms_per_step": ms, "per_layer_us": ms * 1e3 / L, "tokens_per_s": T / (ms * 1e-3),
            "w13_us": d13 * 1e6, "w13_gbs": b13 / d13 / 1e9, "w13_frac": b13 / d13 / 1e9 / peak,
            "w2_us": d2 * 1e6, "w2_gbs": b2 / d2 / 1e9, "w2_frac": b2 / d2 / 1e9 / peak,
            "whole_step_frac": (b13 + b2) * L / (ms * 1e-3) / 1e9 / peak,
            "token_tiles": ("whole stack: 64-row token tiles (nested absmax, FAST operand, > 20 rows per expert on average); the w13 / w2 "
                            "columns time the stand-alone entry points, which keep 32-row tiles"
                            if nested and eff == pkg.REGIME_GEMM_FAST and T * TOP_K > 20 * E else "32-row token tiles" if use_tc else "gemv")}


def oracle_check_of_benched_output(pkg, stack, x, logits0, out0, nested, sample=8):
    """The output the timed step produced for layer 0, compared with the CPU oracle on `sample` tokens (checker only; outside
    every timed region).  Returns (max_abs, rel_l2, ok)."""
    import numpy as np
    import torch
    from oracle import nf4_oracle as orc
    w13, w2 = stack[0]
    T = x.shape[0]
    pick = np.unique(np.linspace(0, T - 1, sample).astype(np.int64))
    lg = logits0[pick].cpu().numpy()
    ids, _ = orc.route_topk(lg, TOP_K)
    E = w13.num_experts
    d13 = np.zeros((E, w13.rows, w13.cols), np.float16)
    d2 = np.zeros((E, w2.rows, w2.cols), np.float16)
    for e in np.unique(ids):
        d13[e], d2[e] = orc.dequant_state(state_to_cpu(w13, int(e))), orc.dequant_state(state_to_cpu(w2, int(e)))
    ref = orc.moe_forward_ref(x[pick].cpu().numpy(), lg, d13, d2, k=TOP_K).astype(np.float32)
    got = out0[pick].float().cpu().numpy()
    max_abs = float(np.abs(got - ref).max())
    rel = float(np.linalg.norm(got - ref) / max(np.linalg.norm(ref), 1e-12))
    return max_abs, rel, bool(max_abs <= 1e-2 and rel <= 2e-3)


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module("gpt-oss-vllm_b200")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: there is no CPU fallback on the product path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # rank 0's stdout carries ONE JSON line: NCCL's own banner / debug lines (NCCL_DEBUG set on the box) go to a file instead
        os.environ.setdefault("NCCL_DEBUG_FILE", "/tmp/nf4moe_nccl_%h_%p.log")
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE {world}", file=sys.stderr)

    E, T, L, K, W = args.experts, args.tokens, args.layers, args.steps, max(3, args.warmup)
    nested = args.absmax == "nested"
    code2 = pkg.dynamic_code_map()          # the product's own table generator: the GPU arm needs nothing from oracle/
    if world > 1:
        return run_ep(args, pkg, dev, rank, world, code2)

    regime = {"auto": pkg.REGIME_AUTO, "gemv": pkg.REGIME_GEMV, "gemm": pkg.REGIME_GEMM, "gemm_fast": pkg.REGIME_GEMM_FAST}[args.regime]
    eff = pkg.ops.auto_regime(T, TOP_K, E, nested) if regime == pkg.REGIME_AUTO else regime     # what AUTO resolves to
    eff_name = {pkg.REGIME_GEMV: "gemv", pkg.REGIME_GEMM: "gemm_exact", pkg.REGIME_GEMM_FAST: "gemm_fast"}[eff]
    use_tc = eff != pkg.REGIME_GEMV
    dec_kernel = use_tc and (T * TOP_K <= 40 * E or (E <= 8 and T * TOP_K <= 128 * E))   # gemm_tile_tokens == 32
    act_fused = True   # the bench runs the gpt-oss clamped SwiGLU, which the decode kernel fuses (so w13 and w2 share one launch)
    stack = make_stack(pkg, dev, L, E, code2)
    g = torch.Generator(device=dev).manual_seed(SEED + 1)
    x = torch.randn(T, H, generator=g, device=dev).half()
    logits = [torch.randn(T, E, generator=g, device=dev).half() for _ in range(L)]
    outs = [torch.empty(T, H, dtype=torch.float16, device=dev) for _ in range(L)]
    ws = pkg.ops.get_workspace(dev, T, E, H, I, TOP_K)

    def stack_forward():
        for l, (w13, w2) in enumerate(stack):
            pkg.ops.moe_forward(x, w13, w2, TOP_K, router_logits=logits[l], out=outs[l], workspace=ws, nested=nested,
                                regime=regime)

    stack_forward()
    launches_per_layer = pkg.ops.last_launch_count()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        stack_forward()
        torch.cuda.synchronize()
        with torch.cuda.graph(graph, stream=side):
            stack_forward()
    torch.cuda.synchronize()

    # ---- value: device-resident, CUDA-graph replay, CUDA events ----
    for _ in range(W):
        graph.replay()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    t_wall0 = time.time()
    torch.cuda.synchronize()
    ev[0].record()
    for i in range(K):
        graph.replay()
        ev[i + 1].record()
    torch.cuda.synchronize()
    t_wall1 = time.time()
    step_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(K)]
    total_ms = ev[0].elapsed_time(ev[K])
    ms_per_step = total_ms / K
    value = T / (ms_per_step * 1e-3)

    # ---- e2e: public API, host buffers, H2D + D2H inside the timed region ----
    method = pkg.B200NF4MoEMethod(top_k=TOP_K, nested=nested, regime=regime)
    layers_mod = []
    for w13, w2 in stack:
        m = torch.nn.Module()
        method.bind(m, w13, w2)
        layers_mod.append(m)
    hx = x.cpu().pin_memory()
    hlg = torch.stack([l.cpu() for l in logits]).pin_memory()
    hout = torch.empty(T, H, dtype=torch.float16).pin_memory()

    def e2e_step():
        dx = hx.to(dev, non_blocking=True)
        dlg = hlg.to(dev, non_blocking=True)
        o = None
        for l, m in enumerate(layers_mod):
            o = method.apply_legacy(m, dx, dlg[l], TOP_K)
        hout.copy_(o, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(W):
        e2e_step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        e2e_step()
    e1.record()
    torch.cuda.synchronize()
    e2e_eager_ms = e0.elapsed_time(e1) / K

    # the same call the way a serving engine issues it (vLLM replays decode steps from CUDA graphs): pkg.GraphedForward captures
    # the 24 apply_legacy calls once; each step = H2D of x + logits from pinned memory, one replay, D2H of the output
    def stack_apply(dx, dlg):
        o = None
        for l, m in enumerate(layers_mod):
            o = method.apply_legacy(m, dx, dlg[l], TOP_K)
        return o

    graphed = pkg.GraphedForward(stack_apply, [x, torch.stack(logits)])
    for _ in range(W):
        graphed(hx, hlg, out=hout)
    torch.cuda.synchronize()
    t_e0 = time.perf_counter()
    for _ in range(K):
        graphed(hx, hlg, out=hout)          # returns after its stream has synchronised: host wall clock is the honest timer
    e2e_ms = (time.perf_counter() - t_e0) * 1e3 / K
    e2e_check = torch.equal(hout.to(dev), outs[L - 1])
    clocks = sampler.stop(t_wall0, time.time())

    # ---- roofline of the dominant kernel (w13 GEMV), timed live with CUDA events on the launching stream ----
    peak, peak_src = measured_peaks()
    # The kernels are launched back to back from CUDA graphs (one graph = the w13 kernel of all L layers, one = the w2 kernels,
    # each layer with its own routing), so the event interval holds kernel time only, not ctypes/launch gaps; weights rotate
    # over L layers (9.6 GB), so no launch finds its weights in L2.
    prep = []
    for l, (w13, w2) in enumerate(stack):
        ids, wts = pkg.ops.route_topk(logits[l], TOP_K)
        b = pkg.ops.bucket(ids, E, want_tiles=use_tc)
        prep.append((b, pkg.ops.permute_rows(x, b, TOP_K) if use_tc else x))
    eact = [int(b["meta"][0].item()) for b, _ in prep]

    def run13():
        return [pkg.ops.expert_w13(xin, w13, b, T, TOP_K, nested=nested, regime=eff) for (w13, _), (b, xin) in zip(stack, prep)]

    def run2(hs):
        for hbuf, (_, w2), (b, _) in zip(hs, stack, prep):
            pkg.ops.expert_w2(hbuf, w2, b, T, TOP_K, nested=nested, regime=eff)

    hs = run13(); run2(hs); torch.cuda.synchronize()
    g13, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g13, stream=side):
            hs = run13()
        with torch.cuda.graph(g2, stream=side):
            run2(hs)
    torch.cuda.synchronize()
    t13, t2 = [], []
    for rep in range(4):
        a0, a1, a2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        a0.record(); g13.replay(); a1.record(); g2.replay(); a2.record()
        torch.cuda.synchronize()
        if rep > 0:
            t13.append(a0.elapsed_time(a1) / L)
            t2.append(a1.elapsed_time(a2) / L)
    # The decode tcgen05 regime serves w13 + SwiGLU and w2 in ONE launch (dec_kernel<2>): that launch is the step's dominant
    # kernel.  Timed alone the same way; its tile lists (which carry the w13-done counters) are rewritten by an untimed replay of
    # the bucket kernels before every timed replay.
    t_both = []
    if dec_kernel and act_fused:
        hy = [pkg.ops.expert_w13_w2(xin, w13, w2, b, T, TOP_K, nested=nested, regime=eff) for (w13, w2), (b, xin) in zip(stack, prep)]
        idss = [pkg.ops.route_topk(logits[l], TOP_K)[0] for l in range(L)]
        torch.cuda.synchronize()
        g_reset, g_both = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(g_reset, stream=side):
                for l, (b, _) in enumerate(prep):
                    pkg.ops.bucket(idss[l], E, want_tiles=True, into=b)
            with torch.cuda.graph(g_both, stream=side):
                for (w13, w2), (b, xin), (hb, yb) in zip(stack, prep, hy):
                    pkg.ops.expert_w13_w2(xin, w13, w2, b, T, TOP_K, nested=nested, regime=eff, h=hb, y=yb)
        torch.cuda.synchronize()
        for rep in range(4):
            g_reset.replay()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(); g_both.replay(); a1.record()
            torch.cuda.synchronize()
            if rep > 0:
                t_both.append(a0.elapsed_time(a1) / L)
    ea = sum(eact) / len(eact)
    # algorithmic bytes: E_act * (packed + absmax) + T * (2 * H * 2 + E * 2); the h [T k, I] and y [T k, H] intermediates
    # are NOT algorithmic bytes -- they count against the kernel, not for it
    b13 = ea * stack[0][0].nbytes(nested) + T * H * 2
    b2 = ea * stack[0][1].nbytes(nested) + T * H * 2
    d13, d2 = statistics.mean(t13) * 1e-3, statistics.mean(t2) * 1e-3
    layer_bytes = b13 + b2 + T * E * 2
    traffic, traffic_src = None, None
    try:   # dram bytes of this kernel at this configuration from the committed `ncu --set full` capture (never measured under ncu here)
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            tj = json.load(f)
        key = "dec" if dec_kernel else eff_name
        if T == 64 and E == 32 and nested and key in tj:
            traffic, traffic_src = tj[key]["traffic_bytes_per_launch"], tj[key]["source"]
    except Exception:
        pass
    if t_both:   # the merged launch is what the step runs: report it as the dominant kernel, the split kernels beside it
        d_both = statistics.mean(t_both) * 1e-3
        key = "dec2"
        traffic, traffic_src = None, None
        try:
            if T == 64 and E == 32 and nested and key in tj:
                traffic, traffic_src = tj[key]["traffic_bytes_per_launch"], tj[key]["source"]
        except Exception:
            pass
        roof = {"bound": "hbm", "kernel": "dec_kernel<SIDES=2> (decode-regime tcgen05 kernel: NF4 -> TMEM A operand; w13 + SwiGLU and w2 in one persistent launch)",
                "achieved": layer_bytes / d_both / 1e9, "peak": peak, "unit": "GB/s", "frac": layer_bytes / d_both / 1e9 / peak,
                "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src, "algorithmic_bytes_per_launch": layer_bytes,
                "avg_launch_us": d_both * 1e6, "frac_of_nominal_8TBs": layer_bytes / d_both / 8e12,
                "split_launches": {"w13": {"achieved": b13 / d13 / 1e9, "frac": b13 / d13 / 1e9 / peak, "avg_launch_us": d13 * 1e6,
                                           "algorithmic_bytes_per_launch": b13},
                                   "w2": {"achieved": b2 / d2 / 1e9, "frac": b2 / d2 / 1e9 / peak, "avg_launch_us": d2 * 1e6,
                                          "algorithmic_bytes_per_launch": b2},
                                   "note": "the same work as two launches of the same kernel template (SIDES = 0 / 1), for comparison"},
                "whole_step": {"achieved": layer_bytes * L / (ms_per_step * 1e-3) / 1e9,
                               "frac": layer_bytes * L / (ms_per_step * 1e-3) / 1e9 / peak, "experts_active_mean": ea}}
    kname = ("dec_kernel<W13> (decode-regime tcgen05 kernel: NF4 -> TMEM A operand, w13 + SwiGLU)" if dec_kernel else
             "gemm_kernel<W13> (tcgen05 grouped GEMM, w13 + SwiGLU)" if use_tc else "gemv_kernel<NT=2,W13> (w13 NF4 GEMV + SwiGLU)")
    roof = roof if t_both else {"bound": "hbm", "kernel": kname, "achieved": b13 / d13 / 1e9, "peak": peak,
            "unit": "GB/s", "frac": b13 / d13 / 1e9 / peak, "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
            "algorithmic_bytes_per_launch": b13, "avg_launch_us": d13 * 1e6, "frac_of_nominal_8TBs": b13 / d13 / 8e12,
            "w2_kernel": {"achieved": b2 / d2 / 1e9, "frac": b2 / d2 / 1e9 / peak, "avg_launch_us": d2 * 1e6,
                          "algorithmic_bytes_per_launch": b2},
            "whole_step": {"achieved": layer_bytes * L / (ms_per_step * 1e-3) / 1e9,
                           "frac": layer_bytes * L / (ms_per_step * 1e-3) / 1e9 / peak, "experts_active_mean": ea}}

    line = {"metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": 1, "steps": K, "warmup": W,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "nf4 weights -> f16 mma operands, f32 accumulate, f16 activations", "data": "synthetic (seeded N(0,0.02^2) weights quantised on device, N(0,1) tokens and logits)",
            "config": {"workload": f"gpt-oss-20b {L}-layer NF4 MoE stack decode, T={T} (BASELINE.json configs[1])", "hidden": H,
                       "intermediate": I, "experts": E, "top_k": TOP_K, "layers": L, "tokens": T, "absmax": args.absmax,
                       "regime": args.regime, "regime_resolved": eff_name, "l2": "9.6 GB of weights streamed per step >> 126 MB L2 (no flush needed)",
                       "cuda_graph": True, "per_layer_us": ms_per_step * 1e3 / L},
            "clocks": clocks,
            "e2e": {"value": T / (e2e_ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": hx.numel() * 2 + hlg.numel() * 2,
                    "d2h_bytes_per_step": hout.numel() * 2, "ms_per_step": e2e_ms,
                    "api": "GraphedForward(B200NF4MoEMethod.apply_legacy(layer, x, router_logits, top_k) x 24)(host x, host logits) -> "
                           "host out: pinned H2D, one CUDA-graph replay, D2H, stream sync; host wall clock",
                    "matches_device_run": bool(e2e_check),
                    "eager": {"value": T / (e2e_eager_ms * 1e-3), "ms_per_step": e2e_eager_ms,
                              "api": "the same 24 apply_legacy calls launched eagerly from Python (ctypes), CUDA events"}},
            "gpu_launches": launches_per_layer * L * K, "launches_per_layer": launches_per_layer,
            "roofline": roof, "step_ms_min_max": [min(step_ms), max(step_ms)]}

    mx, rel, ok = oracle_check_of_benched_output(pkg, stack, x, logits[0], outs[0], nested)
    line["parity_of_benched_output"] = {"layer": 0, "tokens_checked": 8, "max_abs": mx, "rel_l2": rel, "within_1e-2_2e-3": ok,
                                        "checker": "oracle (numpy) on the routed experts, outside the timed region"}
    pkg.ops.check_status()
    if not args.no_sweep:
        # configs[1] says batch 1-64: the whole range in the driver-visible line (regime = what apply() uses at each T)
        line["decode_sweep"] = [stack_and_kernel_times(pkg, dev, stack, t, E, pkg.REGIME_AUTO, nested, 100 + t) for t in (1, 2, 4, 8, 16, 32, 64)]
        # the bit-exact-dequant operand through the same decode-regime kernel (north_star: "a bit-exact-dequant NF4 MoE layer")
        line["exact_operand_decode"] = stack_and_kernel_times(pkg, dev, stack, T, E, pkg.REGIME_GEMM, nested, 201)
        # the vLLM-bound layout (flat fp32 absmax, SURVEY 8(d) "A = 1,555,200"): same kernels, more algorithmic bytes per expert
        line["flat_absmax_decode"] = stack_and_kernel_times(pkg, dev, stack, T, E, regime, False, 202)
        # the 1-GPU anchor of BASELINE.json configs[3] (256 tokens global)
        line["strong_256"] = stack_and_kernel_times(pkg, dev, stack, 256, E, regime, nested, 203)
    if not args.no_prefill:
        line["prefill"] = prefill_section(pkg, dev, stack[0], E, nested)

    if not args.no_cpu_baseline:
        st13 = [state_to_cpu(stack[0][0], e) for e in range(E)]
        st2 = [state_to_cpu(stack[0][1], e) for e in range(E)]
        Tc = args.cpu_tokens or T                # the same batch as the timed GPU step
        times = cpu_layer_seconds(st13, st2, Tc, E, reps=2, warm=0)
        sec = min(times)
        line["cpu_baseline"] = {"value": Tc / (sec * L), "unit": "tokens/s", "cores": os.cpu_count(), "kind": "port",
                                "sample": f"oracle port of the reference path, ONE of the {L} layers, T={Tc} tokens, all {E} experts "
                                          f"dequantised then fp16 matmul per expert; best of 2; scaled to the stack by 1/{L}",
                                "layer_seconds": sec}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    # The driver reads ONE JSON line from stdout.  C libraries write to file descriptor 1 behind Python's back (NCCL prints its
    # version banner there), so fd 1 is pointed at stderr and Python's own sys.stdout keeps the original descriptor.
    sys.stdout.flush()
    keep = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(keep, "w", buffering=1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
