"""Expert parallelism for the NF4 fused-MoE path: one process per GPU, experts sharded contiguously, tokens data-parallel.

The reference is single-GPU (P/gpt_oss_server.py:25 pins CUDA_VISIBLE_DEVICES=0; SURVEY 2.1 "Collectives: none"), so this is
new design, not a re-creation:

    owner(e) = e // (E / world)                      experts [r*E/W, (r+1)*E/W) live on rank r, weights never move
    home rank:   route -> bucket -> ep_pack          rows written straight into the dispatch send buffer, grouped by owner
    exchange:    all-to-all (rows) + all-to-all (local expert ids)       fixed capacity per peer: no host read of counts,
    owner rank:  expert_rows (grouped GEMV / tcgen05 GEMM on the rows it received)        so the layer is graph-capturable
    exchange:    all-to-all (result rows back into the same slots)
    home rank:   combine_indexed                     the same fp32 weighted sum, slot ascending, as on one GPU

Because the home rank sums the very fp16 rows a single GPU would have produced, in the same order, the expert-parallel
output is bit-identical to the single-GPU output.

The collective is NCCL over NVLink/NVSwitch through `torch.distributed` (plumbing); every arithmetic step is a kernel of
libnf4moe.  `device_ops` exists so that the choreography (buffer layout, splits, slot maps) can be exercised on CPU with
`gloo` by the tests, which inject an oracle-backed implementation; the product default is `CudaEPOps` and nothing here
falls back to it silently -- CUDA tensors are required unless a `device_ops` object is passed explicitly.
"""
from __future__ import annotations

import os
import statistics
import time
from dataclasses import dataclass

import torch
import torch.distributed as dist

from . import _lib, ops
from .quant import NF4ExpertWeights


@dataclass(frozen=True)
class EPPlan:
    """Static partition of E experts over `world` ranks."""
    rank: int
    world: int
    num_experts: int

    def __post_init__(self):
        if self.world <= 0 or not (0 <= self.rank < self.world):
            raise ValueError("bad rank/world")
        if self.num_experts % self.world:
            raise ValueError(f"{self.num_experts} experts do not divide over {self.world} ranks")

    @property
    def experts_per_rank(self) -> int:
        return self.num_experts // self.world

    def owner(self, expert: int) -> int:
        return expert // self.experts_per_rank

    @property
    def local_experts(self) -> range:
        return range(self.rank * self.experts_per_rank, (self.rank + 1) * self.experts_per_rank)

    def capacity(self, max_tokens: int, top_k: int, factor: float | None = None) -> int:
        """Rows reserved per peer.  Default: the worst case (every pair of this rank goes to one peer), which can never
        overflow.  `factor` trades bytes on the wire for a (detected, fatal) overflow risk: factor * mean load."""
        worst = max_tokens * top_k
        if factor is None:
            return worst
        mean = -(-worst // self.world)
        return min(worst, max(16, int(factor * mean + 15) // 16 * 16))


_COALESCE_DISPATCH = os.environ.get("NF4MOE_EP_COALESCE", "1") == "1"


class CudaEPOps:
    """The product implementation of the five device steps: kernels of libnf4moe via the C ABI."""

    def __init__(self, nested: bool | None = None, act: int = _lib.ACT_SWIGLU_OAI):
        self.nested, self.act = nested, act
        _lib.load_library()

    def route(self, logits, k):
        return ops.route_topk(logits, k)

    def bucket(self, ids, E):
        return ops.bucket(ids, E, want_tiles=False)

    def pack(self, x, ids, b, E, world, cap, send_rows, send_ids, pair_slot, status):
        ops.ep_pack(x, ids, b, E, world, cap, send_rows, send_ids, pair_slot, status)

    def expert_rows(self, rows, row_expert, w13, w2, y_rows, regime):
        ops.expert_rows(rows, row_expert, w13, w2, y_rows, act=self.act, regime=regime, nested=self.nested)

    def combine(self, back_rows, weights, pair_slot, out):
        return ops.combine_indexed(back_rows, weights, pair_slot, out)


class EPMoELayer:
    """One MoE layer, expert-parallel.  `w13` / `w2` hold only this rank's experts (in global order)."""

    def __init__(self, w13, w2, plan: EPPlan, top_k: int, hidden_size: int, max_tokens: int, device,
                 group=None, device_ops=None, capacity_factor: float | None = None, regime: int | None = None):
        self.w13, self.w2, self.plan, self.k, self.H = w13, w2, plan, top_k, hidden_size
        self.max_tokens, self.group = max_tokens, group
        self.device = torch.device(device)
        if device_ops is None:
            if self.device.type != "cuda":
                raise RuntimeError("EPMoELayer needs a CUDA device: there is no CPU fallback on the product path")
            device_ops = CudaEPOps()
        self.ops = device_ops
        W = plan.world
        self.cap = plan.capacity(max_tokens, top_k, capacity_factor)
        n = W * self.cap
        # default: the library's own choice for (W * cap rows, this rank's experts) -- GEMV for a handful of rows, the decode
        # tcgen05 kernel while token tiles stay <= 32 rows, the prefill grouped GEMM above (nf4moe_auto_regime)
        self.regime = regime if regime is not None else _lib.REGIME_AUTO
        self.regime_resolved = self.regime
        if self.regime == _lib.REGIME_AUTO and self.device.type == "cuda":
            self.regime_resolved = int(_lib.load_library().nf4moe_auto_regime(n, 1, w13.num_experts))
        f16 = dict(dtype=torch.float16, device=self.device)
        i32 = dict(dtype=torch.int32, device=self.device)
        self.send_rows = torch.zeros(n, hidden_size, **f16)
        self.recv_rows = torch.zeros(n, hidden_size, **f16)
        self.y_rows = torch.zeros(n, hidden_size, **f16)
        self.back_rows = torch.zeros(n, hidden_size, **f16)
        self.send_ids = torch.full((n,), -1, **i32)
        self.recv_ids = torch.full((n,), -1, **i32)
        self.pair_slot = torch.full((max_tokens * top_k,), -1, **i32)
        self.status = torch.zeros(4, **i32)

    # The forward is split in phases so that tests can run several virtual ranks in one process and play the exchange.
    def prepare(self, x, router_logits=None, topk_ids=None, topk_weights=None):
        T = x.shape[0]
        if T > self.max_tokens:
            raise ValueError(f"{T} tokens exceed max_tokens={self.max_tokens} this layer's buffers were sized for")
        if router_logits is not None:
            topk_ids, topk_weights = self.ops.route(router_logits, self.k)
        elif topk_ids is None or topk_weights is None:
            raise ValueError("give router_logits or (topk_ids, topk_weights)")
        b = self.ops.bucket(topk_ids, self.plan.num_experts)
        self.ops.pack(x, topk_ids, b, self.plan.num_experts, self.plan.world, self.cap, self.send_rows, self.send_ids,
                      self.pair_slot, self.status)
        self._weights, self._T = topk_weights, T
        return topk_ids, topk_weights

    def dispatch(self):
        if _COALESCE_DISPATCH and self.plan.world > 1 and self.device.type == "cuda" and torch.cuda.is_current_stream_capturing():
            # under CUDA-graph capture: ids and rows to every peer inside ONE NCCL group (one collective kernel instead of two;
            # +4 % at 2 GPUs).  Eagerly the 4W Python P2POps per layer cost more than the second collective, so not there.
            W, cap, ops_ = self.plan.world, self.cap, []
            for p in range(W):
                sl = slice(p * cap, (p + 1) * cap)
                ops_.append(dist.P2POp(dist.isend, self.send_ids[sl], p, self.group))
                ops_.append(dist.P2POp(dist.isend, self.send_rows[sl], p, self.group))
                ops_.append(dist.P2POp(dist.irecv, self.recv_ids[sl], p, self.group))
                ops_.append(dist.P2POp(dist.irecv, self.recv_rows[sl], p, self.group))
            for req in dist.batch_isend_irecv(ops_):
                req.wait()
            return
        dist.all_to_all_single(self.recv_ids, self.send_ids, group=self.group)
        dist.all_to_all_single(self.recv_rows, self.send_rows, group=self.group)

    def compute(self):
        self.ops.expert_rows(self.recv_rows, self.recv_ids, self.w13, self.w2, self.y_rows, self.regime)

    def ret(self):
        dist.all_to_all_single(self.back_rows, self.y_rows, group=self.group)

    def finish(self, out=None):
        return self.ops.combine(self.back_rows, self._weights, self.pair_slot[: self._T * self.k], out)

    def forward(self, x, router_logits=None, topk_ids=None, topk_weights=None, out=None):
        self.prepare(x, router_logits, topk_ids, topk_weights)
        self.dispatch()
        self.compute()
        self.ret()
        return self.finish(out)

    __call__ = forward

    def check_overflow(self) -> None:
        """Host-side check (synchronises): raise if a capacity_factor < worst case dropped pairs."""
        if int(self.status[0].item()) & 1:
            raise RuntimeError("expert-parallel dispatch overflowed its per-peer capacity; raise capacity_factor")

    def wire_bytes(self) -> dict:
        """Bytes this rank puts on the wire per forward (fixed-capacity buffers, (W-1)/W of them leave the GPU)."""
        W, n = self.plan.world, self.plan.world * self.cap
        off = (W - 1) / W
        return {"dispatch": int(n * (self.H * 2 + 4) * off), "combine": int(n * self.H * 2 * off)}


def shard_experts(w: NF4ExpertWeights, plan: EPPlan) -> NF4ExpertWeights:
    """This rank's slice of a full expert set (views, no copy)."""
    r = plan.local_experts
    s = slice(r.start, r.stop)
    pick = lambda t: None if t is None else t[s]
    code2 = w.code2 if (w.code2 is None or w.code2.dim() == 1) else w.code2[s]
    return NF4ExpertWeights(w.packed[s], w.rows, w.cols, pick(w.absmax), pick(w.absmax_u8), pick(w.absmax2), code2,
                            pick(w.offset))


# ------------------------------------------------------------------------------------------------------------- bench
def bench_ep(args, pkg, dev, rank, world, code2, ClockSampler, measured_peaks, metric):
    """bench.py's N>1 arm: the 24-layer decode stack with experts sharded over `world` GPUs, `--tokens` tokens per rank
    (weak scaling in tokens; each rank streams only its E/world experts).  Max-over-ranks CUDA-event timing."""
    import json
    H = I = 2880
    K_TOP = 4
    E, T, L, K, Wm = args.experts, args.tokens, args.layers, args.steps, max(3, args.warmup)
    nested = args.absmax == "nested"
    plan = EPPlan(rank, world, E)
    El = plan.experts_per_rank
    # fixed per-peer capacity keeps the exchange graph-capturable (no host read of counts); 2 x the mean load is > 6 sigma of the
    # per-peer count for near-uniform routing, and check_overflow() below makes a dropped pair fatal rather than silent
    cap_factor = getattr(args, "ep_capacity_factor", 0.0) or None
    g = torch.Generator(device=dev).manual_seed(20250805 + 17 * rank)
    layers = []
    for _ in range(L):
        w13 = NF4ExpertWeights.quantize((torch.randn(El, 2 * I, H, generator=g, device=dev) * 0.02).half(), code2)
        w2 = NF4ExpertWeights.quantize((torch.randn(El, H, I, generator=g, device=dev) * 0.02).half(), code2)
        layers.append(EPMoELayer(w13, w2, plan, K_TOP, H, T, dev, device_ops=CudaEPOps(nested=nested), capacity_factor=cap_factor))
    x = torch.randn(T, H, generator=g, device=dev).half()
    logits = [torch.randn(T, E, generator=g, device=dev).half() for _ in range(L)]
    outs = [torch.empty(T, H, dtype=torch.float16, device=dev) for _ in range(L)]

    def stack_forward():
        for l, layer in enumerate(layers):
            layer.forward(x, router_logits=logits[l], out=outs[l])

    stack_forward()
    torch.cuda.synchronize()
    dist.barrier()
    graph, use_graph = None, os.environ.get("NF4MOE_EP_GRAPH", "1") == "1"
    if use_graph:
        try:
            graph = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            with torch.cuda.stream(side):
                stack_forward()
                torch.cuda.synchronize()
                with torch.cuda.graph(graph, stream=side):
                    stack_forward()
            torch.cuda.synchronize()
        except Exception as exc:  # NCCL capture unsupported in this build: time eager launches instead
            graph = None
            if rank == 0:
                print(f"note: CUDA-graph capture of the EP step failed ({type(exc).__name__}); timing eager", flush=True)
    flags = torch.tensor([1 if graph is not None else 0], device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if int(flags.item()) == 0:
        graph = None
    step = graph.replay if graph is not None else stack_forward
    for _ in range(Wm):
        step()
    torch.cuda.synchronize()
    dist.barrier()
    sampler = ClockSampler(dev.index or 0)
    sampler.start()
    time.sleep(0.25)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.time()
    e0.record()
    for _ in range(K):
        step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1) / K], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms.item())

    # e2e: host buffers, H2D of x + logits and D2H of the output inside the timed region, public layer call
    hx = x.cpu().pin_memory()
    hlg = torch.stack([l.cpu() for l in logits]).pin_memory()
    hout = torch.empty(T, H, dtype=torch.float16).pin_memory()

    def e2e_step():
        dx = hx.to(dev, non_blocking=True)
        dlg = hlg.to(dev, non_blocking=True)
        o = None
        for l, layer in enumerate(layers):
            o = layer.forward(dx, router_logits=dlg[l])
        hout.copy_(o, non_blocking=True)
        torch.cuda.current_stream().synchronize()

    for _ in range(Wm):
        e2e_step()
    dist.barrier()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(K):
        e2e_step()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    ms2 = torch.tensor([e0.elapsed_time(e1) / K], device=dev, dtype=torch.float64)
    dist.all_reduce(ms2, op=dist.ReduceOp.MAX)
    e2e_eager_ms = float(ms2.item())
    e2e_ms, e2e_api = e2e_eager_ms, "EPMoELayer.forward(x, router_logits) x layers launched eagerly, pinned host buffers, per rank"
    if graph is not None:
        # the serving-engine form of the same call: GraphedForward captures the L EPMoELayer.forward calls (collectives included);
        # a step = H2D of x + logits from pinned memory, one replay, D2H of the output, stream sync; host wall clock, max over ranks
        from .graphed import GraphedForward

        def stack_apply(dx, dlg):
            o = None
            for l, layer in enumerate(layers):
                o = layer.forward(dx, router_logits=dlg[l])
            return o

        graphed = GraphedForward(stack_apply, [x, torch.stack(logits)])
        for _ in range(Wm):
            graphed(hx, hlg, out=hout)
        torch.cuda.synchronize()
        dist.barrier()
        t_e0 = time.perf_counter()
        for _ in range(K):
            graphed(hx, hlg, out=hout)
        ms3 = torch.tensor([(time.perf_counter() - t_e0) * 1e3 / K], device=dev, dtype=torch.float64)
        dist.all_reduce(ms3, op=dist.ReduceOp.MAX)
        e2e_ms = float(ms3.item())
        e2e_api = ("GraphedForward(EPMoELayer.forward(x, router_logits) x layers)(host x, host logits) -> host out per rank: "
                   "pinned H2D, one CUDA-graph replay (NCCL all-to-alls inside), D2H, stream sync; host wall clock, max over ranks")
    clocks = sampler.stop(t0, time.time())
    for layer in layers:
        layer.check_overflow()

    # per-rank compute-only time of the expert kernels (roofline of this rank's weight stream)
    # (the L layers' expert chains replayed from one CUDA graph between two events: kernel time, no Python launch gaps)
    tk = []
    try:
        cg, cs = torch.cuda.CUDAGraph(), torch.cuda.Stream()
        with torch.cuda.stream(cs):
            for layer in layers:
                layer.compute()
            cs.synchronize()
            with torch.cuda.graph(cg, stream=cs):
                for layer in layers:
                    layer.compute()
        torch.cuda.synchronize()
        for rep in range(4):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            cg.replay()
            a1.record()
            torch.cuda.synchronize()
            if rep:
                tk.append(a0.elapsed_time(a1) / len(layers))
    except Exception:
        tk = []
    if not tk:
        for rep in range(3):
            for layer in layers:
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                layer.compute()
                a1.record()
                torch.cuda.synchronize()
                if rep:
                    tk.append(a0.elapsed_time(a1))
    peak, peak_src = measured_peaks()
    wbytes = El * (layers[0].w13.nbytes(nested) + layers[0].w2.nbytes(nested))
    d = statistics.mean(tk) * 1e-3
    wire = layers[0].wire_bytes()
    if rank == 0:
        T_all = T * world
        line = {"metric": metric, "value": T_all / (ms_per_step * 1e-3), "unit": "tokens/s", "n_gpus": world, "steps": K,
                "warmup": Wm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "nf4 weights -> f16 mma operands, f32 accumulate, f16 activations",
                "data": "synthetic (seeded N(0,0.02^2) weights quantised on device, N(0,1) tokens and logits)",
                "config": {"workload": f"gpt-oss-20b {L}-layer NF4 MoE stack decode, {T} tokens per rank ({T_all} global), "
                                       f"{E} experts expert-parallel over {world} GPUs ({El} per GPU), NCCL all-to-all dispatch/"
                                       f"combine (BASELINE.json configs[3] shape)", "hidden": H, "intermediate": I, "experts": E,
                           "top_k": K_TOP, "layers": L, "tokens_per_rank": T, "absmax": args.absmax, "parallelism": f"ep{world}",
                           "regime": {_lib.REGIME_GEMV: "gemv", _lib.REGIME_GEMM: "gemm_exact",
                                      _lib.REGIME_GEMM_FAST: "gemm_fast (decode tcgen05 kernel)"}.get(layers[0].regime_resolved, "auto"),
                           "capacity_rows_per_peer": layers[0].cap,
                           "capacity": ("worst case (T*k rows per peer)" if cap_factor is None else
                                        f"{cap_factor} x mean load per peer, overflow checked after the run (none)"),
                           "cuda_graph": graph is not None,
                           "l2": f"{wbytes * L / 1e9:.2f} GB of weights streamed per rank per step >> 126 MB L2",
                           "per_layer_us": ms_per_step * 1e3 / L},
                "clocks": clocks,
                "e2e": {"value": T_all / (e2e_ms * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": hx.numel() * 2 + hlg.numel() * 2,
                        "d2h_bytes_per_step": hout.numel() * 2, "ms_per_step": e2e_ms,
                        "api": e2e_api, "eager": {"value": T_all / (e2e_eager_ms * 1e-3), "ms_per_step": e2e_eager_ms}},
                "gpu_launches": (5 + (1 if layers[0].regime_resolved == _lib.REGIME_GEMV else 2)) * L * K,
                "collectives_per_layer": 3, "wire_bytes_per_rank_per_layer": wire,
                "roofline": {"bound": "hbm", "kernel": "expert_rows chain (bucket + w13 + w2) on this rank's experts",
                             "achieved": wbytes / d / 1e9, "peak": peak, "unit": "GB/s", "frac": wbytes / d / 1e9 / peak,
                             "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": wbytes,
                             "avg_launch_us": d * 1e6}}
        print(json.dumps(line), flush=True)
    # Leave without NCCL teardown: destroying a communicator that a live CUDA graph captured can block forever, and a
    # benchmark process has nothing left to release.  Every rank has finished its collectives at this point.
    torch.cuda.synchronize()
    import sys
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)
