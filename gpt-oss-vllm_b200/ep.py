"""Expert parallelism for the NF4 fused-MoE path: one process per GPU, experts sharded contiguously, tokens data-parallel.

The reference is single-GPU (P/gpt_oss_server.py:25 pins CUDA_VISIBLE_DEVICES=0; SURVEY 2.1 "Collectives: none"), so this is
new design, not a re-creation:

    owner(e) = e // (E / world)                      experts [r*E/W, (r+1)*E/W) live on rank r, weights never move
    home rank:   route -> bucket -> ep_pack          rows written straight into the dispatch send buffer, grouped by owner
    exchange:    rows + local expert ids to the owner                    fixed capacity per peer: no host read of counts,
    owner rank:  expert_rows (grouped GEMV / tcgen05 GEMM on the rows it received)        so the layer is graph-capturable
    exchange:    result rows back into the same slots
    home rank:   combine_indexed                     the same fp32 weighted sum, slot ascending, as on one GPU

Because the home rank sums the very fp16 rows a single GPU would have produced, in the same order, the expert-parallel
output is bit-identical to the single-GPU output.

Two exchange back ends:
  * `PeerExchange` (default when torch symmetric memory can be set up): the dispatch kernel stores each row straight into the
    OWNER's receive buffer over NVLink peer memory and publishes an epoch flag; the owner's one-warp wait kernel spins on the
    flags; the return path is the mirror image (csrc/ep.cuh).  No collective library on the data path, only rows that exist
    cross the link, so the capacity is the worst case (T*k rows per peer) at no cost and NO pair is ever dropped.
  * `NcclExchange`: fixed-capacity `all_to_all_single` over NCCL (round 1's path; kept as the fallback and for A/B).
`ReplicatedMoELayer` is the other mode SURVEY 8(e) lists: activations already replicated on every rank, no dispatch, one
all-reduce of the partial outputs.
Every arithmetic step is a kernel of libnf4moe.  `device_ops` exists so that the choreography (buffer layout, splits, slot maps) can be exercised on CPU with
`gloo` by the tests, which inject an oracle-backed implementation; the product default is `CudaEPOps` and nothing here
falls back to it silently -- CUDA tensors are required unless a `device_ops` object is passed explicitly.
"""
from __future__ import annotations

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

    # nf4moe_bucket sorts at most this many (row, expert) pairs per call (csrc/nf4moe.cu: the pair list must fit shared memory)
    MAX_OWNER_SLOTS = 204800

    def check_owner_slots(self, cap: int, hidden_size: int = 0) -> int:
        """An owner runs its experts over ALL world * cap receive slots (most of them empty), so that product -- not the token
        count -- is what the bucket kernel's pair limit and the owner's scratch (recv / back / y rows and the expert workspace,
        O(world * cap * hidden_size)) scale with.  Raises when the slots exceed what nf4moe_bucket takes: 8 ranks at the
        worst-case capacity stop at 6,400 tokens per rank, so prefill-sized chunks need a capacity_factor.  Returns the slots."""
        slots = self.world * cap
        if slots > self.MAX_OWNER_SLOTS:
            gb = f" (and about {slots * hidden_size * 2 * 6 / 2**30:.1f} GiB of owner-side buffers)" if hidden_size else ""
            raise ValueError(
                f"expert parallel: {self.world} ranks x capacity {cap} = {slots} receive slots per owner exceed the "
                f"{self.MAX_OWNER_SLOTS} pairs nf4moe_bucket sorts{gb}; pass capacity_factor (e.g. 2.0 = twice the mean load; an "
                f"overflow is detected, never silent) or feed fewer tokens per call")
        return slots


_COALESCE_DISPATCH = True


class CudaEPOps:
    """The product implementation of the device steps: kernels of libnf4moe via the C ABI."""

    def __init__(self, nested: bool | None = None, act: int = _lib.ACT_SWIGLU_OAI):
        self.nested, self.act = nested, act
        _lib.load_library()

    def route(self, logits, k):
        return ops.route_topk(logits, k)

    def bucket(self, ids, E):
        return ops.bucket(ids, E, want_tiles=False)

    def pack(self, x, ids, b, E, world, cap, send_rows, send_ids, pair_slot, status):
        ops.ep_pack(x, ids, b, E, world, cap, send_rows, send_ids, pair_slot, status)

    def expert_rows(self, rows, row_expert, w13, w2, y_rows, regime, w13_bias=None, rows_hint=0):
        ops.expert_rows(rows, row_expert, w13, w2, y_rows, act=self.act, regime=regime, nested=self.nested, w13_bias=w13_bias,
                        rows_hint=rows_hint)

    def combine(self, back_rows, weights, pair_slot, out, ids=None, w2_bias=None):
        return ops.combine_indexed(back_rows, weights, pair_slot, out, ids=ids, w2_bias=w2_bias)


# --------------------------------------------------------------------------------------------------------- exchanges
class NcclExchange:
    """Fixed-capacity all-to-all through torch.distributed (NCCL on GPUs, gloo in the CPU tests).  Buffers: send/recv rows and
    ids [W*cap], y_rows / back_rows [W*cap, H]."""
    kind = "nccl"

    def __init__(self, plan: EPPlan, cap: int, hidden_size: int, device, group=None):
        self.plan, self.cap, self.group, self.device = plan, cap, group, torch.device(device)
        n = plan.world * cap
        f16 = dict(dtype=torch.float16, device=self.device)
        i32 = dict(dtype=torch.int32, device=self.device)
        self.send_rows, self.recv_rows = torch.zeros(n, hidden_size, **f16), torch.zeros(n, hidden_size, **f16)
        self.y_rows, self.back_rows = torch.zeros(n, hidden_size, **f16), torch.zeros(n, hidden_size, **f16)
        self.send_ids, self.recv_ids = torch.full((n,), -1, **i32), torch.full((n,), -1, **i32)

    def dispatch(self):
        W, cap = self.plan.world, self.cap
        if _COALESCE_DISPATCH and W > 1 and self.device.type == "cuda" and torch.cuda.is_current_stream_capturing():
            # under CUDA-graph capture: ids and rows to every peer inside ONE NCCL group (one collective kernel instead of two)
            ops_ = []
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

    def ret(self):
        dist.all_to_all_single(self.back_rows, self.y_rows, group=self.group)

    collectives_per_layer = 3


class PeerExchange:
    """Peer-memory exchange (csrc/ep.cuh): every rank's receive / return buffers live in memory its peers can address.

    `peers` is the list, indexed by rank, of the PeerExchange-shaped objects whose buffers this rank writes into:
      * one process per GPU: build with `PeerExchange.symmetric(...)` -- torch symmetric memory hands out the peer pointers;
      * several VIRTUAL ranks in one process (tests): build each with `PeerExchange.local(...)` then `connect([...])`.
    One instance serves every layer of a stack: layers run one after another on the stream, and the protocol needs no extra
    handshake to reuse the buffers (see csrc/ep.cuh)."""
    kind = "p2p"
    collectives_per_layer = 0

    def __init__(self, plan: EPPlan, cap: int, hidden_size: int, device, storage: torch.Tensor):
        self.plan, self.cap, self.H, self.device = plan, cap, hidden_size, torch.device(device)
        self.storage = storage
        W = plan.world
        o = self.layout(W, cap, hidden_size)
        view = lambda name, dt, shape: storage[o[name][0]:o[name][0] + o[name][1]].view(dt).view(*shape)
        self.recv_rows = view("recv_rows", torch.float16, (W * cap, hidden_size))
        self.back_rows = view("back_rows", torch.float16, (W * cap, hidden_size))
        self.recv_ids = view("recv_ids", torch.int32, (W * cap,))
        self.flags = view("flags", torch.int64, (2, W))
        self.y_rows = torch.zeros(W * cap, hidden_size, dtype=torch.float16, device=self.device)
        self.ctr = torch.zeros(8, dtype=torch.int32, device=self.device)     # [0:2] dispatch, [2:4] return, [4] / [6] the two waits
        self.status = torch.zeros(4, dtype=torch.int32, device=self.device)
        self.recv_ids.fill_(-1)
        self.flags.zero_()
        self._tables = None
        self._keep = None
        self.fused_wait = False   # one process per GPU (symmetric()): the dispatch / return kernels wait for the peers themselves

    @staticmethod
    def layout(W: int, cap: int, H: int) -> dict:
        al = lambda v: (v + 255) // 256 * 256
        o, off = {}, 0
        for name, nbytes in (("recv_rows", W * cap * H * 2), ("back_rows", W * cap * H * 2), ("recv_ids", W * cap * 4), ("flags", 2 * W * 8)):
            o[name] = (off, nbytes)
            off += al(nbytes)
        o["total"] = off
        return o

    @classmethod
    def local(cls, plan, cap, hidden_size, device):
        total = cls.layout(plan.world, cap, hidden_size)["total"]
        return cls(plan, cap, hidden_size, device, torch.zeros(total, dtype=torch.uint8, device=device))

    @classmethod
    def symmetric(cls, plan, cap, hidden_size, device, group=None):
        """One process per GPU: allocate the buffers in torch symmetric memory and exchange peer pointers.  Raises if symmetric
        memory is unavailable (the caller then falls back to NcclExchange and says so)."""
        import torch.distributed._symmetric_memory as symm_mem
        total = cls.layout(plan.world, cap, hidden_size)["total"]
        storage = symm_mem.empty(total, dtype=torch.uint8, device=device)
        storage.zero_()
        grp = group if group is not None else dist.group.WORLD
        try:   # older torch builds want the group registered first; newer ones do it inside rendezvous
            symm_mem.enable_symm_mem_for_group(grp.group_name)
        except Exception:   # noqa: BLE001
            pass
        hdl = symm_mem.rendezvous(storage, grp)
        self = cls(plan, cap, hidden_size, device, storage)
        self._keep = hdl
        self.fused_wait = True
        torch.cuda.synchronize(device)
        dist.barrier(group=group)          # every rank has zeroed its flags before anyone writes one
        self._set_tables([int(p) for p in hdl.buffer_ptrs])
        return self

    def connect(self, peers: list["PeerExchange"]) -> None:
        self._set_tables([int(p.storage.data_ptr()) for p in peers])

    def _set_tables(self, bases: list[int]) -> None:
        import ctypes as C
        W = self.plan.world
        if len(bases) != W:
            raise ValueError("one base pointer per rank is needed")
        o = self.layout(W, self.cap, self.H)
        arr = lambda name, extra=0: (C.c_void_p * W)(*[b + o[name][0] + extra for b in bases])
        self._tables = {"recv_rows": arr("recv_rows"), "back_rows": arr("back_rows"), "recv_ids": arr("recv_ids"),
                        "flags_dispatch": arr("flags"), "flags_return": arr("flags", W * 8)}

    # the two halves of each direction are separate calls so that tests can interleave virtual ranks
    def dispatch_from(self, x, ids, b, E, k, pair_slot):
        ops.ep_dispatch_p2p(x, ids, b, E, self.plan.world, self.plan.rank, self.cap, self._tables["recv_rows"], self._tables["recv_ids"],
                            self._tables["flags_dispatch"], pair_slot, self.ctr[0:2], self.status,
                            *((self.flags[0], self.ctr[4:6]) if self.fused_wait else ()))

    def wait_dispatch(self):
        if not self.fused_wait:
            ops.ep_wait(self.flags[0], self.plan.world, self.ctr[4:6], self.status)

    def ret(self):
        ops.ep_return_p2p(self.y_rows, self.recv_ids, self.plan.world, self.plan.rank, self.cap, self._tables["back_rows"],
                          self._tables["flags_return"], self.ctr[2:4],
                          *((self.flags[1], self.ctr[6:8], self.status) if self.fused_wait else ()))

    def wait_return(self):
        if not self.fused_wait:
            ops.ep_wait(self.flags[1], self.plan.world, self.ctr[6:8], self.status)


class EPMoELayer:
    """One MoE layer, expert-parallel.  `w13` / `w2` hold only this rank's experts (in global order); `w13_bias` is this rank's
    slice fp16 [E_local, 2I], `w2_bias` the FULL table fp16 [E, H] (added at home so the result equals the single-GPU layer)."""

    def __init__(self, w13, w2, plan: EPPlan, top_k: int, hidden_size: int, max_tokens: int, device,
                 group=None, device_ops=None, capacity_factor: float | None = None, regime: int | None = None,
                 w13_bias=None, w2_bias=None, exchange=None):
        self.w13, self.w2, self.plan, self.k, self.H = w13, w2, plan, top_k, hidden_size
        self.max_tokens, self.group = max_tokens, group
        self.w13_bias, self.w2_bias = w13_bias, w2_bias
        self.device = torch.device(device)
        if device_ops is None:
            if self.device.type != "cuda":
                raise RuntimeError("EPMoELayer needs a CUDA device: there is no CPU fallback on the product path")
            device_ops = CudaEPOps()
        self.ops = device_ops
        if exchange is None:
            exchange = NcclExchange(plan, plan.capacity(max_tokens, top_k, capacity_factor), hidden_size, self.device, group)
        self.x = exchange
        self.cap = exchange.cap
        plan.check_owner_slots(self.cap, hidden_size)
        # The owner's kernels are chosen for the rows it EXPECTS (every rank sends about T*k/W rows to each peer, so T*k arrive),
        # not for the W*cap slots of its receive buffer, which is mostly empty
        self.rows_hint = min(plan.world * self.cap, max_tokens * top_k)
        self.regime = regime if regime is not None else _lib.REGIME_AUTO
        self.regime_resolved = self.regime
        if self.regime == _lib.REGIME_AUTO and self.device.type == "cuda" and w13 is not None:
            nested = getattr(device_ops, "nested", None)
            nested_both = all((nested if nested is not None else w.absmax is None) for w in (w13, w2))
            fused_act = getattr(device_ops, "act", _lib.ACT_SWIGLU_OAI) == _lib.ACT_SWIGLU_OAI
            self.regime_resolved = ops.auto_regime(self.rows_hint, 1, w13.num_experts, nested_both and fused_act)
        self.pair_slot = torch.full((max_tokens * top_k,), -1, dtype=torch.int32, device=self.device)
        self.status = getattr(exchange, "status", None)
        if self.status is None:
            self.status = torch.zeros(4, dtype=torch.int32, device=self.device)

    # buffer views the tests and the bench reach for
    recv_rows = property(lambda self: self.x.recv_rows)
    recv_ids = property(lambda self: self.x.recv_ids)
    y_rows = property(lambda self: self.x.y_rows)
    back_rows = property(lambda self: self.x.back_rows)
    send_rows = property(lambda self: self.x.send_rows)
    send_ids = property(lambda self: self.x.send_ids)

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
        if self.x.kind == "p2p":
            self.x.dispatch_from(x, topk_ids, b, self.plan.num_experts, self.k, self.pair_slot)
        else:
            self.ops.pack(x, topk_ids, b, self.plan.num_experts, self.plan.world, self.cap, self.x.send_rows, self.x.send_ids,
                          self.pair_slot, self.status)
        self._weights, self._ids, self._T = topk_weights, topk_ids, T
        return topk_ids, topk_weights

    def dispatch(self):
        if self.x.kind == "p2p":
            self.x.wait_dispatch()
        else:
            self.x.dispatch()

    def compute(self):
        self.ops.expert_rows(self.x.recv_rows, self.x.recv_ids, self.w13, self.w2, self.x.y_rows, self.regime,
                             **({"w13_bias": self.w13_bias, "rows_hint": self.rows_hint} if isinstance(self.ops, CudaEPOps) else {}))

    def ret(self):
        self.x.ret()
        if self.x.kind == "p2p":
            self.x.wait_return()

    def finish(self, out=None):
        extra = {"ids": self._ids, "w2_bias": self.w2_bias} if self.w2_bias is not None else {}
        return self.ops.combine(self.x.back_rows, self._weights, self.pair_slot[: self._T * self.k], out, **extra)

    def forward(self, x, router_logits=None, topk_ids=None, topk_weights=None, out=None):
        self.prepare(x, router_logits, topk_ids, topk_weights)
        self.dispatch()
        self.compute()
        self.ret()
        return self.finish(out)

    __call__ = forward

    def check_overflow(self) -> None:
        """Host-side check (synchronises): raise if a capacity below the worst case dropped pairs, or a peer never arrived."""
        st = int(self.status[0].item())
        if st & 1:
            raise RuntimeError("expert-parallel dispatch overflowed its per-peer capacity; raise capacity_factor")
        if st & 8:
            raise RuntimeError("expert-parallel peer exchange timed out waiting for a peer's flag")

    def wire_bytes(self) -> dict:
        """Bytes this rank puts on the wire per forward.  NCCL: the fixed-capacity buffers, (W-1)/W of them leave the GPU;
        peer memory: only the rows that exist (T*k rows out, T*k back, (W-1)/W of them remote) plus ids and flags."""
        W = self.plan.world
        off = (W - 1) / W
        if self.x.kind == "p2p":
            rows = self.max_tokens * self.k
            return {"dispatch": int((rows * self.H * 2 + W * self.cap * 4 + W * 8) * off), "combine": int((rows * self.H * 2 + W * 8) * off)}
        n = W * self.cap
        return {"dispatch": int(n * (self.H * 2 + 4) * off), "combine": int(n * self.H * 2 * off)}


class ReplicatedMoELayer:
    """SURVEY 8(e)'s second mode, for callers whose activations are ALREADY replicated on every rank (tensor-parallel attention):
    no dispatch at all -- every rank routes all T tokens (the router is replicated, so all ranks agree), runs only ITS experts on
    the pairs routed to them (non-local pairs are masked with -1, which the bucket kernel drops), and the partial [T, H] outputs are
    summed across ranks with one all-reduce (NCCL; in-switch NVLS reduction where the fabric offers it).  The sum is over fp16
    partials, so unlike `EPMoELayer` the result is NOT bit-identical to one GPU -- it meets the MoE tolerance (tests).  Wire bytes:
    T*H*2 per rank each way, against T*k/W rows for the dispatch mode; it pays only when the all-gather it saves the attention side
    is worth more, which is why it is the fallback and not the default.  `w2_bias`: this rank's slice [E_local, H]."""

    def __init__(self, w13, w2, plan: EPPlan, top_k: int, device, group=None, device_ops=None, nested: bool | None = None,
                 regime: int | None = None, w13_bias=None, w2_bias=None):
        self.w13, self.w2, self.plan, self.k, self.group = w13, w2, plan, top_k, group
        self.device = torch.device(device)
        if device_ops is None and self.device.type != "cuda":
            raise RuntimeError("ReplicatedMoELayer needs a CUDA device: there is no CPU fallback on the product path")
        self.ops, self.nested = device_ops, nested
        self.regime = regime if regime is not None else _lib.REGIME_AUTO
        self.w13_bias, self.w2_bias = w13_bias, w2_bias

    def local_ids(self, topk_ids: torch.Tensor) -> torch.Tensor:
        lo, n = self.plan.local_experts.start, self.plan.experts_per_rank
        loc = topk_ids.to(torch.int32) - lo
        return torch.where((loc >= 0) & (loc < n), loc, torch.full_like(loc, -1))

    def partial(self, x, router_logits=None, topk_ids=None, topk_weights=None, out=None):
        """This rank's share of the layer output (before the all-reduce): fp16 [T, H], zero rows for tokens none of whose experts
        live here."""
        if router_logits is not None:
            topk_ids, topk_weights = (self.ops.route(router_logits, self.k) if self.ops is not None else ops.route_topk(router_logits, self.k))
        elif topk_ids is None or topk_weights is None:
            raise ValueError("give router_logits or (topk_ids, topk_weights)")
        loc = self.local_ids(topk_ids)
        if self.ops is not None:
            return self.ops.local_forward(x, loc, topk_weights, out)
        return ops.moe_forward(x, self.w13, self.w2, self.k, topk_ids=loc, topk_weights=topk_weights, regime=self.regime,
                               nested=self.nested, out=out, w13_bias=self.w13_bias, w2_bias=self.w2_bias)

    def forward(self, x, router_logits=None, topk_ids=None, topk_weights=None, out=None):
        part = self.partial(x, router_logits, topk_ids, topk_weights, out)
        if self.plan.world > 1:
            dist.all_reduce(part, op=dist.ReduceOp.SUM, group=self.group)
        return part

    __call__ = forward


def make_exchange(plan: EPPlan, top_k: int, hidden_size: int, max_tokens: int, device, group=None, backend: str = "auto",
                  capacity_factor: float | None = None):
    """The exchange a stack of EP layers shares.  backend: "p2p" (peer memory), "nccl", or "auto" = p2p when torch symmetric
    memory can be set up on every rank, else nccl.  Returns (exchange, note).
    capacity_factor applies to BOTH backends.  None = the worst case max_tokens * top_k rows per peer: drop-free, and over peer
    memory free on the wire (only rows that exist cross the link) -- but every owner then buckets and runs world * max_tokens * top_k
    slots and holds O(that * hidden_size) bytes of buffers, which is fine at decode sizes and not at prefill sizes (see
    EPPlan.check_owner_slots).  A factor bounds both at factor * mean load; pairs past it are dropped AND flagged
    (NF4MOE_STATUS_EP_OVERFLOW -> EPMoELayer.check_overflow raises)."""
    if backend not in ("auto", "p2p", "nccl"):
        raise ValueError("backend must be auto, p2p or nccl")
    plan.check_owner_slots(plan.capacity(max_tokens, top_k, capacity_factor), hidden_size)
    note = ""
    if backend in ("auto", "p2p") and torch.device(device).type == "cuda":
        ok = 1
        try:
            ex = PeerExchange.symmetric(plan, plan.capacity(max_tokens, top_k, capacity_factor), hidden_size, device, group)
        except Exception as exc:   # noqa: BLE001 -- any failure to set up peer memory selects the collective path, loudly
            ok, note = 0, f"peer memory unavailable ({type(exc).__name__}: {exc})"
            if backend == "p2p":
                raise
        flag = torch.tensor([ok], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 1:
            return ex, "peer memory (torch symmetric memory over NVLink)" + (
                f"; capacity {ex.cap} rows per peer (factor {capacity_factor})" if capacity_factor is not None else "")
        note = note or "peer memory unavailable on another rank"
    return NcclExchange(plan, plan.capacity(max_tokens, top_k, capacity_factor), hidden_size, device, group), ("nccl all-to-all" + (f"; {note}" if note else ""))


def shard_experts(w: NF4ExpertWeights, plan: EPPlan) -> NF4ExpertWeights:
    """This rank's slice of a full expert set (views, no copy)."""
    r = plan.local_experts
    s = slice(r.start, r.stop)
    pick = lambda t: None if t is None else t[s]
    code2 = w.code2 if (w.code2 is None or w.code2.dim() == 1) else w.code2[s]
    return NF4ExpertWeights(w.packed[s], w.rows, w.cols, pick(w.absmax), pick(w.absmax_u8), pick(w.absmax2), code2,
                            pick(w.offset))
