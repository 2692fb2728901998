"""Tensor-level wrappers over the C ABI (one Python function per entry point of include/nf4moe.h).

All tensors must live on one CUDA device; calls are asynchronous on torch's current stream of that device and do not
synchronise.  Nothing here falls back to PyTorch math.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .quant import NF4ExpertWeights, _require_cuda, _stream_ptr

class WorkspacePool:
    """Scratch buffers for `moe_forward` / `expert_rows`, one per (device, CUDA stream, layer geometry).

    The C ABI's threading contract is "each caller passes its own stream and workspace" (include/nf4moe.h): two streams that ran
    MoE layers through one shared buffer would race, so the pool is keyed by the stream a call is issued on.  A buffer is sized
    for the next power of two of the token count (>= 64).  When a larger batch arrives a NEW buffer is allocated and the old
    one is kept alive: a CUDA graph captured earlier has the old pointer baked in, and freeing it would let later replays
    scribble over memory the caching allocator has handed to someone else.  Buffers are zero-filled once: their first word is
    the status word the kernels OR error bits into (`workspace_status`)."""

    def __init__(self):
        self._live: dict[tuple, tuple[int, torch.Tensor]] = {}
        self._retired: list[torch.Tensor] = []

    def get(self, device, T: int, E: int, H: int, I: int, k: int) -> torch.Tensor:
        device = torch.device(device)
        key = (device.index if device.index is not None else torch.cuda.current_device(),
               torch.cuda.current_stream(device).cuda_stream, E, H, I, k)
        cap = 64
        while cap < T:
            cap *= 2
        cur = self._live.get(key)
        if cur is None or cur[0] < cap:
            if cur is not None:
                self._retired.append(cur[1])
            buf = torch.zeros(workspace_bytes(cap, E, H, I, k), dtype=torch.uint8, device=device)
            self._live[key] = (cap, buf)
            return buf
        return cur[1]

    def buffers(self) -> list[torch.Tensor]:
        return [b for _, b in self._live.values()] + list(self._retired)

    def release_retired(self) -> None:
        """Drop superseded buffers (only safe when no CUDA graph captured with them is replayed any more)."""
        self._retired.clear()


_pool = WorkspacePool()


def workspace_bytes(T: int, E: int, H: int, I: int, k: int) -> int:
    return int(_lib.load_library().nf4moe_workspace_bytes(T, E, H, I, k))


def get_workspace(device, T: int, E: int, H: int, I: int, k: int) -> torch.Tensor:
    """The calling stream's scratch buffer for this geometry (see WorkspacePool)."""
    return _pool.get(device, T, E, H, I, k)


def new_workspace(device, T: int, E: int, H: int, I: int, k: int) -> torch.Tensor:
    """A private, zero-filled workspace for callers that manage their own (one per captured CUDA graph, say)."""
    return torch.zeros(workspace_bytes(T, E, H, I, k), dtype=torch.uint8, device=device)


def workspace_status(workspace: torch.Tensor) -> int:
    """NF4MOE_STATUS_* bits a kernel raised in this workspace (synchronises).  0 = every kernel ran."""
    return int(workspace[:4].view(torch.int32).item())


def check_status(workspace: torch.Tensor | None = None) -> None:
    """Raise if a kernel refused to run (workspace=None: every pooled buffer).  Synchronises; call it where you sync anyway."""
    bufs = [workspace] if workspace is not None else _pool.buffers()
    for b in bufs:
        st = workspace_status(b)
        if st:
            raise _lib.Nf4MoeError(-100, f"kernel status word = {st:#x} (2: decode tcgen05 kernel could not place its pipeline in "
                                         f"shared memory, 4: prefill tcgen05 kernel likewise, 16: the merged decode launch was not fully "
                                         f"resident and gave up waiting); the output is not valid")


def route_topk(logits: torch.Tensor, k: int) -> tuple[torch.Tensor, torch.Tensor]:
    """(ids int32 [T,k], weights fp32 [T,k]); P/unsloth_to_vllm_converter.py:291-292."""
    _require_cuda(logits, "router_logits")
    T, E = logits.shape
    dt = _lib.LOGITS_DTYPE[str(logits.dtype).replace("torch.", "")]
    ids = torch.empty(T, k, dtype=torch.int32, device=logits.device)
    w = torch.empty(T, k, dtype=torch.float32, device=logits.device)
    with torch.cuda.device(logits.device):
        _lib.check(_lib.load_library().nf4moe_route_topk(logits.data_ptr(), dt, T, E, k, ids.data_ptr(), w.data_ptr(),
                                                         _stream_ptr(logits.device)))
    return ids, w


def bucket(ids: torch.Tensor, E: int, want_tiles: bool = True, tile_rows: int = 0,
           into: dict[str, torch.Tensor] | None = None) -> dict[str, torch.Tensor]:
    """Stable counting sort of the (token,slot) pairs by expert -> offsets / sorted_pairs / active / meta / tiles.
    `into`: a dict returned by an earlier call of the same shape, rewritten in place (CUDA-graph replays; the merged w13+w2
    launch needs its tile list rewritten before every call)."""
    _require_cuda(ids, "topk_ids")
    T, k = ids.shape
    dev = ids.device
    n = T * k
    out = into if into is not None else {
        "expert_offsets": torch.empty(E + 1, dtype=torch.int32, device=dev),
        "sorted_pairs": torch.full(((n + 255) // 256 * 256 + 256,), -1, dtype=torch.int32, device=dev),
        "active_experts": torch.empty(E, dtype=torch.int32, device=dev),
        "meta": torch.zeros(16, dtype=torch.int32, device=dev),
        "tile_list": torch.full((4 * (n // 16 + E + 2),), -1, dtype=torch.int32, device=dev) if want_tiles else None,
    }
    with torch.cuda.device(dev):
        _lib.check(_lib.load_library().nf4moe_bucket(
            ids.data_ptr(), T, k, E, out["expert_offsets"].data_ptr(), out["sorted_pairs"].data_ptr(),
            out["active_experts"].data_ptr(), out["meta"].data_ptr(),
            out["tile_list"].data_ptr() if want_tiles else None, tile_rows, _stream_ptr(dev)))
    return out


def auto_regime(T: int, k: int, E: int, nested_absmax: bool = False) -> int:
    """The regime REGIME_AUTO resolves to (nf4moe_auto_regime[_for]).  nested_absmax: both weight sets are passed with the nested
    absmax and the activation is the fused gpt-oss one (the decode kernel then has 64-row token tiles and a wider regime)."""
    return int(_lib.load_library().nf4moe_auto_regime_for(T, k, E, 1 if nested_absmax else 0))


def permute_rows(x: torch.Tensor, b: dict, k: int) -> torch.Tensor:
    T, H = x.shape
    xs = torch.empty((T * k + 255) // 256 * 256 + 256, H, dtype=torch.float16, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load_library().nf4moe_permute_rows(x.data_ptr(), b["sorted_pairs"].data_ptr(), b["meta"].data_ptr(),
                                                           T, k, H, xs.data_ptr(), _stream_ptr(x.device)))
    return xs


def combine(y: torch.Tensor, weights: torch.Tensor, ids: torch.Tensor | None, E: int) -> torch.Tensor:
    T, k = weights.shape
    H = y.shape[-1]
    out = torch.empty(T, H, dtype=torch.float16, device=y.device)
    with torch.cuda.device(y.device):
        _lib.check(_lib.load_library().nf4moe_combine(y.data_ptr(), weights.data_ptr(),
                                                      None if ids is None else ids.data_ptr(), E, T, k, H,
                                                      out.data_ptr(), _stream_ptr(y.device)))
    return out


def expert_w13(x: torch.Tensor, w13: NF4ExpertWeights, b: dict, T: int, k: int, act: int = _lib.ACT_SWIGLU_OAI,
               nested: bool | None = None, regime: int = _lib.REGIME_GEMV) -> torch.Tensor:
    """h_sorted fp16 [rows_padded, I] (act fused) or [rows_padded, 2I] raw for ACT_CHUNK_SILU."""
    lib = _lib.load_library()
    E = w13.num_experts
    width = w13.rows // 2 if act == _lib.ACT_SWIGLU_OAI else w13.rows
    h = torch.zeros((T * k + 255) // 256 * 256 + 256, width, dtype=torch.float16, device=x.device)
    q = w13.c_struct(nested)
    with torch.cuda.device(x.device):
        if regime in (_lib.REGIME_GEMM, _lib.REGIME_GEMM_FAST):
            mode = _lib.GEMM_FAST if regime == _lib.REGIME_GEMM_FAST else _lib.GEMM_EXACT
            _lib.check(lib.nf4moe_gemm_w13_swiglu(x.data_ptr(), C.byref(q), b["expert_offsets"].data_ptr(),
                                                  b["meta"].data_ptr(), b["tile_list"].data_ptr(), T, k, E, act, mode,
                                                  h.data_ptr(), _stream_ptr(x.device)))
        else:
            _lib.check(lib.nf4moe_gemv_w13_swiglu(x.data_ptr(), C.byref(q), b["expert_offsets"].data_ptr(),
                                                  b["sorted_pairs"].data_ptr(), b["active_experts"].data_ptr(),
                                                  b["meta"].data_ptr(), T, k, E, act, h.data_ptr(),
                                                  _stream_ptr(x.device)))
    return h


def expert_w2(h: torch.Tensor, w2: NF4ExpertWeights, b: dict, T: int, k: int, nested: bool | None = None,
              regime: int = _lib.REGIME_GEMV) -> torch.Tensor:
    """y fp16 [T*k, H] in (token, slot) order."""
    lib = _lib.load_library()
    E = w2.num_experts
    y = torch.zeros(T * k, w2.rows, dtype=torch.float16, device=h.device)
    q = w2.c_struct(nested)
    with torch.cuda.device(h.device):
        if regime in (_lib.REGIME_GEMM, _lib.REGIME_GEMM_FAST):
            mode = _lib.GEMM_FAST if regime == _lib.REGIME_GEMM_FAST else _lib.GEMM_EXACT
            _lib.check(lib.nf4moe_gemm_w2(h.data_ptr(), C.byref(q), b["expert_offsets"].data_ptr(),
                                          b["sorted_pairs"].data_ptr(), b["meta"].data_ptr(), b["tile_list"].data_ptr(),
                                          T, k, E, mode, y.data_ptr(), _stream_ptr(h.device)))
        else:
            _lib.check(lib.nf4moe_gemv_w2(h.data_ptr(), C.byref(q), b["expert_offsets"].data_ptr(),
                                          b["sorted_pairs"].data_ptr(), b["active_experts"].data_ptr(),
                                          b["meta"].data_ptr(), T, k, E, y.data_ptr(), _stream_ptr(h.device)))
    return y


def expert_w13_w2(xs: torch.Tensor, w13: NF4ExpertWeights, w2: NF4ExpertWeights, b: dict, T: int, k: int,
                  nested: bool | None = None, regime: int = _lib.REGIME_GEMM_FAST, w13_bias: torch.Tensor | None = None,
                  h: torch.Tensor | None = None, y: torch.Tensor | None = None) -> tuple[torch.Tensor, torch.Tensor]:
    """Both expert products in ONE launch of the decode-regime tcgen05 kernel (nf4moe_gemm_w13_w2): h fp16 [T*k, I] in sorted
    order, y fp16 [T*k, H] in (token, slot) order.  `b` must come fresh from bucket() (tile rows <= 32) for every call."""
    lib = _lib.load_library()
    E = w13.num_experts
    I = w13.rows // 2
    h = torch.zeros(T * k, I, dtype=torch.float16, device=xs.device) if h is None else h
    y = torch.zeros(T * k, w2.rows, dtype=torch.float16, device=xs.device) if y is None else y
    q13, q2 = w13.c_struct(nested), w2.c_struct(nested)
    mode = _lib.GEMM_FAST if regime == _lib.REGIME_GEMM_FAST else _lib.GEMM_EXACT
    with torch.cuda.device(xs.device):
        _lib.check(lib.nf4moe_gemm_w13_w2(xs.data_ptr(), C.byref(q13), None if w13_bias is None else w13_bias.data_ptr(),
                                          h.data_ptr(), C.byref(q2), y.data_ptr(), b["sorted_pairs"].data_ptr(),
                                          b["meta"].data_ptr(), b["tile_list"].data_ptr(), T, k, E, _lib.ACT_SWIGLU_OAI, mode,
                                          _stream_ptr(xs.device)))
    return h, y


def cast_f16_bf16(src: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp16 -> bf16 (round to nearest even) or bf16 -> fp16 (exact inside the fp16 range, saturating outside): nf4moe_cast_f16_bf16."""
    _require_cuda(src, "src")
    if src.dtype not in (torch.float16, torch.bfloat16) or not src.is_contiguous():
        raise TypeError("cast_f16_bf16 takes a contiguous fp16 or bf16 tensor")
    to_bf16 = src.dtype == torch.float16
    if out is None:
        out = torch.empty_like(src, dtype=torch.bfloat16 if to_bf16 else torch.float16)
    with torch.cuda.device(src.device):
        _lib.check(_lib.load_library().nf4moe_cast_f16_bf16(src.data_ptr(), out.data_ptr(), src.numel(), 1 if to_bf16 else 0,
                                                            _stream_ptr(src.device)))
    return out


def moe_forward(x: torch.Tensor, w13: NF4ExpertWeights, w2: NF4ExpertWeights, top_k: int,
                router_logits: torch.Tensor | None = None, topk_ids: torch.Tensor | None = None,
                topk_weights: torch.Tensor | None = None, act: int = _lib.ACT_SWIGLU_OAI,
                regime: int = _lib.REGIME_AUTO, nested: bool | None = None, out: torch.Tensor | None = None,
                workspace: torch.Tensor | None = None, return_routing: bool = False,
                w13_bias: torch.Tensor | None = None, w2_bias: torch.Tensor | None = None):
    """The whole layer (nf4moe_moe_forward[_bias]): x fp16 [T,H] -> fp16 [T,H].  w13_bias fp16 [E, 2I] / w2_bias fp16 [E, H]
    (optional) are gpt-oss's expert biases."""
    _require_cuda(x, "hidden_states")
    if x.dtype == torch.bfloat16:
        # bf16 callers (P/bnb_moe_4bit.py:35-37 lists bf16 beside fp16): widened exactly to fp16, computed as for fp16 (fp32 accumulation),
        # rounded to bf16 once at the end -- two elementwise kernels of this library around the same launch chain
        if out is not None and out.dtype != torch.bfloat16:
            raise TypeError("out must be bf16 for bf16 hidden_states")
        res = moe_forward(cast_f16_bf16(x.contiguous()), w13, w2, top_k, router_logits=router_logits, topk_ids=topk_ids,
                          topk_weights=topk_weights, act=act, regime=regime, nested=nested, workspace=workspace,
                          return_routing=return_routing, w13_bias=w13_bias, w2_bias=w2_bias)
        o16 = res[0] if return_routing else res
        out = cast_f16_bf16(o16, out)
        return (out, res[1], res[2]) if return_routing else out
    if x.dtype != torch.float16:
        raise TypeError("hidden_states must be fp16 or bf16 (the reference serves with --dtype float16)")
    lib = _lib.load_library()
    T, H = x.shape
    E, I = w13.num_experts, w13.rows // 2
    dev = x.device
    if out is None:
        out = torch.empty(T, H, dtype=torch.float16, device=dev)
    if workspace is None:
        workspace = get_workspace(dev, T, E, H, I, top_k)
    ldt = 0
    if router_logits is not None:
        _require_cuda(router_logits, "router_logits")
        ldt = _lib.LOGITS_DTYPE[str(router_logits.dtype).replace("torch.", "")]
    else:
        if topk_ids is None or topk_weights is None:
            raise ValueError("give router_logits or (topk_ids, topk_weights)")
        if topk_ids.dtype != torch.int32:
            topk_ids = topk_ids.to(torch.int32)
        if topk_weights.dtype != torch.float32:
            topk_weights = topk_weights.to(torch.float32)
        topk_ids, topk_weights = topk_ids.contiguous(), topk_weights.contiguous()
    ids_out = w_out = None
    if return_routing and router_logits is not None:
        ids_out = torch.empty(T, top_k, dtype=torch.int32, device=dev)
        w_out = torch.empty(T, top_k, dtype=torch.float32, device=dev)
    q13, q2 = w13.c_struct(nested), w2.c_struct(nested)
    for name, b, shape in (("w13_bias", w13_bias, (E, 2 * I)), ("w2_bias", w2_bias, (E, H))):
        if b is not None:
            _require_cuda(b, name)
            if b.dtype != torch.float16 or tuple(b.shape) != shape or not b.is_contiguous():
                raise ValueError(f"{name} must be a contiguous fp16 tensor of shape {shape}")
    with torch.cuda.device(dev):
        _lib.check(lib.nf4moe_moe_forward_bias(
            x.data_ptr(), None if router_logits is None else router_logits.data_ptr(), ldt,
            None if topk_ids is None else topk_ids.data_ptr(), None if topk_weights is None else topk_weights.data_ptr(),
            C.byref(q13), C.byref(q2), None if w13_bias is None else w13_bias.data_ptr(),
            None if w2_bias is None else w2_bias.data_ptr(), out.data_ptr(), T, E, H, I, top_k, act, regime, workspace.data_ptr(),
            workspace.numel(), None if ids_out is None else ids_out.data_ptr(),
            None if w_out is None else w_out.data_ptr(), _stream_ptr(dev)))
    if return_routing:
        return out, (ids_out if ids_out is not None else topk_ids), (w_out if w_out is not None else topk_weights)
    return out


def last_launch_count() -> int:
    return int(_lib.load_library().nf4moe_last_launch_count())


def set_merged_decode(enabled: bool) -> bool:
    """nf4moe_set_merged_decode: False = the decode regime issues w13 and w2 as two launches (safe on a GPU this process shares --
    MPS percentage, SM partitions, a concurrent kernel that waits for this stream -- where the single merged launch may not get all
    its CTAs resident and would time out and trap); True (default) = one launch.  Results are bit-identical.  Returns the previous
    setting; process-wide; set it before capturing CUDA graphs."""
    return bool(_lib.load_library().nf4moe_set_merged_decode(1 if enabled else 0))


# ---- expert parallelism (nf4moe_ep_pack / nf4moe_expert_rows / nf4moe_combine_indexed) ---------------------------------
def ep_pack(x: torch.Tensor, ids: torch.Tensor, b: dict, E: int, world: int, cap: int, send_rows: torch.Tensor,
            send_ids: torch.Tensor, pair_slot: torch.Tensor, status: torch.Tensor) -> None:
    """Write the dispatch all-to-all's send buffer straight from x (no separate pack pass)."""
    T, H = x.shape
    k = ids.shape[1]
    with torch.cuda.device(x.device):
        _lib.check(_lib.load_library().nf4moe_ep_pack(
            x.data_ptr(), ids.data_ptr(), b["expert_offsets"].data_ptr(), b["sorted_pairs"].data_ptr(), T, k, E, H, world,
            cap, send_rows.data_ptr(), send_ids.data_ptr(), pair_slot.data_ptr(), status.data_ptr(), _stream_ptr(x.device)))


def expert_rows(rows: torch.Tensor, row_expert: torch.Tensor, w13: NF4ExpertWeights, w2: NF4ExpertWeights,
                y_rows: torch.Tensor, act: int = _lib.ACT_SWIGLU_OAI, regime: int = _lib.REGIME_AUTO,
                nested: bool | None = None, workspace: torch.Tensor | None = None,
                w13_bias: torch.Tensor | None = None, rows_hint: int = 0) -> torch.Tensor:
    """y_rows[r] = W2[e] act(W13[e] rows[r] + w13_bias[e]) for e = row_expert[r] in [0, E_local); other rows are left untouched.
    w13_bias: fp16 [E_local, 2I] or None."""
    _require_cuda(rows, "rows")
    n, H = rows.shape
    E, I = w13.num_experts, w13.rows // 2
    if workspace is None:
        workspace = get_workspace(rows.device, n, E, H, I, 1)
    if w13_bias is not None:
        _require_cuda(w13_bias, "w13_bias")
        if w13_bias.dtype != torch.float16 or tuple(w13_bias.shape) != (E, 2 * I):
            raise ValueError(f"w13_bias must be a contiguous fp16 tensor of shape {(E, 2 * I)}")
    q13, q2 = w13.c_struct(nested), w2.c_struct(nested)
    with torch.cuda.device(rows.device):
        _lib.check(_lib.load_library().nf4moe_expert_rows_bias(
            rows.data_ptr(), row_expert.data_ptr(), C.byref(q13), C.byref(q2), None if w13_bias is None else w13_bias.data_ptr(),
            y_rows.data_ptr(), n, rows_hint, E, H, I, act, regime, workspace.data_ptr(), workspace.numel(), _stream_ptr(rows.device)))
    return y_rows


def combine_indexed(y: torch.Tensor, weights: torch.Tensor, pair_slot: torch.Tensor, out: torch.Tensor | None = None,
                    ids: torch.Tensor | None = None, w2_bias: torch.Tensor | None = None):
    """out[t] = fp16(sum_slot w * (y[pair_slot] + w2_bias[ids])); w2_bias is the FULL fp16 [E, H] table, ids the global topk ids."""
    T, k = weights.shape
    H = y.shape[-1]
    if out is None:
        out = torch.empty(T, H, dtype=torch.float16, device=y.device)
    E = 0
    if w2_bias is not None:
        if ids is None:
            raise ValueError("w2_bias needs the global topk ids")
        _require_cuda(w2_bias, "w2_bias")
        if w2_bias.dtype != torch.float16 or w2_bias.dim() != 2 or w2_bias.shape[1] != H:
            raise ValueError("w2_bias must be fp16 [E, H]")
        E = int(w2_bias.shape[0])
    with torch.cuda.device(y.device):
        _lib.check(_lib.load_library().nf4moe_combine_indexed_bias(
            y.data_ptr(), weights.data_ptr(), pair_slot.data_ptr(), None if w2_bias is None else ids.data_ptr(),
            None if w2_bias is None else w2_bias.data_ptr(), E, T, k, H, out.data_ptr(), _stream_ptr(y.device)))
    return out


# ---- peer-memory exchange (nf4moe_ep_dispatch_p2p / nf4moe_ep_wait / nf4moe_ep_return_p2p) ---------------------------------
def ep_dispatch_p2p(x: torch.Tensor, ids: torch.Tensor, b: dict, E: int, world: int, rank: int, cap: int, peer_recv_rows,
                    peer_recv_ids, peer_flags, pair_slot: torch.Tensor, ctr: torch.Tensor, status: torch.Tensor,
                    wait_flags: torch.Tensor | None = None, wait_ctr: torch.Tensor | None = None) -> None:
    """Store this rank's routed rows straight into their owners' receive buffers (peer_* are ctypes arrays of device pointers).
    wait_flags / wait_ctr: also wait, in the same launch, for every peer's rows to land here (one process per GPU only)."""
    _require_cuda(x, "x")
    T, H = x.shape
    k = ids.shape[1]
    with torch.cuda.device(x.device):
        _lib.check(_lib.load_library().nf4moe_ep_dispatch_p2p(
            x.data_ptr(), ids.data_ptr(), b["expert_offsets"].data_ptr(), b["sorted_pairs"].data_ptr(), T, k, E, H, world, rank, cap,
            C.cast(peer_recv_rows, C.c_void_p), C.cast(peer_recv_ids, C.c_void_p), C.cast(peer_flags, C.c_void_p),
            pair_slot.data_ptr(), ctr.data_ptr(), status.data_ptr(), None if wait_flags is None else wait_flags.data_ptr(),
            None if wait_ctr is None else wait_ctr.data_ptr(), _stream_ptr(x.device)))


def ep_return_p2p(y_rows: torch.Tensor, recv_ids: torch.Tensor, world: int, rank: int, cap: int, peer_back_rows, peer_flags,
                  ctr: torch.Tensor, wait_flags: torch.Tensor | None = None, wait_ctr: torch.Tensor | None = None,
                  status: torch.Tensor | None = None) -> None:
    H = y_rows.shape[-1]
    with torch.cuda.device(y_rows.device):
        _lib.check(_lib.load_library().nf4moe_ep_return_p2p(
            y_rows.data_ptr(), recv_ids.data_ptr(), H, world, rank, cap, C.cast(peer_back_rows, C.c_void_p),
            C.cast(peer_flags, C.c_void_p), ctr.data_ptr(), None if wait_flags is None else wait_flags.data_ptr(),
            None if wait_ctr is None else wait_ctr.data_ptr(), None if status is None else status.data_ptr(), _stream_ptr(y_rows.device)))


def ep_wait(flags: torch.Tensor, world: int, ctr: torch.Tensor, status: torch.Tensor) -> None:
    with torch.cuda.device(flags.device):
        _lib.check(_lib.load_library().nf4moe_ep_wait(flags.data_ptr(), world, ctr.data_ptr(), status.data_ptr(),
                                                      _stream_ptr(flags.device)))
