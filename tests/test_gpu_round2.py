"""GPU parity, round 2: the benchmarked shapes against the oracle (full-size 20b layer, T = 64 / 128 / 256, flat and nested
absmax, uniform and skewed routing), the CUDA dequant pinned to the reference-held golden vectors, the EXACT operand of the
decode-regime tcgen05 kernel, the status word, and a REAL vLLM FusedMoE driving the plugin's method.  Every call goes through
the C ABI.  Bars (north_star): dequant bit-exact, router ids bit-exact, MoE output max-abs <= 1e-2 and rel <= 2e-3."""
import hashlib
import importlib
import importlib.util
import os

import numpy as np
import pytest
import torch

from oracle import nf4_oracle as orc

HERE = os.path.dirname(os.path.abspath(__file__))


def _load(name, path):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


make_golden = _load("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
_parity = _load("_gpu_parity_helpers", os.path.join(HERE, "test_gpu_parity.py"))
assert_moe_close, make_layer = _parity.assert_moe_close, _parity.make_layer

pytestmark = pytest.mark.gpu

E20, H20, I20, K20 = 32, 2880, 2880, 4


def _state(w, e):
    return {"packed": w.packed[e].cpu().numpy().reshape(-1), "absmax_u8": w.absmax_u8[e].cpu().numpy(),
            "absmax2": w.absmax2[e].cpu().numpy(), "code2": w.code2.cpu().numpy(), "offset": w.offset[e].item(),
            "shape": (w.rows, w.cols), "blocksize": 64}


@pytest.fixture(scope="module")
def full_layer(pkg, cuda_device):
    """One full-size gpt-oss-20b MoE layer quantised on the device + the oracle's dequantised fp16 copy of every expert (made
    once: ~1.6 GB of host memory, a few seconds)."""
    g = torch.Generator(device=cuda_device).manual_seed(20250805)
    code2 = torch.from_numpy(orc.create_dynamic_map())
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E20, 2 * I20, H20, generator=g, device=cuda_device) * 0.02).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E20, H20, I20, generator=g, device=cuda_device) * 0.02).half(), code2)
    d13 = np.stack([orc.dequant_state(_state(w13, e)) for e in range(E20)])
    d2 = np.stack([orc.dequant_state(_state(w2, e)) for e in range(E20)])
    return w13, w2, d13, d2


def _inputs(dev, T, skewed, seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    x = torch.randn(T, H20, generator=g, device=dev).half()
    lg = torch.randn(T, E20, generator=g, device=dev).float()
    if skewed:   # logits += 2 * onehot(t mod 4): load imbalance, experts 0..3 get most of the rows
        lg[torch.arange(T, device=dev), torch.arange(T, device=dev) % 4] += 2.0
    return x, lg.half()


@pytest.mark.parametrize("routing", ["uniform", "skewed"])
@pytest.mark.parametrize("absmax", ["flat", "nested"])
@pytest.mark.parametrize("T", [64, 128, 256])
def test_full_size_decode_vs_oracle(pkg, cuda_device, full_layer, T, absmax, routing):
    """The shapes the bench times, in the regime the drop-in apply() uses (AUTO -> decode tcgen05 kernel, FAST operand), on BOTH
    absmax forms -- flat fp32 is what vLLM binds (V/model_executor/model_loader/bitsandbytes_loader.py:585-617) and is the form
    on which round 1's kernel silently did nothing once an expert had more than 16 rows.  Compared with the oracle on every
    token; the largest per-expert count is asserted to exceed 16 so the N = 32 tile path is really taken."""
    w13, w2, d13, d2 = full_layer
    x, lg = _inputs(cuda_device, T, routing == "skewed", 1000 + T)
    nested = absmax == "nested"
    ws = pkg.ops.new_workspace(cuda_device, T, E20, H20, I20, K20)
    out, ids, wts = pkg.ops.moe_forward(x, w13, w2, K20, router_logits=lg, nested=nested, workspace=ws, return_routing=True)
    assert pkg.ops.workspace_status(ws) == 0
    rid, _ = orc.route_topk(lg.cpu().numpy(), K20)
    assert np.array_equal(ids.cpu().numpy(), rid)
    counts = np.bincount(rid.reshape(-1), minlength=E20)
    if T >= 128 or routing == "skewed":
        assert counts.max() > 16, counts.max()
    ref = orc.moe_forward_ref(x.cpu().numpy(), lg.cpu().numpy(), d13, d2, k=K20)
    assert_moe_close(out.cpu().numpy(), ref)
    assert float(out.float().abs().max()) > 1e-3          # and it is not a buffer of zeros / stale scratch
    # EXACT operand through the same decode-regime kernel: also inside the tolerance, and not identical to FAST
    out_e = pkg.ops.moe_forward(x, w13, w2, K20, topk_ids=ids, topk_weights=wts, nested=nested, workspace=ws, regime=pkg.REGIME_GEMM)
    assert pkg.ops.workspace_status(ws) == 0
    assert_moe_close(out_e.cpu().numpy(), ref)


@pytest.mark.parametrize("absmax", ["flat", "nested"])
@pytest.mark.parametrize("T", [1, 2, 4, 8])
def test_full_size_gemv_regime_vs_oracle(pkg, cuda_device, full_layer, T, absmax):
    """Batch 1..8 at the REAL shape (E = 32, H = I = 2880), straight against the oracle: this is what a gpt-oss decode step and the
    bench's decode_sweep execute -- the compile-time K = 2880 instantiation of gemv_kernel and, with router logits given, the
    single-launch route_bucket_small_kernel -- and B*S = 1 is literally the reference's loop (8 GEMVs per token,
    P/unsloth_to_vllm_converter.py:298-323).  Router ids bit-exact, output inside the MoE tolerance on both absmax forms, and the
    AUTO regime must really be the GEMV here."""
    w13, w2, d13, d2 = full_layer
    assert pkg.ops.auto_regime(T, K20, E20, absmax == "nested") == pkg.REGIME_GEMV
    x, lg = _inputs(cuda_device, T, False, 4000 + T)
    ws = pkg.ops.new_workspace(cuda_device, T, E20, H20, I20, K20)
    out, ids, wts = pkg.ops.moe_forward(x, w13, w2, K20, router_logits=lg, nested=absmax == "nested", workspace=ws, return_routing=True)
    assert pkg.ops.workspace_status(ws) == 0
    rid, rw = orc.route_topk(lg.cpu().numpy(), K20)
    assert np.array_equal(ids.cpu().numpy(), rid)
    assert np.allclose(wts.cpu().numpy(), rw, rtol=1e-5, atol=1e-6)
    ref = orc.moe_forward_ref(x.cpu().numpy(), lg.cpu().numpy(), d13, d2, k=K20)
    assert_moe_close(out.cpu().numpy(), ref)
    assert float(out.float().abs().max()) > 1e-3
    # the vLLM-style call (routing pre-computed: separate bucket launch instead of the fused router + bucket) gives the same bits
    out2 = pkg.ops.moe_forward(x, w13, w2, K20, topk_ids=ids, topk_weights=wts, nested=absmax == "nested", workspace=ws)
    assert torch.equal(out, out2)


def test_apply_on_vllm_bound_layer_full_size(pkg, cuda_device, full_layer):
    """B200NF4MoEMethod.apply() on a layer whose parameters carry vLLM's fused, de-nested `bnb_quant_state` (flat fp32 absmax,
    `NF4ExpertWeights.from_vllm_param`), T = 128, full size: the exact call path round 1 got wrong."""
    w13, w2, d13, d2 = full_layer
    T = 128
    method = pkg.B200NF4MoEConfig().get_quant_method(None, "model.layers.0.mlp.experts")
    layer = torch.nn.Module()
    method.create_weights(layer, E20, H20, I20, torch.float16)
    layer.to(cuda_device)
    layer.w13_weight.data = w13.packed
    layer.w2_weight.data = w2.packed

    class QS:   # V/model_executor/model_loader/bitsandbytes_loader.py:673-692
        def __init__(self, absmax):
            self.absmax, self.blocksize, self.quant_type, self.nested = absmax.reshape(-1), 64, "nf4", False
    layer.w13_weight.bnb_quant_state, layer.w2_weight.bnb_quant_state = QS(w13.absmax), QS(w2.absmax)
    method.process_weights_after_loading(layer)
    assert layer._nf4_w13.absmax is not None and not layer._nf4_w13.has_nested
    x, lg = _inputs(cuda_device, T, True, 77)
    ids, wts = pkg.ops.route_topk(lg, K20)
    out = method.apply(layer, x, wts, ids, None, None)
    pkg.ops.check_status()
    ref = orc.moe_forward_ref(x.cpu().numpy(), lg.cpu().numpy(), d13, d2, k=K20)
    assert_moe_close(out.cpu().numpy(), ref)
    assert torch.equal(out, method.apply_legacy(layer, x, lg, K20))


def test_prefill_8192_sample_vs_oracle(pkg, cuda_device, full_layer):
    """BASELINE.json configs[2] at full size: T = 8192 through the prefill grouped GEMM (EXACT operand, what AUTO picks), 64
    sampled tokens against the oracle."""
    w13, w2, d13, d2 = full_layer
    T = 8192
    g = torch.Generator(device=cuda_device).manual_seed(8192)
    x = torch.randn(T, H20, generator=g, device=cuda_device).half()
    lg = torch.randn(T, E20, generator=g, device=cuda_device).half()
    assert pkg.ops.auto_regime(T, K20, E20) == pkg.REGIME_GEMM
    ws = pkg.ops.new_workspace(cuda_device, T, E20, H20, I20, K20)
    out = pkg.ops.moe_forward(x, w13, w2, K20, router_logits=lg, nested=True, workspace=ws)
    assert pkg.ops.workspace_status(ws) == 0
    pick = np.linspace(0, T - 1, 64).astype(np.int64)
    ref = orc.moe_forward_ref(x[pick].cpu().numpy(), lg[pick].cpu().numpy(), d13, d2, k=K20)
    assert_moe_close(out[pick].cpu().numpy(), ref)
    out_f = pkg.ops.moe_forward(x, w13, w2, K20, router_logits=lg, nested=False, workspace=ws, regime=pkg.REGIME_GEMM_FAST)
    assert pkg.ops.workspace_status(ws) == 0
    assert_moe_close(out_f[pick].cpu().numpy(), ref)


def test_expert_parallel_prefill_regime_full_size_k(pkg, cuda_device, full_layer):
    """BASELINE.json configs[4]'s prefill leg, the regime it runs in: EPMoELayer with 160-row token tiles at K = 2880 (two virtual
    ranks, 16 full-size experts each, 400 tokens per rank -> ~100 rows per expert, AUTO resolves to the prefill grouped GEMM
    with the exact operand).  Bit-identical to the single-GPU layer in the same regime, and a sample against the oracle."""
    ep = importlib.import_module("gpt-oss-vllm_b200.ep")
    w13, w2, d13, d2 = full_layer
    world, T = 2, 400
    layers, xs, lgs = [], [], []
    for r in range(world):
        plan = ep.EPPlan(r, world, E20)
        layers.append(ep.EPMoELayer(ep.shard_experts(w13, plan), ep.shard_experts(w2, plan), plan, K20, H20, T, cuda_device))
        assert layers[r].regime_resolved == pkg.REGIME_GEMM, "AUTO must pick the prefill grouped GEMM for ~100 rows per expert"
        x, lg = _inputs(cuda_device, T - 7 * r, False, 4096 + r)          # ragged token counts per rank
        xs.append(x); lgs.append(lg)
    for r in range(world):
        layers[r].prepare(xs[r], router_logits=lgs[r])
    cap = layers[0].cap

    def exchange(src_name, dst_name):
        for dst in range(world):
            for src in range(world):
                getattr(layers[dst], dst_name)[src * cap:(src + 1) * cap].copy_(getattr(layers[src], src_name)[dst * cap:(dst + 1) * cap])
    exchange("send_ids", "recv_ids")
    exchange("send_rows", "recv_rows")
    for l in layers:
        l.compute()
    exchange("y_rows", "back_rows")
    for r in range(world):
        out = layers[r].finish()
        layers[r].check_overflow()
        Tr = xs[r].shape[0]
        ws = pkg.ops.new_workspace(cuda_device, Tr, E20, H20, I20, K20)
        single = pkg.ops.moe_forward(xs[r], w13, w2, K20, router_logits=lgs[r], regime=pkg.REGIME_GEMM, workspace=ws)
        assert pkg.ops.workspace_status(ws) == 0
        assert torch.equal(out, single), f"rank {r}: expert-parallel prefill regime differs from the single-GPU layer"
        pick = np.linspace(0, Tr - 1, 12).astype(np.int64)
        ref = orc.moe_forward_ref(xs[r][pick].cpu().numpy(), lgs[r][pick].cpu().numpy(), d13, d2, k=K20)
        assert_moe_close(out[pick].cpu().numpy(), ref)


# ------------------------------------------------------------------------------------------------- golden pins
def _swap_nibbles(packed: np.ndarray) -> np.ndarray:
    return ((packed >> 4) | ((packed & 0x0F) << 4)).astype(np.uint8)


@pytest.mark.parametrize("name,shape,which", [("dequant_unsloth_expert_cpu/gate_up_vec", (5760, 2880), "streaming"),
                                              ("dequant_unsloth_expert_cpu/down_scalar", (2880, 2880), "streaming"),
                                              ("load_and_dequantize_unsloth_weight/down_vec", (2880, 2880), "direct")])
def test_cuda_dequant_reproduces_reference_held_vectors(pkg, cuda_device, golden, name, shape, which):
    """Direct pin of the CUDA dequant kernel to outputs of the reference's OWN Python (P/streaming_unsloth_bnb.py:109-230,
    P/unsloth_direct_loader.py:51-142, run with bnb stubbed by tests/golden/make_golden.py): the reference unpacks low nibble
    first and scales by its own (offset-free) absmax, so with the packed bytes nibble-swapped and that absmax handed over as
    the flat fp32 form, `nf4moe_dequant` must reproduce the committed SHA-256 of the reference's fp16 output bit for bit --
    the NF4 table, the block indexing, the fp32 multiply and the single rounding are thereby pinned to reference-held data,
    independently of the oracle's bnb-semantics mode."""
    _, meta = golden
    m = meta[name]
    st = make_golden.synth_state(m["seed"], m["n_weights"], m["scalar_nested"])
    code = st["nested_quant_map"].astype(np.float32)[st["absmax_u8"].astype(np.int64)]
    if which == "direct":
        absmax = orc.dequantize_absmax_reference(st["absmax_u8"], st["nested_absmax"], st["nested_quant_map"])
    elif st["nested_absmax"].size == 1:
        absmax = (code * np.float32(st["nested_absmax"][0])).astype(np.float32)
    else:   # shape mismatch branch of P/streaming_unsloth_bnb.py:158-171: the mean of nested_absmax (torch fp32 reduction)
        mean = torch.from_numpy(np.ascontiguousarray(st["nested_absmax"], dtype=np.float32)).mean()
        absmax = (torch.from_numpy(code.copy()) * mean).numpy()
    state = {"packed": _swap_nibbles(st["packed"]), "absmax_f32": absmax.astype(np.float32), "shape": shape, "blocksize": 64}
    w = pkg.NF4ExpertWeights.from_states([state], cuda_device)
    out = w.dequantize(nested=False).cpu().numpy()[0]
    flat = np.ascontiguousarray(out).reshape(-1)
    assert list(out.shape) == m["shape"]
    assert flat.view(np.uint16)[np.asarray(m["sample_pos"])].tolist() == m["sample_bits"]
    assert hashlib.sha256(flat.tobytes()).hexdigest() == m["sha256"]


def test_bitsandbytes_cross_check_or_recorded_absence(pkg, cuda_device):
    """SURVEY 8(c) optional cross-check: if bitsandbytes imports on this box, the oracle's bnb-semantics dequant (the mode the
    CUDA kernels are graded against) must equal bnb.functional.dequantize_4bit bit for bit; if it does not import, the absence is
    the recorded outcome and the test is SKIPPED with that reason -- never a green
    pass that checked nothing."""
    try:
        import bitsandbytes as bnb   # noqa: F401
        from bitsandbytes.functional import dequantize_4bit, quantize_4bit
    except Exception as exc:
        pytest.skip(f"bitsandbytes unavailable on this box ({type(exc).__name__}: {exc}); recorded in BASELINE.md section 4")
    g = torch.Generator(device=cuda_device).manual_seed(5)
    w = (torch.randn(256, 512, generator=g, device=cuda_device) * 0.02).half()
    q, state = quantize_4bit(w, blocksize=64, compress_statistics=True, quant_type="nf4")
    ref = dequantize_4bit(q, state).cpu().numpy()
    st = {"packed": q.cpu().numpy().reshape(-1), "absmax_u8": state.absmax.cpu().numpy(),
          "absmax2": state.state2.absmax.cpu().numpy(), "code2": state.state2.code.cpu().numpy(),
          "offset": float(state.offset.item()), "shape": (256, 512), "blocksize": 64}
    assert np.array_equal(orc.dequant_state(st).view(np.uint16), ref.view(np.uint16))
    got = pkg.NF4ExpertWeights.from_states([st], cuda_device).dequantize(nested=True).cpu().numpy()[0]
    assert np.array_equal(got.view(np.uint16), ref.view(np.uint16))


# ------------------------------------------------------------------------------------------------- decode kernel, EXACT operand
@pytest.mark.parametrize("nested", [False, True])
def test_decode_kernel_exact_operand_is_bit_exact_dequant(pkg, cuda_device, nested):
    """One-hot activations through the DECODE-regime tcgen05 kernel (token tiles <= 32 rows) in EXACT mode return the
    dequantised weight columns themselves: the A operand must be bit-identical to bnb's dequantize_4bit; K = 320 (5 k-blocks:
    one full 4-block stage + a padded one), 3 row tiles."""
    E, H, I, k = 2, 320, 192, 1
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 157, E, H, I)
    T = 24                                        # 12 tokens per expert -> 32-row tiles -> dec_kernel
    d13 = np.stack([orc.dequant_state(s) for s in st13])      # [E, 2I, H]
    for c0 in range(0, H, T):
        cols = np.arange(c0, min(c0 + T, H))
        x = np.zeros((T, H), np.float16)
        x[np.arange(cols.size), cols] = 1
        ids = (np.arange(T) % E).astype(np.int32).reshape(T, 1)
        b = pkg.ops.bucket(torch.from_numpy(ids).to(cuda_device), E)
        assert int(b["meta"][3].item()) <= 32
        xs = pkg.ops.permute_rows(torch.from_numpy(x).to(cuda_device), b, k)
        raw = pkg.ops.expert_w13(xs, w13, b, T, k, act=pkg.ACT_CHUNK_SILU, regime=pkg.REGIME_GEMM, nested=nested).cpu().numpy()
        assert int(b["meta"][4].item()) == 0
        sp = b["sorted_pairs"].cpu().numpy()[:T]
        for s, pair in enumerate(sp):
            if pair < cols.size:
                assert np.array_equal(raw[s].view(np.uint16), d13[ids[pair, 0]][:, cols[pair]].view(np.uint16))


@pytest.mark.parametrize("T,E,H,I,k", [(9, 8, 128, 128, 2), (40, 8, 256, 192, 4), (100, 4, 192, 128, 2), (64, 32, 320, 128, 4)])
def test_decode_kernel_exact_vs_fast_vs_oracle(pkg, cuda_device, T, E, H, I, k):
    """REGIME_GEMM (EXACT operand, what vLLM `operand_mode="exact"` selects) and REGIME_GEMM_FAST through the decode-regime
    kernel, flat and nested: all inside the north-star tolerance, and flat == nested bit for bit in each mode."""
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 163, E, H, I)
    rng = np.random.default_rng(T * 7)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    xd, lgd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(lg).to(cuda_device)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k)
    for regime in (pkg.REGIME_GEMM, pkg.REGIME_GEMM_FAST):
        o_f = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=regime, nested=False)
        o_n = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=regime, nested=True)
        assert torch.equal(o_f, o_n)
        assert_moe_close(o_f.cpu().numpy(), ref)
    pkg.ops.check_status()


@pytest.mark.parametrize("T,E,H,I,k,skew", [(256, 32, 320, 128, 4, False), (200, 8, 256, 192, 4, False), (160, 4, 192, 128, 2, True),
                                            (192, 32, 2880, 128, 4, True)])
def test_decode_kernel_64_row_token_tiles(pkg, cuda_device, T, E, H, I, k, skew):
    """Round 2: with nested absmax (and the FAST operand) the decode-regime kernel takes token tiles of up to 64 rows (64-column
    accumulators, 64-row activation boxes, the epilogue reads the accumulator 32 columns at a time), so an expert with 33..64 routed
    rows is decoded once instead of twice.  Shapes with tiles of 33..48 and 49..64 rows, experts that need several 64-row tiles,
    K = 2880 (45 k-blocks, 3-stage activation ring), with and without the fused up-projection bias; against the oracle and
    against the GEMV regime.  AUTO resolves to this path for these shapes."""
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 640 + T, E, H, I, 0.02 if H > 1000 else 0.05)
    rng = np.random.default_rng(T + 3 * E)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    if skew:
        lg[np.arange(T), np.arange(T) % 2] += 3.0        # two hot experts: > 64 rows each, the rest ragged
    xd, lgd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(lg).to(cuda_device)
    assert pkg.ops.auto_regime(T, k, E, True) == pkg.REGIME_GEMM_FAST
    ids, _ = orc.route_topk(lg, k)
    counts = np.bincount(ids.reshape(-1), minlength=E)
    assert counts.max() > 32, "the case must need tiles wider than 32 rows"
    ws = pkg.ops.new_workspace(cuda_device, T, E, H, I, k)
    out = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, nested=True, workspace=ws)           # AUTO
    assert pkg.ops.workspace_status(ws) == 0
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k)
    assert_moe_close(out.cpu().numpy(), ref)
    out_f = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, nested=True, regime=pkg.REGIME_GEMM_FAST, workspace=ws)
    assert torch.equal(out, out_f)
    out_v = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, nested=True, regime=pkg.REGIME_GEMV)
    assert_moe_close(out.cpu().numpy(), out_v.cpu().numpy())
    # flat absmax keeps 32-row tiles (the 64-row plan does not fit next to 46 KB of absmax stages): same operand, same arithmetic
    # per (row, token) -> bit-identical
    out_flat = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, nested=False, regime=pkg.REGIME_GEMM_FAST)
    assert torch.equal(out, out_flat)
    # fused up-projection bias in the 64-column epilogue
    b13 = (rng.standard_normal((E, 2 * I)) * 0.5).astype(np.float16)
    b2 = (rng.standard_normal((E, H)) * 0.5).astype(np.float16)
    out_b = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, nested=True, w13_bias=torch.from_numpy(b13).to(cuda_device),
                                w2_bias=torch.from_numpy(b2).to(cuda_device), workspace=ws)
    assert pkg.ops.workspace_status(ws) == 0
    assert_moe_close(out_b.cpu().numpy(), orc.moe_forward_from_states(x, lg, st13, st2, k=k, b13=b13, b2=b2))


def test_bf16_activations_are_widened_exactly(pkg, cuda_device):
    """P/bnb_moe_4bit.py:35-37 lists bf16 beside fp16.  nf4moe_cast_f16_bf16: bf16 -> fp16 exact inside the fp16 range (saturating
    outside, NaN kept), fp16 -> bf16 round-to-nearest-even -- against torch's own conversions bit for bit; a bf16 layer call equals
    the fp16 call on the widened input, rounded once, and meets the MoE tolerance against the oracle on the widened input."""
    g = torch.Generator(device=cuda_device).manual_seed(5)
    v = (torch.randn(100003, generator=g, device=cuda_device) * 40).to(torch.bfloat16)
    v[:6] = torch.tensor([0.0, -0.0, 70000.0, -1e30, float("inf"), 3e-6], device=cuda_device).to(torch.bfloat16)
    h = pkg.ops.cast_f16_bf16(v)
    ref = v.float().clamp(-65504.0, 65504.0).to(torch.float16)
    assert h.dtype == torch.float16 and torch.equal(h.view(torch.int16), ref.view(torch.int16))
    nan = pkg.ops.cast_f16_bf16(torch.full((8,), float("nan"), device=cuda_device, dtype=torch.bfloat16))
    assert bool(torch.isnan(nan).all())
    f = (torch.randn(4099, generator=g, device=cuda_device) * 100).half()
    b = pkg.ops.cast_f16_bf16(f)
    assert b.dtype == torch.bfloat16 and torch.equal(b.view(torch.int16), f.to(torch.bfloat16).view(torch.int16))
    # the layer
    T, E, H, I, k = 40, 8, 256, 192, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 77, E, H, I)
    rng = np.random.default_rng(17)
    xb = torch.from_numpy(rng.standard_normal((T, H)).astype(np.float32)).to(cuda_device).to(torch.bfloat16)
    lg = torch.from_numpy(rng.standard_normal((T, E)).astype(np.float16)).to(cuda_device)
    out_b = pkg.ops.moe_forward(xb, w13, w2, k, router_logits=lg)
    x16 = xb.to(torch.float16)                         # exact: |x| < 65504
    out_h = pkg.ops.moe_forward(x16, w13, w2, k, router_logits=lg)
    assert out_b.dtype == torch.bfloat16 and torch.equal(out_b.view(torch.int16), out_h.to(torch.bfloat16).view(torch.int16))
    ref = orc.moe_forward_from_states(x16.cpu().numpy(), lg.cpu().numpy(), st13, st2, k=k)
    got = out_b.float().cpu().numpy()
    assert np.abs(got - ref.astype(np.float32)).max() <= 1e-2 + 2.0 ** -8 * np.abs(ref.astype(np.float32)).max()   # + one bf16 rounding


@pytest.mark.parametrize("T,E,k", [(1, 32, 4), (2, 32, 4), (5, 32, 4), (8, 32, 4), (8, 128, 4), (3, 8, 2), (8, 255, 4), (6, 40, 5)])
def test_router_and_bucket_in_one_launch_for_tiny_batches(pkg, cuda_device, T, E, k):
    """Batch-1..8 decode: nf4moe_moe_forward routes and buckets in ONE single-CTA launch (route_bucket_small_kernel).  The routing it
    returns is bit-identical to the stand-alone router (ties included), the layer output is bit-identical to the same layer driven
    with that routing through the ordinary bucket kernel, and it meets the tolerance against the oracle."""
    H, I = 128, 64
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 900 + E + T, E, H, I)
    rng = np.random.default_rng(T * 1000 + E)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    lg[0, : min(E, 6)] = lg[0, 0]                       # ties: the lowest expert index wins
    xd, lgd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(lg).to(cuda_device)
    out, ids, wts = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, return_routing=True)
    launches = pkg.ops.last_launch_count()
    rid, rw = pkg.ops.route_topk(lgd, k)
    assert torch.equal(ids, rid) and torch.equal(wts, rw)
    oid, _ = orc.route_topk(lg, k)
    assert np.array_equal(ids.cpu().numpy(), oid)
    out2 = pkg.ops.moe_forward(xd, w13, w2, k, topk_ids=rid, topk_weights=rw)      # ordinary bucket kernel
    assert torch.equal(out, out2)
    assert launches == pkg.ops.last_launch_count(), "one launch for router + bucket = as many launches as the pre-routed call"
    assert_moe_close(out.cpu().numpy(), orc.moe_forward_from_states(x, lg, st13, st2, k=k))


def test_per_expert_code2_tables(pkg, cuda_device):
    """code2_stride = 256 (one double-quant table per expert): the decode kernel then stages the expert's table per item
    instead of using the per-lane copy in the decode table's holes; results must not change."""
    E, H, I, k, T = 4, 256, 128, 2, 20
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 171, E, H, I)
    x = torch.randn(T, H, generator=torch.Generator().manual_seed(1)).half().to(cuda_device)
    lg = torch.randn(T, E, generator=torch.Generator().manual_seed(2)).half().to(cuda_device)
    base = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, nested=True, regime=pkg.REGIME_GEMM_FAST)
    for w in (w13, w2):
        w.code2 = w.code2.reshape(1, 256).repeat(E, 1).contiguous()
    assert w13.c_struct(True).code2_stride == 256
    per = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, nested=True, regime=pkg.REGIME_GEMM_FAST)
    assert torch.equal(base, per)


# ------------------------------------------------------------------------------------------------- failure is loud
def test_shape_the_pipeline_cannot_place_is_refused(pkg, cuda_device):
    """K = 4096 with flat fp32 absmax does not fit the decode kernel's shared-memory plan (two 32 KB absmax stages + rings):
    the C ABI returns an error instead of launching a kernel that would do nothing (round 1 returned 0 and garbage)."""
    E, H, I, k, T = 2, 4096, 64, 1, 40
    packed = torch.zeros(E, 2 * I * H // 2, 1, dtype=torch.uint8, device=cuda_device)
    w13 = pkg.NF4ExpertWeights(packed, 2 * I, H, absmax=torch.ones(E, 2 * I * H // 64, device=cuda_device))
    ids = torch.zeros(T, 1, dtype=torch.int32, device=cuda_device)
    b = pkg.ops.bucket(ids, E)
    xs = pkg.ops.permute_rows(torch.zeros(T, H, dtype=torch.float16, device=cuda_device), b, k)
    try:
        pkg.ops.expert_w13(xs, w13, b, T, k, regime=pkg.REGIME_GEMM_FAST, nested=False)
    except pkg.Nf4MoeError as exc:
        assert "shared-memory plan" in str(exc)
    else:   # if a future plan does fit, it must really have run: no status bit, finite output
        assert int(b["meta"][4].item()) == 0


def test_merged_launch_refuses_a_tile_list_built_for_larger_tiles(pkg, cuda_device):
    """nf4moe_gemm_w13_w2 serves token tiles of at most 32 rows.  Handed the list nf4moe_bucket builds by default at a prefill-sized
    batch (160-row tiles; meta[5] records the size), the kernel used to clamp its MMA N and drop the rest of every tile silently.
    Now the whole launch is refused: NF4MOE_STATUS_DEC_NOFIT in meta[4], h and y untouched.  The same batch bucketed with 32-row
    tiles runs and raises nothing."""
    E, H, I, k, T = 4, 256, 128, 2, 320                      # 640 pairs on 4 experts: 160 rows per expert on average
    g = torch.Generator(device=cuda_device).manual_seed(5)
    code2 = torch.linspace(-1, 1, 256)
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=cuda_device) * 0.05).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=cuda_device) * 0.05).half(), code2)
    x = torch.randn(T, H, generator=g, device=cuda_device).half()
    ids = torch.stack([torch.arange(T, device=cuda_device) % E, (torch.arange(T, device=cuda_device) + 1) % E], 1).to(torch.int32)
    b = pkg.ops.bucket(ids, E)                               # tile_rows = 0: what the prefill grouped GEMM wants here
    assert int(b["meta"][5].item()) > 32
    xs = pkg.ops.permute_rows(x, b, k)
    h = torch.full((T * k, I), 3.0, dtype=torch.float16, device=cuda_device)
    y = torch.full((T * k, H), 3.0, dtype=torch.float16, device=cuda_device)
    pkg.ops.expert_w13_w2(xs, w13, w2, b, T, k, nested=True, h=h, y=y)
    torch.cuda.synchronize(cuda_device)
    assert int(b["meta"][4].item()) & pkg._lib.STATUS_DEC_NOFIT
    assert bool((h == 3.0).all()) and bool((y == 3.0).all())
    b32 = pkg.ops.bucket(ids, E, tile_rows=32)
    assert int(b32["meta"][5].item()) == 32
    h2, y2 = pkg.ops.expert_w13_w2(pkg.ops.permute_rows(x, b32, k), w13, w2, b32, T, k, nested=True)
    torch.cuda.synchronize(cuda_device)
    assert int(b32["meta"][4].item()) == 0
    ref = pkg.ops.moe_forward(x, w13, w2, k, topk_ids=ids, topk_weights=torch.ones(T, k, device=cuda_device), nested=True,
                              regime=pkg.REGIME_GEMM_FAST)
    got = y2.view(T, k, H).float().sum(1)
    assert torch.allclose(got, ref.float(), atol=2e-2, rtol=1e-2)


def test_split_decode_launches_for_shared_gpus_are_bit_identical(pkg, cuda_device):
    """ops.set_merged_decode(False) (nf4moe_set_merged_decode): the decode regime runs w13 and w2 as two launches, which need no
    co-residency of the whole grid -- the setting for MPS / partitioned GPUs.  Same kernel template, so the layer output is
    bit-identical to the merged launch, with exactly one launch more; the switch returns the previous setting and is restored."""
    E, H, I, k, T = 8, 256, 128, 4, 48
    g = torch.Generator(device=cuda_device).manual_seed(9)
    code2 = torch.linspace(-1, 1, 256)
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=cuda_device) * 0.05).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=cuda_device) * 0.05).half(), code2)
    x = torch.randn(T, H, generator=g, device=cuda_device).half()
    lg = torch.randn(T, E, generator=g, device=cuda_device).half()
    merged = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, nested=True, regime=pkg.REGIME_GEMM_FAST)
    n_merged = pkg.ops.last_launch_count()
    assert pkg.ops.set_merged_decode(False) is True
    try:
        split = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, nested=True, regime=pkg.REGIME_GEMM_FAST)
        n_split = pkg.ops.last_launch_count()
    finally:
        assert pkg.ops.set_merged_decode(True) is False
    assert torch.equal(merged, split)
    assert n_split == n_merged + 1
    assert pkg.ops.last_launch_count() == n_split and pkg.ops.set_merged_decode(True) is True


# ------------------------------------------------------------------------------------------------- real vLLM layer
def test_real_vllm_fused_moe_forward(pkg, cuda_device):
    """What P/patch_6_add_bnb_moe.py:75-107 and P/patch_9_fusedmoe_unsloth.py:19-56 were trying to reach: a REAL
    vllm FusedMoE built with quant_config = the plugin's config, packed weights + `bnb_quant_state` bound the way vLLM's
    bitsandbytes loader leaves them, then the layer's own quant_method.apply() (the call the MoE runner makes) and, when the
    runner's custom-op path is usable on this box, the layer's forward."""
    pytest.importorskip("vllm")
    build_real_fused_moe = _load("_loader_plugin_helpers", os.path.join(HERE, "test_loader_plugin.py")).build_real_fused_moe
    plug = importlib.import_module("gpt-oss-vllm_b200.vllm_plugin")
    try:
        cls = plug.register()
        E, H, I, k, T = 8, 256, 128, 4, 48
        layer, vc = build_real_fused_moe(cls, device="cuda", E=E, H=H, I=I, top_k=k)
    except Exception as exc:
        pytest.skip(f"vLLM cannot build a FusedMoE on this box: {type(exc).__name__}: {exc}")
    from vllm.model_executor.layers.fused_moe.fused_moe_method_base import FusedMoEMethodBase
    assert isinstance(layer.quant_method, FusedMoEMethodBase)
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 181, E, H, I)
    layer.to(cuda_device)
    layer.w13_weight.data.copy_(w13.packed)
    layer.w2_weight.data.copy_(w2.packed)

    class QS:
        def __init__(self, absmax):
            self.absmax, self.blocksize, self.quant_type, self.nested = absmax.reshape(-1), 64, "nf4", False
    layer.w13_weight.bnb_quant_state, layer.w2_weight.bnb_quant_state = QS(w13.absmax), QS(w2.absmax)
    rng = np.random.default_rng(5)
    b13 = (rng.standard_normal((E, 2 * I)) * 0.3).astype(np.float16)
    b2 = (rng.standard_normal((E, H)) * 0.3).astype(np.float16)
    layer.w13_bias.data.copy_(torch.from_numpy(b13))
    layer.w2_bias.data.copy_(torch.from_numpy(b2))
    layer.quant_method.process_weights_after_loading(layer)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    xd, lgd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(lg).to(cuda_device)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k, b13=b13, b2=b2)
    ids, wts = pkg.ops.route_topk(lgd, k)
    out = layer.quant_method.apply(layer, xd, wts, ids, None, None)
    pkg.ops.check_status()
    assert_moe_close(out.cpu().numpy(), ref)
    # the whole layer through vLLM's runner (router + custom op + our apply); a vLLM-internal failure is recorded, not hidden
    try:
        from vllm.forward_context import set_forward_context
        with set_forward_context(None, vc):
            full = layer(xd.clone(), lgd)
    except Exception as exc:
        print(f"vLLM runner path not usable on this box ({type(exc).__name__}: {exc}); quant_method.apply() verified instead")
        return
    assert_moe_close(full.cpu().numpy(), ref)


def test_apply_pass_throughs(pkg, cuda_device):
    """`expert_map` / `global_num_experts` and `apply_router_weight_on_input`, read from the layer as vLLM's bnb method reads
    them (V/model_executor/layers/quantization/bitsandbytes.py:493-505)."""
    E, H, I, T = 8, 128, 128, 12
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 191, E, H, I)
    rng = np.random.default_rng(9)
    x = rng.standard_normal((T, H)).astype(np.float16)
    xd = torch.from_numpy(x).to(cuda_device)
    # (a) expert_map: 16 global experts, this rank owns global 8..15 as local 0..7
    lg16 = rng.standard_normal((T, 16)).astype(np.float16)
    gids, gw = orc.route_topk(lg16, 4)
    method = pkg.B200NF4MoEMethod(top_k=4)
    layer = torch.nn.Module()
    method.bind(layer, w13, w2)
    emap = torch.full((16,), -1, dtype=torch.int32, device=cuda_device)
    emap[8:] = torch.arange(8, dtype=torch.int32, device=cuda_device)
    layer.expert_map, layer.global_num_experts = emap, 16
    out = method.apply(layer, xd, torch.from_numpy(gw).to(cuda_device), torch.from_numpy(gids).to(cuda_device), None, None)
    local = gids >= 8
    lids = np.where(local, gids - 8, 0).astype(np.int32)
    ref = orc.moe_forward_from_states(x, lg16[:, :8], st13, st2, k=4, topk=(lids, np.where(local, gw, 0).astype(np.float32)))
    assert_moe_close(out.cpu().numpy(), ref)
    layer.global_num_experts = 12
    with pytest.raises(ValueError):
        method.apply(layer, xd, torch.from_numpy(gw).to(cuda_device), torch.from_numpy(gids).to(cuda_device), None, None)
    # (b) apply_router_weight_on_input, top_k = 1
    layer2 = torch.nn.Module()
    method.bind(layer2, w13, w2)
    layer2.apply_router_weight_on_input = True
    lg = rng.standard_normal((T, E)).astype(np.float16)
    ids1, w1 = orc.route_topk(lg, 1)
    w1 = (w1 * 0.5).astype(np.float32)          # softmax over one logit is 1.0; scale it so the test can tell where it is applied
    out1 = method.apply(layer2, xd, torch.from_numpy(w1).to(cuda_device), torch.from_numpy(ids1).to(cuda_device), None, None)
    x_scaled = (torch.from_numpy(x) * torch.from_numpy(w1).half()).numpy()
    ref1 = orc.moe_forward_from_states(x_scaled, lg, st13, st2, k=1, topk=(ids1, np.ones_like(w1)))
    assert_moe_close(out1.cpu().numpy(), ref1)
    with pytest.raises(ValueError):
        layer2.apply_router_weight_on_input = True
        method.apply(layer2, xd, torch.from_numpy(gw).to(cuda_device), torch.from_numpy(np.minimum(gids, 7)).to(cuda_device), None, None)


def test_workspace_pool_is_per_stream_and_keeps_captured_buffers(pkg, cuda_device):
    pool = pkg.ops.WorkspacePool()
    a = pool.get(cuda_device, 16, 8, 128, 128, 2)
    assert pool.get(cuda_device, 64, 8, 128, 128, 2) is a                  # same capacity bucket
    s = torch.cuda.Stream(device=cuda_device)
    with torch.cuda.stream(s):
        b = pool.get(cuda_device, 16, 8, 128, 128, 2)
    assert b is not a and b.data_ptr() != a.data_ptr()                     # another stream, another buffer
    big = pool.get(cuda_device, 300, 8, 128, 128, 2)
    assert big is not a and any(t is a for t in pool.buffers())            # the superseded buffer stays alive (captured graphs)
    assert pkg.ops.workspace_status(big) == 0 and int(a[:4].view(torch.int32).item()) == 0
