"""GPU parity tests proper: every call goes through the C ABI (ctypes -> libnf4moe.so) and is compared with the CPU
oracle on the same seeded inputs.  Bars (BASELINE.json north_star): NF4 dequant bit-exact, router ids bit-exact,
MoE output max-abs <= 1e-2 and rel <= 2e-3 (fp16 storage, fp32 accumulate)."""
import numpy as np
import pytest
import torch

from oracle import nf4_oracle as orc

pytestmark = pytest.mark.gpu

ATOL, RTOL = 1e-2, 2e-3   # north_star tolerance for MoE outputs


def assert_moe_close(got: np.ndarray, ref: np.ndarray, intermediate: bool = False):
    """north-star bar for MoE outputs.  `intermediate`: the activation output h can reach tens in magnitude, where one fp16
    ulp already exceeds the absolute bar, so only the relative-L2 and the mixed per-element bounds apply to it."""
    got32, ref32 = got.astype(np.float32), ref.astype(np.float32)
    diff = np.abs(got32 - ref32)
    assert np.isfinite(got32).all()
    assert intermediate or diff.max() <= ATOL, f"max-abs {diff.max()}"
    rel = np.linalg.norm(got32 - ref32) / max(np.linalg.norm(ref32), 1e-12)
    assert rel <= RTOL, f"rel {rel}"
    assert np.all(diff <= ATOL + RTOL * np.abs(ref32))


def make_states(rng, E, rows, cols, std=0.02):
    return [orc.quantize_nf4((rng.standard_normal((rows, cols)) * std).astype(np.float16)) for _ in range(E)]


def make_layer(pkg, dev, seed, E, H, I, std=0.05):
    rng = np.random.default_rng(seed)
    st13, st2 = make_states(rng, E, 2 * I, H, std), make_states(rng, E, H, I, std)
    return st13, st2, pkg.NF4ExpertWeights.from_states(st13, dev), pkg.NF4ExpertWeights.from_states(st2, dev)


# ------------------------------------------------------------------------------------------------- dequant

def expected_tile_rows(T, k, E):
    """gemm_tile_tokens (gemm_tc.cuh): 32-row token tiles while the mean rows per expert stay <= 40 (or <= 128 with <= 8 experts)."""
    return 32 if (T * k <= 40 * E or (E <= 8 and T * k <= 128 * E)) else 160

@pytest.mark.parametrize("nested", [True, False])
def test_dequant_bit_exact_random(pkg, cuda_device, nested):
    rng = np.random.default_rng(11)
    sts = make_states(rng, 3, 96, 320)
    w = pkg.NF4ExpertWeights.from_states(sts, cuda_device)
    got = w.dequantize(nested=nested).cpu().numpy()
    ref = np.stack([orc.dequant_state(s, nested=nested) for s in sts])
    assert np.array_equal(got.view(np.uint16), ref.view(np.uint16))
    # expert sub-range
    got1 = w.dequantize(nested=nested, e0=1, ne=2).cpu().numpy()
    assert np.array_equal(got1.view(np.uint16), ref[1:].view(np.uint16))


def test_dequant_bit_exact_adversarial(pkg, cuda_device):
    """every byte value x absmax extremes (zero, denormal, huge, inf) -- raw packed bytes, no quantiser involved."""
    rng = np.random.default_rng(12)
    rows, cols = 64, 256
    nb = rows * cols // 64
    packed = np.tile(np.arange(256, dtype=np.uint8), rows * cols // 2 // 256)
    rng.shuffle(packed)
    absmax = (rng.random(nb, dtype=np.float32) * 0.1).astype(np.float32)
    absmax[:8] = [0.0, 1e-42, 6.0e-8, 5.9e-8, 65504.0, 70000.0, np.inf, 1.0]
    st = {"packed": packed, "absmax_f32": absmax, "shape": (rows, cols), "blocksize": 64}
    w = pkg.NF4ExpertWeights.from_states([st], cuda_device)
    got = w.dequantize().cpu().numpy()[0]
    with np.errstate(all="ignore"):
        ref = orc.dequant_nf4(packed, absmax, (rows, cols))
    assert np.array_equal(got.view(np.uint16), ref.view(np.uint16))
    # known answer: byte 0x0F -> (-absmax, +absmax): bnb's high-nibble-first order
    st2 = {"packed": np.full(32, 0x0F, np.uint8), "absmax_f32": np.array([0.5], np.float32), "shape": (1, 64)}
    d = pkg.NF4ExpertWeights.from_states([st2], cuda_device).dequantize().cpu().numpy().reshape(-1)
    assert d[0] == np.float16(-0.5) and d[1] == np.float16(0.5)


def test_dequant_full_size_expert_and_denest(pkg, cuda_device):
    """one full gpt-oss w13 expert [5760, 2880]: 259,200 blocks, 1,013 nested blocks with a partial tail."""
    rng = np.random.default_rng(13)
    st = orc.quantize_nf4((rng.standard_normal((5760, 2880)) * 0.02).astype(np.float16))
    assert st["absmax2"].size == 1013
    w = pkg.NF4ExpertWeights.from_states([st], cuda_device)
    flat_given = w.absmax.clone()
    denested = w.denest().cpu().numpy()[0]
    assert np.array_equal(denested.view(np.uint32), st["absmax_f32"].view(np.uint32))
    assert torch.equal(flat_given, w.absmax)
    got = w.dequantize(nested=True).cpu().numpy()[0]
    assert np.array_equal(got.view(np.uint16), orc.dequant_state(st).view(np.uint16))


def test_device_quantizer_matches_oracle(pkg, cuda_device):
    rng = np.random.default_rng(14)
    E, rows, cols = 2, 160, 448           # 1120 blocks per expert -> 5 nested blocks, last partial
    w = (rng.standard_normal((E, rows, cols)) * 0.02).astype(np.float16)
    w[0, 0, :64] = 0                      # all-zero block
    w[1, 3, 70] = 2.5                     # outlier
    code2 = orc.create_dynamic_map()
    q = pkg.NF4ExpertWeights.quantize(torch.from_numpy(w).to(cuda_device), torch.from_numpy(code2))
    for e in range(E):
        st = orc.quantize_nf4(w[e], code2=code2)
        assert np.array_equal(q.packed[e].cpu().numpy().reshape(-1), st["packed"])
        assert np.array_equal(q.absmax2[e].cpu().numpy(), st["absmax2"])
        assert np.float32(q.offset[e].item()) == st["offset"]
        assert np.array_equal(q.absmax_u8[e].cpu().numpy(), st["absmax_u8"])
        assert np.array_equal(q.absmax[e].cpu().numpy(), st["absmax_f32"])
    # round trip through the device dequantiser
    d = q.dequantize(nested=True).cpu().numpy()
    assert np.array_equal(d[0], orc.dequant_state(orc.quantize_nf4(w[0], code2=code2)))


# ------------------------------------------------------------------------------------------------- router / bucketing
@pytest.mark.parametrize("E,k,dtype", [(32, 4, torch.float16), (128, 4, torch.float16), (32, 4, torch.float32),
                                       (64, 2, torch.bfloat16), (256, 8, torch.float16)])
def test_route_topk_bit_exact_ids(pkg, cuda_device, E, k, dtype):
    g = torch.Generator().manual_seed(21)
    lg = (torch.randn(257, E, generator=g) * 2).to(dtype)
    lg[3, 5] = lg[3, 9] = lg[3].max() + 1   # tie for first place
    lg[7, :] = 0.5                           # fully tied row
    ids, w = pkg.ops.route_topk(lg.to(cuda_device), k)
    rid, rw = orc.route_topk(lg.float().numpy(), k)
    assert np.array_equal(ids.cpu().numpy(), rid)
    np.testing.assert_allclose(w.cpu().numpy(), rw, rtol=0, atol=1e-6)


def test_route_topk_golden_fixture(pkg, cuda_device, golden):
    arrays, _ = golden
    lg = arrays["route/logits_f16"]
    ids, w = pkg.ops.route_topk(torch.from_numpy(lg).to(cuda_device), 4)
    ids = ids.cpu().numpy()
    tied = np.array([np.unique(r).size < r.size for r in lg.astype(np.float32)])
    assert np.array_equal(ids[~tied], arrays["route/ids"][~tied])
    np.testing.assert_allclose(w.cpu().numpy(), arrays["route/weights"], rtol=0, atol=1e-6)


@pytest.mark.parametrize("T,k,E", [(1, 4, 32), (7, 4, 32), (64, 4, 32), (300, 4, 128), (8192, 4, 32), (5, 1, 32),
                                   (513, 4, 32), (2001, 4, 128), (4099, 2, 8)])   # >= 2048 pairs: the 8-CTA cluster kernel
def test_bucket_stable_sort(pkg, cuda_device, T, k, E):
    rng = np.random.default_rng(22)
    ids = np.stack([rng.choice(E, size=k, replace=False) for _ in range(T)]).astype(np.int32)
    if T >= 7:
        ids[2, 1] = -1            # dropped pair (expert-parallel mask)
        ids[5, 0] = E + 3
    b = pkg.ops.bucket(torch.from_numpy(ids).to(cuda_device), E)
    flat = ids.reshape(-1)
    keep = (flat >= 0) & (flat < E)
    order = np.argsort(np.where(keep, flat, E), kind="stable")[:keep.sum()]
    counts = np.bincount(flat[keep], minlength=E)
    off = np.concatenate([[0], np.cumsum(counts)])
    assert np.array_equal(b["expert_offsets"].cpu().numpy(), off)
    assert np.array_equal(b["sorted_pairs"].cpu().numpy()[:keep.sum()], order)
    meta = b["meta"].cpu().numpy()
    active = np.nonzero(counts)[0]
    assert meta[0] == active.size and meta[1] == keep.sum() and meta[3] == counts.max()
    assert np.array_equal(b["active_experts"].cpu().numpy()[:active.size], active)
    tile_rows = expected_tile_rows(T, k, E)               # what the grouped GEMM expects for this (T, k, E)
    tiles = b["tile_list"].cpu().numpy()[:4 * meta[2]].reshape(-1, 4)
    exp_tiles = []
    for e in active:
        c = int(counts[e])
        ntl = -(-c // tile_rows)
        sz = (-(-c // ntl) + 15) // 16 * 16
        exp_tiles += [[e, off[e] + r, min(sz, c - r), 0] for r in range(0, c, sz)]
    assert tiles.tolist() == exp_tiles


def test_permute_and_combine(pkg, cuda_device):
    rng = np.random.default_rng(23)
    T, k, E, H = 37, 4, 32, 128
    ids = np.stack([rng.choice(E, size=k, replace=False) for _ in range(T)]).astype(np.int32)
    x = rng.standard_normal((T, H)).astype(np.float16)
    b = pkg.ops.bucket(torch.from_numpy(ids).to(cuda_device), E)
    xs = pkg.ops.permute_rows(torch.from_numpy(x).to(cuda_device), b, k).cpu().numpy()
    sp = b["sorted_pairs"].cpu().numpy()[:T * k]
    assert np.array_equal(xs[:T * k], x[sp // k])
    y = rng.standard_normal((T * k, H)).astype(np.float16)
    w = rng.random((T, k)).astype(np.float32)
    out = pkg.ops.combine(torch.from_numpy(y).to(cuda_device), torch.from_numpy(w).to(cuda_device), None, E).cpu().numpy()
    ref = np.zeros((T, H), np.float32)
    for s in range(k):
        ref += w[:, s:s + 1] * y.reshape(T, k, H)[:, s].astype(np.float32)
    assert np.array_equal(out, ref.astype(np.float16))


# ------------------------------------------------------------------------------------------------- expert products
@pytest.mark.parametrize("T", [1, 3, 8, 9, 16, 17, 40])
@pytest.mark.parametrize("nested", [False, True])
def test_gemv_stages_vs_oracle(pkg, cuda_device, T, nested):
    """w13+SwiGLU and w2 separately against the oracle's per-expert fp16 matmul (P/unsloth_to_vllm_converter.py:309-318)."""
    E, H, I, k = 8, 256, 192, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 31, E, H, I)
    rng = np.random.default_rng(100 + T)
    x = rng.standard_normal((T, H)).astype(np.float16)
    ids = np.stack([rng.choice(E, size=k, replace=False) for _ in range(T)]).astype(np.int32)
    b = pkg.ops.bucket(torch.from_numpy(ids).to(cuda_device), E)
    xd = torch.from_numpy(x).to(cuda_device)
    h = pkg.ops.expert_w13(xd, w13, b, T, k, nested=nested)
    y = pkg.ops.expert_w2(h, w2, b, T, k, nested=nested).cpu().numpy()
    h = h.cpu().numpy()
    sp = b["sorted_pairs"].cpu().numpy()[:T * k]
    off = b["expert_offsets"].cpu().numpy()
    d13 = np.stack([orc.dequant_state(s) for s in st13])
    d2 = np.stack([orc.dequant_state(s) for s in st2])
    for e in range(E):
        rows = sp[off[e]:off[e + 1]]
        if rows.size == 0:
            continue
        xe = x[rows // k]
        h_ref = orc.swiglu_oai(orc._mm_f16_f32acc(xe, d13[e]).astype(np.float16).astype(np.float32)).astype(np.float16)
        assert_moe_close(h[off[e]:off[e + 1]], h_ref, intermediate=True)
        y_ref = orc._mm_f16_f32acc(h[off[e]:off[e + 1]], d2[e]).astype(np.float16)
        assert_moe_close(y[rows], y_ref)


@pytest.mark.parametrize("T,E,H,I,k", [(1, 32, 128, 128, 4), (2, 32, 192, 64, 4), (5, 8, 64, 320, 2), (16, 32, 256, 256, 4),
                                       (33, 32, 128, 192, 4), (64, 32, 320, 128, 4), (70, 128, 128, 128, 4),
                                       (200, 8, 128, 128, 4)])
def test_moe_forward_vs_oracle(pkg, cuda_device, T, E, H, I, k):
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 41, E, H, I)
    rng = np.random.default_rng(T)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    out, ids, wts = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k,
                                        router_logits=torch.from_numpy(lg).to(cuda_device), return_routing=True,
                                        regime=pkg.REGIME_GEMV)
    rid, rw = orc.route_topk(lg, k)
    assert np.array_equal(ids.cpu().numpy(), rid)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k)
    assert_moe_close(out.cpu().numpy(), ref)
    # precomputed routing (vLLM 0.22 `apply` shape) gives the same answer bit for bit
    out2 = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k, topk_ids=ids, topk_weights=wts,
                               regime=pkg.REGIME_GEMV)
    assert torch.equal(out, out2)
    # nested-absmax kernels read the double-quant state directly: same numbers
    out3 = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k, topk_ids=ids, topk_weights=wts,
                               nested=True, regime=pkg.REGIME_GEMV)
    assert torch.equal(out, out3)


def test_moe_forward_reference_activation_mode(pkg, cuda_device):
    """chunk + silu, the reference wrapper's own activation (P/unsloth_to_vllm_converter.py:312-315)."""
    T, E, H, I, k = 9, 8, 128, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 43, E, H, I)
    rng = np.random.default_rng(7)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    out = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k,
                              router_logits=torch.from_numpy(lg).to(cuda_device), act=pkg.ACT_CHUNK_SILU,
                              regime=pkg.REGIME_GEMV)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k, act="chunk_silu")
    assert_moe_close(out.cpu().numpy(), ref)


def test_moe_forward_swiglu_saturation_and_zero_blocks(pkg, cuda_device):
    """adversarial: activations large enough to hit the +-7 clamps, an all-zero weight block, skewed routing."""
    T, E, H, I, k = 12, 32, 128, 128, 4
    rng = np.random.default_rng(44)
    w13f = (rng.standard_normal((E, 2 * I, H)) * 0.3).astype(np.float16)
    w13f[:, 0, :64] = 0
    w2f = (rng.standard_normal((E, H, I)) * 0.05).astype(np.float16)
    st13 = [orc.quantize_nf4(w) for w in w13f]
    st2 = [orc.quantize_nf4(w) for w in w2f]
    w13, w2 = pkg.NF4ExpertWeights.from_states(st13, cuda_device), pkg.NF4ExpertWeights.from_states(st2, cuda_device)
    x = (rng.standard_normal((T, H)) * 2).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    lg[:, :4] += 6                      # everyone picks experts 0..3
    out = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k,
                              router_logits=torch.from_numpy(lg).to(cuda_device), regime=pkg.REGIME_GEMV)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k)
    got = out.cpu().numpy().astype(np.float32)
    # outputs here are O(10): scale the absolute bar with the magnitude, keep the relative one
    diff = np.abs(got - ref.astype(np.float32))
    assert np.all(diff <= ATOL * max(1.0, np.abs(ref).max()) + RTOL * np.abs(ref.astype(np.float32)))


def test_moe_forward_masked_experts(pkg, cuda_device):
    """ids outside [0,E) are dropped end to end (what an expert-parallel rank does with non-local experts)."""
    T, E, H, I, k = 10, 8, 128, 128, 4
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 45, E, H, I)
    rng = np.random.default_rng(8)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    ids, w = orc.route_topk(lg, k)
    local = ids < 4
    ids_m = np.where(local, ids, -1).astype(np.int32)
    out = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k,
                              topk_ids=torch.from_numpy(ids_m).to(cuda_device),
                              topk_weights=torch.from_numpy(w).to(cuda_device), regime=pkg.REGIME_GEMV)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k, topk=(ids, np.where(local, w, 0).astype(np.float32)))
    assert_moe_close(out.cpu().numpy(), ref)


def test_config1_full_size_layer_16_tokens(pkg, cuda_device):
    """BASELINE.json configs[0]: one gpt-oss-20b MoE layer (H=I=2880, 32 experts, top-4), 16 tokens, weights made by
    the DEVICE quantiser (so 32 full-size experts cost seconds, not minutes) and checked against the oracle's
    dequant-then-matmul on the experts actually routed to."""
    E, H, I, k, T = 32, 2880, 2880, 4, 16
    g = torch.Generator(device=cuda_device).manual_seed(20250805)
    code2 = torch.from_numpy(orc.create_dynamic_map())
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=cuda_device) * 0.02).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=cuda_device) * 0.02).half(), code2)
    x = torch.randn(T, H, generator=g, device=cuda_device).half()
    lg = torch.randn(T, E, generator=g, device=cuda_device).half()
    out, ids, wts = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, return_routing=True)
    out_n = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, nested=True)
    assert torch.equal(out, out_n)
    rid, rw = orc.route_topk(lg.cpu().numpy(), k)
    assert np.array_equal(ids.cpu().numpy(), rid)
    used = np.unique(rid)
    d13 = np.zeros((E, 2 * I, H), np.float16)
    d2 = np.zeros((E, H, I), np.float16)
    for e in used:   # oracle dequant from the raw quant state the kernels read
        for dst, w in ((d13, w13), (d2, w2)):
            st = {"packed": w.packed[e].cpu().numpy().reshape(-1), "absmax_u8": w.absmax_u8[e].cpu().numpy(),
                  "absmax2": w.absmax2[e].cpu().numpy(), "code2": w.code2.cpu().numpy(), "offset": w.offset[e].item(),
                  "shape": (w.rows, w.cols), "blocksize": 64}
            dst[e] = orc.dequant_state(st)
        # the device dequantiser agrees bit for bit on these full-size experts too
        assert np.array_equal(w13.dequantize(e0=int(e), ne=1).cpu().numpy()[0].view(np.uint16), d13[e].view(np.uint16))
    ref = orc.moe_forward_ref(x.cpu().numpy(), lg.cpu().numpy(), d13, d2, k=k)
    assert_moe_close(out.cpu().numpy(), ref)


def test_size_independent_properties_full_size(pkg, cuda_device):
    """At BASELINE's decode sizes (T=64, full 20b layer) the oracle is too slow to run per test; check properties:
    (a) token-permutation equivariance, (b) each token's row equals the same token run alone (T=1),
    (c) scaling the routing weights scales the output (combine linearity), (d) determinism."""
    E, H, I, k, T = 32, 2880, 2880, 4, 64
    g = torch.Generator(device=cuda_device).manual_seed(99)
    code2 = torch.from_numpy(orc.create_dynamic_map())
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=cuda_device) * 0.02).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=cuda_device) * 0.02).half(), code2)
    x = torch.randn(T, H, generator=g, device=cuda_device).half()
    lg = torch.randn(T, E, generator=g, device=cuda_device).half()
    out, ids, wts = pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg, return_routing=True)
    assert torch.equal(out, pkg.ops.moe_forward(x, w13, w2, k, router_logits=lg))                     # (d)
    perm = torch.randperm(T, generator=torch.Generator().manual_seed(1)).to(cuda_device)
    out_p = pkg.ops.moe_forward(x[perm].contiguous(), w13, w2, k, router_logits=lg[perm].contiguous())
    assert_moe_close(out_p.cpu().numpy(), out[perm].cpu().numpy())                                    # (a)
    for t in (0, 17, 63):                                                                             # (b)
        o1 = pkg.ops.moe_forward(x[t:t + 1].contiguous(), w13, w2, k, router_logits=lg[t:t + 1].contiguous())
        assert_moe_close(o1.cpu().numpy(), out[t:t + 1].cpu().numpy())
    out_h = pkg.ops.moe_forward(x, w13, w2, k, topk_ids=ids, topk_weights=wts * 0.5)                   # (c)
    assert_moe_close(out_h.cpu().numpy() * 2, out.cpu().numpy())


# ------------------------------------------------------------------------------------------------- reference interface
def test_method_entry_points_agree(pkg, cuda_device):
    """create_weights registers vLLM's parameter set; apply / apply_monolithic / legacy / wrapper-shaped forward agree."""
    E, H, I, k, T = 32, 128, 128, 4, 10
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 51, E, H, I)
    method = pkg.B200NF4MoEConfig().get_quant_method(None, "model.layers.0.mlp.experts")
    layer = torch.nn.Module()
    method.create_weights(layer, E, H, I, torch.float16)
    assert tuple(layer.w13_weight.shape) == (E, H * 2 * I // 2, 1) and layer.w13_weight.dtype == torch.uint8
    assert tuple(layer.w2_weight.shape) == (E, H * I // 2, 1)
    assert layer.w13_weight.experts_shape == (E, 2 * I, H) and layer.w2_weight.use_bitsandbytes_4bit
    layer.to(cuda_device)
    layer.w13_weight.data.copy_(w13.packed)
    layer.w2_weight.data.copy_(w2.packed)

    class QS:  # the fields vLLM's fused bnb_quant_state carries (V/.../bitsandbytes_loader.py:673-692)
        def __init__(self, absmax):
            self.absmax, self.blocksize, self.quant_type, self.nested = absmax.reshape(-1), 64, "nf4", False
    layer.w13_weight.bnb_quant_state = QS(w13.absmax)
    layer.w2_weight.bnb_quant_state = QS(w2.absmax)
    layer.top_k = k
    rng = np.random.default_rng(3)
    x = torch.from_numpy(rng.standard_normal((T, H)).astype(np.float16)).to(cuda_device)
    lg = torch.from_numpy(rng.standard_normal((T, E)).astype(np.float16)).to(cuda_device)
    a = method.apply_monolithic(layer, x, lg)
    b = method.apply_legacy(layer, x, lg, k, renormalize=True)
    ids, wts = pkg.ops.route_topk(lg, k)
    c = method.apply(layer, x, wts, ids, None, None)
    wrapper = pkg.NF4MoELayer(w13, w2, top_k=k)
    d = wrapper(x.reshape(2, 5, H), lg.reshape(2, 5, E), top_k=k).reshape(T, H)
    assert torch.equal(a, b) and torch.equal(a, c) and torch.equal(a, d)
    ref = orc.moe_forward_from_states(x.cpu().numpy(), lg.cpu().numpy(), st13, st2, k=k)
    assert_moe_close(a.cpu().numpy(), ref)


def test_wrapper_golden_shapes(pkg, cuda_device):
    """the reference's own self-test shapes: randn(1,10,2880) / randn(1,10,32) -> same shape out
    (P/unsloth_to_vllm_converter.py:356-361), here with numerical parity on top."""
    E, H, I, k = 32, 2880, 2880, 4
    g = torch.Generator(device=cuda_device).manual_seed(5)
    code2 = torch.from_numpy(orc.create_dynamic_map())
    w13 = pkg.NF4ExpertWeights.quantize((torch.randn(E, 2 * I, H, generator=g, device=cuda_device) * 0.02).half(), code2)
    w2 = pkg.NF4ExpertWeights.quantize((torch.randn(E, H, I, generator=g, device=cuda_device) * 0.02).half(), code2)
    layer = pkg.NF4MoELayer(w13, w2, top_k=k)
    hs = torch.randn(1, 10, H, generator=g, device=cuda_device).half()
    lg = torch.randn(1, 10, E, generator=g, device=cuda_device).half()
    out = layer(hs, lg, top_k=4)
    assert out.shape == hs.shape and out.dtype == torch.float16 and torch.isfinite(out).all()


def test_errors_are_loud(pkg, cuda_device):
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 61, 8, 128, 128)
    x = torch.zeros(4, 128, dtype=torch.float16)
    with pytest.raises(RuntimeError):          # CPU tensor: no fallback
        pkg.ops.moe_forward(x, w13, w2, 4, router_logits=torch.zeros(4, 8, dtype=torch.float16))
    with pytest.raises(pkg.Nf4MoeError):       # wrong shapes
        pkg.ops.moe_forward(torch.zeros(4, 64, dtype=torch.float16, device=cuda_device), w13, w2, 4,
                            router_logits=torch.zeros(4, 8, dtype=torch.float16, device=cuda_device))
    with pytest.raises(pkg.Nf4MoeError):       # more experts than the router kernel supports
        pkg.ops.route_topk(torch.zeros(4, 300, dtype=torch.float16, device=cuda_device), 4)


# ------------------------------------------------------------------------------------------------- prefill regime (tcgen05)
def _gemm_stage_check(pkg, dev, T, E, H, I, k, nested, seed, regime=None, std=0.05):
    regime = pkg.REGIME_GEMM if regime is None else regime
    st13, st2, w13, w2 = make_layer(pkg, dev, seed, E, H, I, std)
    rng = np.random.default_rng(200 + T)
    x = rng.standard_normal((T, H)).astype(np.float16)
    ids = np.stack([rng.choice(E, size=k, replace=False) for _ in range(T)]).astype(np.int32)
    b = pkg.ops.bucket(torch.from_numpy(ids).to(dev), E)
    xd = torch.from_numpy(x).to(dev)
    xs = pkg.ops.permute_rows(xd, b, k)
    h = pkg.ops.expert_w13(xs, w13, b, T, k, nested=nested, regime=regime)
    y = pkg.ops.expert_w2(h, w2, b, T, k, nested=nested, regime=regime).cpu().numpy()
    h = h.cpu().numpy()
    sp = b["sorted_pairs"].cpu().numpy()[:T * k]
    off = b["expert_offsets"].cpu().numpy()
    d13 = np.stack([orc.dequant_state(s) for s in st13])
    d2 = np.stack([orc.dequant_state(s) for s in st2])
    for e in range(E):
        rows = sp[off[e]:off[e + 1]]
        if rows.size == 0:
            continue
        xe = x[rows // k]
        h_ref = orc.swiglu_oai(orc._mm_f16_f32acc(xe, d13[e]).astype(np.float16).astype(np.float32)).astype(np.float16)
        assert_moe_close(h[off[e]:off[e + 1]], h_ref, intermediate=True)
        y_ref = orc._mm_f16_f32acc(h[off[e]:off[e + 1]], d2[e]).astype(np.float16)
        assert_moe_close(y[rows], y_ref)


@pytest.mark.parametrize("T,E,H,I,k", [(3, 8, 64, 64, 2), (40, 8, 256, 192, 4), (300, 8, 256, 192, 4), (700, 4, 192, 320, 2),
                                       (1100, 4, 128, 64, 1), (100, 4, 192, 128, 2), (200, 8, 128, 192, 4)])
@pytest.mark.parametrize("nested", [False, True])
@pytest.mark.parametrize("fast", [False, True])
def test_gemm_stages_vs_oracle(pkg, cuda_device, T, E, H, I, k, nested, fast):
    """tcgen05 grouped GEMMs (w13+SwiGLU, w2) against the oracle's per-expert fp16 matmul: ragged token tiles, row tiles
    that end inside the next expert (rows % 128 != 0), K tails that are not a multiple of the 4-block raw stage.  Both
    operand modes: EXACT (bit-exact dequant) and FAST (half2 arithmetic) must meet the north-star MoE tolerance."""
    _gemm_stage_check(pkg, cuda_device, T, E, H, I, k, nested, 51, pkg.REGIME_GEMM_FAST if fast else pkg.REGIME_GEMM)


@pytest.mark.parametrize("T", [16, 64, 8192])
def test_gemm_full_size_k_2880(pkg, cuda_device, T):
    """the real K = 2880 (45 k-blocks: 11 raw stages + a 1-block tail, A-stage ring wrapping inside an item) on a few
    experts, decode-sized and prefill-sized token tiles, both operand modes."""
    E, H, I, k = 4, 2880, 256, 2
    for fast in (False, True):
        _gemm_stage_check(pkg, cuda_device, T, E, H, I, k, True, 59, pkg.REGIME_GEMM_FAST if fast else pkg.REGIME_GEMM, std=0.02)


@pytest.mark.parametrize("T,E,H,I,k", [(65, 32, 128, 128, 4), (333, 8, 192, 128, 4), (1024, 32, 256, 128, 4)])
def test_moe_forward_gemm_regime_vs_oracle(pkg, cuda_device, T, E, H, I, k):
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 53, E, H, I)
    rng = np.random.default_rng(T)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    xd, lgd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(lg).to(cuda_device)
    out = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=pkg.REGIME_GEMM)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k)
    assert_moe_close(out.cpu().numpy(), ref)
    out_n = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=pkg.REGIME_GEMM, nested=True)
    assert torch.equal(out, out_n)
    # the two regimes differ only in fp32 summation order
    out_v = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=pkg.REGIME_GEMV)
    assert_moe_close(out.cpu().numpy(), out_v.cpu().numpy())
    # AUTO is exactly the regime nf4moe_auto_regime names: the decode tcgen05 kernel (FAST operand) while token tiles are
    # <= 32 rows, the EXACT prefill GEMM above
    auto = pkg.ops.auto_regime(T, k, E)
    assert auto == (pkg.REGIME_GEMM_FAST if expected_tile_rows(T, k, E) == 32 else pkg.REGIME_GEMM)
    out_a = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=pkg.REGIME_AUTO)
    assert torch.equal(out_a, pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=auto))
    assert_moe_close(out_a.cpu().numpy(), ref)
    # FAST operand mode: same layer, half2-arithmetic dequant, still inside the north-star tolerance of the oracle
    out_f = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=pkg.REGIME_GEMM_FAST)
    assert_moe_close(out_f.cpu().numpy(), ref)


def test_gemm_operand_is_bit_exact_dequant(pkg, cuda_device):
    """With one-hot activations the grouped GEMM returns the dequantised weight columns themselves: the A operand the
    tensor cores see must be bit-identical to bnb's dequantize_4bit (fp32 multiply, one rounding)."""
    E, H, I, k = 2, 128, 64, 1
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 57, E, H, I)
    T = H
    x = np.eye(H, dtype=np.float16)
    ids = np.zeros((T, 1), dtype=np.int32)
    ids[T // 2:] = 1
    b = pkg.ops.bucket(torch.from_numpy(ids).to(cuda_device), E)
    xs = pkg.ops.permute_rows(torch.from_numpy(x).to(cuda_device), b, k)
    raw = pkg.ops.expert_w13(xs, w13, b, T, k, act=pkg.ACT_CHUNK_SILU, regime=pkg.REGIME_GEMM).cpu().numpy()[:T]
    d13 = np.stack([orc.dequant_state(s) for s in st13])      # [E, 2I, H]
    for t in range(T):
        e = 0 if t < T // 2 else 1
        assert np.array_equal(raw[t].view(np.uint16), d13[e][:, t].view(np.uint16))


# ------------------------------------------------------------------------------------------------- loader
def test_loader_to_gpu_forward(pkg, cuda_device, tmp_path):
    """synthetic checkpoint in the reference's on-disk key set -> loader -> kernels: dequant bit-exact, layer within tolerance."""
    import importlib
    import os
    ld = importlib.import_module("gpt-oss-vllm_b200.loader")
    E, H, I, k, T = 8, 128, 64, 2, 9
    rng = np.random.default_rng(77)
    st13, st2 = make_states(rng, E, 2 * I, H, 0.05), make_states(rng, E, H, I, 0.05)
    path = os.path.join(tmp_path, "ckpt.safetensors")
    ld.save_layer_experts(path, 3, st13, st2)
    w13, w2 = ld.load_layer_experts([path], 3, E, H, I, device=cuda_device)
    got = w13.dequantize(nested=True).cpu().numpy()
    assert np.array_equal(got.view(np.uint16), np.stack([orc.dequant_state(s) for s in st13]).view(np.uint16))
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    out = pkg.ops.moe_forward(torch.from_numpy(x).to(cuda_device), w13, w2, k, router_logits=torch.from_numpy(lg).to(cuda_device))
    assert_moe_close(out.cpu().numpy(), orc.moe_forward_from_states(x, lg, st13, st2, k=k))


def test_graphed_forward_matches_eager(pkg, cuda_device):
    """GraphedForward (the serving-engine form: static buffers, one CUDA-graph replay per call, host in / host out) returns
    exactly what the eager apply_legacy calls return, for several different inputs through the same captured graph."""
    E, H, I, k, T = 32, 256, 128, 4, 24
    _, _, w13, w2 = make_layer(pkg, cuda_device, 77, E, H, I)
    method = pkg.B200NF4MoEMethod(top_k=k)
    layer = torch.nn.Module()
    method.bind(layer, w13, w2)
    g = torch.Generator().manual_seed(3)
    mk = lambda: (torch.randn(T, H, generator=g).half().pin_memory(), torch.randn(T, E, generator=g).half().pin_memory())
    x0, lg0 = mk()
    graphed = pkg.GraphedForward(lambda x, lg: method.apply_legacy(layer, x, lg, k), [x0.to(cuda_device), lg0.to(cuda_device)])
    for _ in range(3):
        x, lg = mk()
        got = graphed(x, lg)
        ref = method.apply_legacy(layer, x.to(cuda_device), lg.to(cuda_device), k)
        assert torch.equal(got.to(cuda_device), ref)
    with pytest.raises(ValueError):
        graphed(x[:5], lg)


@pytest.mark.parametrize("T,E,H,I,k,regime", [(5, 8, 128, 128, 2, "gemv"), (40, 8, 256, 192, 4, "auto"), (300, 8, 256, 192, 4, "gemm"),
                                              (64, 32, 320, 128, 4, "auto")])
@pytest.mark.parametrize("act", ["swiglu_oai", "chunk_silu"])
def test_moe_forward_with_expert_biases(pkg, cuda_device, T, E, H, I, k, regime, act):
    """gpt-oss expert biases: w13_bias before the activation, w2_bias before the routing weight, in every
    regime and both activation layouts, against the oracle; zero biases reproduce the unbiased layer bit for bit; the
    quant-method picks the biases up from the layer's vLLM parameter names."""
    st13, st2, w13, w2 = make_layer(pkg, cuda_device, 91, E, H, I)
    rng = np.random.default_rng(T + E)
    x = rng.standard_normal((T, H)).astype(np.float16)
    lg = rng.standard_normal((T, E)).astype(np.float16)
    b13 = (rng.standard_normal((E, 2 * I)) * 0.5).astype(np.float16)
    b2 = (rng.standard_normal((E, H)) * 0.5).astype(np.float16)
    reg = {"gemv": pkg.REGIME_GEMV, "gemm": pkg.REGIME_GEMM, "auto": pkg.REGIME_AUTO}[regime]
    a = pkg.ACT_SWIGLU_OAI if act == "swiglu_oai" else pkg.ACT_CHUNK_SILU
    xd, lgd = torch.from_numpy(x).to(cuda_device), torch.from_numpy(lg).to(cuda_device)
    b13d, b2d = torch.from_numpy(b13).to(cuda_device), torch.from_numpy(b2).to(cuda_device)
    out = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=reg, act=a, w13_bias=b13d, w2_bias=b2d)
    ref = orc.moe_forward_from_states(x, lg, st13, st2, k=k, act=act, b13=b13, b2=b2)
    assert_moe_close(out.cpu().numpy(), ref)
    plain = orc.moe_forward_from_states(x, lg, st13, st2, k=k, act=act)
    assert np.abs(ref.astype(np.float32) - plain.astype(np.float32)).max() > 0.05          # the biases matter in this case
    # each bias alone
    for kw, okw in (({"w13_bias": b13d}, {"b13": b13}), ({"w2_bias": b2d}, {"b2": b2})):
        o1 = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=reg, act=a, **kw)
        assert_moe_close(o1.cpu().numpy(), orc.moe_forward_from_states(x, lg, st13, st2, k=k, act=act, **okw))
    # zero w2 bias == no w2 bias, bit for bit (the w13 path differs by design: fused vs raw + bias kernel)
    o0 = pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=reg, act=a, w2_bias=torch.zeros_like(b2d))
    assert torch.equal(o0, pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, regime=reg, act=a))
    if act == "swiglu_oai":
        method = pkg.B200NF4MoEMethod(top_k=k, regime=reg)
        layer = torch.nn.Module()
        method.bind(layer, w13, w2)
        layer.w13_bias = torch.nn.Parameter(b13d, requires_grad=False)
        layer.w2_bias = torch.nn.Parameter(b2d, requires_grad=False)
        assert torch.equal(method.apply_legacy(layer, xd, lgd, k), out)
    with pytest.raises(ValueError):
        pkg.ops.moe_forward(xd, w13, w2, k, router_logits=lgd, w2_bias=b2d[:, :-8].contiguous())
