"""CPU: the safetensors loader round-trips the reference's on-disk key set without touching the packed
bytes, fails loudly on malformed checkpoints, and the vLLM registration shim (8(f).2) registers a real QuantizationConfig."""
import importlib
import os

import numpy as np
import pytest
import torch

from oracle import nf4_oracle as orc


def _states(rng, E, rows, cols):
    return [orc.quantize_nf4((rng.standard_normal((rows, cols)) * 0.02).astype(np.float16)) for _ in range(E)]


def test_loader_round_trip_keeps_bytes(tmp_path):
    ld = importlib.import_module("gpt-oss-vllm_b200.loader")
    rng = np.random.default_rng(3)
    E, H, I = 4, 128, 64
    st13, st2 = _states(rng, E, 2 * I, H), _states(rng, E, H, I)
    path = os.path.join(tmp_path, "layer7.safetensors")
    ld.save_layer_experts(path, 7, st13, st2)
    from safetensors import safe_open
    with safe_open(path, framework="pt", device="cpu") as f:        # the key set of P/unsloth_bnb_loader.py:196-227
        keys = set(f.keys())
    p = "model.layers.7.mlp.experts.gate_up_projs.0.weight"
    assert {p, p + ".absmax", p + ".nested_absmax", p + ".nested_quant_map", p + ".quant_map"} <= keys
    assert "model.layers.7.mlp.experts.down_projs.3.weight" in keys
    w13, w2 = ld.load_layer_experts([path], 7, E, H, I)
    assert w13.packed.shape == (E, 2 * I * H // 2, 1) and w2.packed.shape == (E, H * I // 2, 1)   # vLLM's fused layout
    for w, sts in ((w13, st13), (w2, st2)):
        for e, s in enumerate(sts):
            assert np.array_equal(w.packed[e].numpy().reshape(-1), s["packed"])          # never unpacked / re-quantised
            assert np.array_equal(w.absmax_u8[e].numpy(), s["absmax_u8"])
            assert np.array_equal(w.absmax2[e].numpy().view(np.uint32), s["absmax2"].view(np.uint32))
            assert np.float32(w.offset[e].item()) == np.float32(s["offset"])
            back = {"packed": w.packed[e].numpy().reshape(-1), "absmax_u8": w.absmax_u8[e].numpy(), "absmax2": w.absmax2[e].numpy(),
                    "code2": w.code2.numpy(), "offset": np.float32(w.offset[e].item()), "shape": s["shape"], "blocksize": 64}
            assert np.array_equal(orc.dequant_state(back).view(np.uint16), orc.dequant_state(s).view(np.uint16))
    sub13, _ = ld.load_layer_experts([path], 7, E, H, I, experts=range(2, 4))              # an EP rank's slice
    assert sub13.num_experts == 2 and np.array_equal(sub13.packed[0].numpy().reshape(-1), st13[2]["packed"])


def test_loader_fails_loudly(tmp_path):
    ld = importlib.import_module("gpt-oss-vllm_b200.loader")
    rng = np.random.default_rng(4)
    st13, st2 = _states(rng, 2, 64, 64), _states(rng, 2, 64, 32 * 2)
    path = os.path.join(tmp_path, "l0.safetensors")
    ld.save_layer_experts(path, 0, st13, st2)
    with pytest.raises(ld.CheckpointFormatError):           # a missing expert is an error, never a silent zero fill
        ld.load_layer_experts([path], 0, 3, 64, 32)
    with pytest.raises(ld.CheckpointFormatError):           # wrong geometry
        ld.load_layer_experts([path], 0, 2, 128, 32)
    good = {"x": torch.zeros(64 * 64 // 2, 1, dtype=torch.uint8), "x.absmax": torch.zeros(64, dtype=torch.uint8),
            "x.nested_absmax": torch.zeros(1), "x.nested_quant_map": torch.zeros(256), "x.quant_map": torch.zeros(16)}
    with pytest.raises(ld.CheckpointFormatError):           # not the NF4 table
        ld.load_expert_matrix(good.get, "x", 64, 64, offset_override=0.0)
    good["x.quant_map"] = torch.tensor(ld.NF4_TABLE)
    with pytest.raises(ld.CheckpointFormatError):           # no offset anywhere (the reference dropped it: SURVEY 0.2b)
        ld.load_expert_matrix(good.get, "x", 64, 64)
    assert ld.load_expert_matrix(good.get, "x", 64, 64, offset_override=0.25)["offset"] == 0.25


def _vllm_or_skip():
    pytest.importorskip("vllm")
    plug = importlib.import_module("gpt-oss-vllm_b200.vllm_plugin")
    try:
        cls = plug.register()
    except Exception as exc:   # vLLM import needs CUDA libs on some boxes
        pytest.skip(f"vLLM could not be imported here: {type(exc).__name__}: {exc}")
    return plug, cls


def test_vllm_plugin_registers_config():
    plug, cls = _vllm_or_skip()
    from vllm.model_executor.layers.fused_moe.fused_moe_method_base import FusedMoEMethodBase
    from vllm.model_executor.layers.quantization import get_quantization_config
    assert get_quantization_config(plug.NAME) is cls and plug.register() is cls
    cfg = cls.from_config({"top_k": 4})
    assert cfg.get_name() == plug.NAME and cfg.get_min_capability() == 100 and cfg.get_supported_act_dtypes() == [torch.float16, torch.bfloat16]
    assert cfg.operand_mode == "fast" and cls.from_config({"operand_mode": "exact"}).operand_mode == "exact"
    with pytest.raises(ValueError):
        cls.from_config({"operand_mode": "sloppy"})

    class FusedMoE(torch.nn.Module):
        pass
    m = cfg.get_quant_method(FusedMoE(), "model.layers.0.mlp.experts")
    assert isinstance(m, FusedMoEMethodBase)            # what V/model_executor/layers/fused_moe/layer.py:363 asserts
    assert isinstance(m, plug.build_method_class()) and m.inner.top_k == 4 and m.is_monolithic is False
    lib = importlib.import_module("gpt-oss-vllm_b200._lib")
    assert m.inner.regime == lib.REGIME_AUTO                                           # operand_mode is honoured, not ignored:
    m2 = cls.from_config({"operand_mode": "exact"}).get_quant_method(FusedMoE(), "model.layers.0.mlp.experts")
    assert m2.inner.regime == lib.REGIME_GEMM                                          # "exact" -> the bit-exact tcgen05 operand
    assert cfg.get_quant_method(torch.nn.Linear(2, 2), "model.layers.0.self_attn.qkv_proj") is None


def build_real_fused_moe(cls, device="cpu", E=4, H=128, I=64, top_k=2, has_bias=True, operand_mode="fast", prefix=None,
                         apply_router_weight_on_input=False):
    """A REAL vllm FusedMoE layer whose quant_config is the plugin's config (what P/patch_6_add_bnb_moe.py:75-107 tried to
    reach with `FusedMoE(quant_config=moe_quant_config)`).  Returns (layer, vllm_config)."""
    from vllm.config import DeviceConfig, VllmConfig, set_current_vllm_config
    from vllm.distributed import init_distributed_environment, initialize_model_parallel
    from vllm.distributed.parallel_state import model_parallel_is_initialized
    from vllm.model_executor.layers.fused_moe import FusedMoE
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    vc = VllmConfig(device_config=DeviceConfig(device=device))
    with set_current_vllm_config(vc):
        if not model_parallel_is_initialized():
            init_distributed_environment(world_size=1, rank=0, local_rank=0, distributed_init_method="tcp://127.0.0.1:29655",
                                         backend="gloo" if device == "cpu" else "nccl")
            initialize_model_parallel(1, 1)
        build_real_fused_moe.n = getattr(build_real_fused_moe, "n", 0) + 1
        layer = FusedMoE(num_experts=E, top_k=top_k, hidden_size=H, intermediate_size=I, params_dtype=torch.float16,
                         renormalize=True, quant_config=cls(operand_mode=operand_mode),
                         prefix=prefix or f"model.layers.{build_real_fused_moe.n}.mlp.experts", activation="swigluoai",
                         has_bias=has_bias, apply_router_weight_on_input=apply_router_weight_on_input)
    return layer, vc


def test_real_vllm_fused_moe_accepts_the_method():
    """vLLM 0.22 constructs FusedMoE with this plugin's quant method (no dummy class): isinstance holds, the parameters are the
    ones V/model_executor/layers/quantization/bitsandbytes.py:508-575 registers, plus the gpt-oss bias pair."""
    plug, cls = _vllm_or_skip()
    from vllm.model_executor.layers.fused_moe.fused_moe_method_base import FusedMoEMethodBase
    try:
        layer, _ = build_real_fused_moe(cls)
    except Exception as exc:
        pytest.skip(f"vLLM cannot build a FusedMoE on this box: {type(exc).__name__}: {exc}")
    assert isinstance(layer.quant_method, FusedMoEMethodBase) and isinstance(layer.quant_method, plug.build_method_class())
    shapes = {n: tuple(p.shape) for n, p in layer.named_parameters()}
    assert shapes == {"w13_weight": (4, 2 * 64 * 128 // 2, 1), "w2_weight": (4, 128 * 64 // 2, 1), "w13_bias": (4, 128),
                      "w2_bias": (4, 128)}
    assert layer.w13_weight.experts_shape == (4, 128, 128) and layer.w13_weight.use_bitsandbytes_4bit
    assert layer.w13_weight.dtype == torch.uint8 and callable(layer.w13_weight.weight_loader)
    assert layer.quant_method.get_fused_moe_quant_config(layer) is None
