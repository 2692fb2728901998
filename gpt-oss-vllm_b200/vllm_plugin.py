"""vLLM registration shim: makes `--quantization b200_nf4_moe` resolve to this package's method, replacing
the reference's string edits of an installed vLLM (P/patch_6_add_bnb_moe.py, P/patch_7_allow_moe_bnb.py,
P/patch_9_fusedmoe_unsloth.py) and its router monkey patch (P/monkey_patch_router_fp16.py) with a plugin that needs no
source edits.  vLLM is imported lazily: the rest of the package never depends on it.

    import nf4moe_b200.vllm_plugin as plug; plug.register()        # e.g. from a vllm.general_plugins entry point

What it registers is a `QuantizationConfig` whose `get_quant_method` hands FusedMoE / RoutedExperts layers a method object that
IS a `FusedMoEMethodBase` (vLLM asserts that, V/model_executor/layers/fused_moe/layer.py:363 -- the reference's
P/bnb_moe_4bit.py:43-48 returned the string "bnb_4bit" there, which is the bug the north-star asks to fix) and leaves every
other layer to whatever the outer configuration decides.
"""
from __future__ import annotations

import torch

from . import _lib
from .moe_method import B200NF4MoEConfig, B200NF4MoEMethod

NAME = "b200_nf4_moe"
OPERAND_MODES = {"fast": _lib.REGIME_AUTO,     # GEMV up to 8 tokens, tcgen05 kernels with the half2 operand above
                 "exact": _lib.REGIME_GEMM}    # tcgen05 kernels with the bit-exact (bnb dequantize_4bit) operand at every T
_registered = None
_method_cls = None


def _is_moe_layer(layer) -> bool:
    names = {c.__name__ for c in type(layer).__mro__}
    return bool(names & {"FusedMoE", "RoutedExperts", "SharedFusedMoE"})


def build_method_class():
    """`VllmNF4MoEMethod(FusedMoEMethodBase)`; needs vLLM importable.  Built once."""
    global _method_cls
    if _method_cls is not None:
        return _method_cls
    from vllm.model_executor.layers.fused_moe.fused_moe_method_base import FusedMoEMethodBase

    class VllmNF4MoEMethod(FusedMoEMethodBase):
        """The method object vLLM's MoE runner calls (V/model_executor/layers/fused_moe/runner/moe_runner.py:498-539), same
        entry points and argument meaning as vLLM's `BitsAndBytesMoEMethod` (V/model_executor/layers/quantization/
        bitsandbytes.py:435-603); the arithmetic is libnf4moe's (no dequantise-everything, no Triton)."""

        def __init__(self, quant_config: B200NF4MoEConfig, moe, operand_mode: str = "fast", top_k: int | None = None):
            super().__init__(moe)
            if operand_mode not in OPERAND_MODES:
                raise ValueError(f"operand_mode must be one of {sorted(OPERAND_MODES)}, got {operand_mode!r}")
            self.quant_config = quant_config
            self.operand_mode = operand_mode
            self.inner = B200NF4MoEMethod(quant_config, moe=moe, top_k=top_k or getattr(moe, "experts_per_token", None),
                                          regime=OPERAND_MODES[operand_mode])

        # vLLM's router (bias, renormalise, grouped top-k ...) picks the experts and hands over (topk_weights, topk_ids)
        @property
        def is_monolithic(self) -> bool:
            return False

        def create_weights(self, layer, num_experts: int, hidden_size: int, intermediate_size_per_partition: int,
                           params_dtype: torch.dtype, **extra_weight_attrs):
            self.inner.create_weights(layer, num_experts, hidden_size, intermediate_size_per_partition, params_dtype,
                                      has_bias=bool(getattr(self.moe, "has_bias", False)), **extra_weight_attrs)

        def get_fused_moe_quant_config(self, layer):   # this method owns its kernels: nothing for vLLM's modular kernels
            return None

        def process_weights_after_loading(self, layer) -> None:
            self.inner.process_weights_after_loading(layer)

        def apply(self, layer, x, topk_weights, topk_ids, shared_experts=None, shared_experts_input=None):
            # shared_experts are run by the MoE runner around this call (NO_OVERLAP / MULTI_STREAM orders); like vLLM's own
            # bnb method this one has no internal overlap mode and leaves them alone
            return self.inner.apply(layer, x, topk_weights, topk_ids)

        def apply_monolithic(self, layer, x, router_logits, input_ids=None):
            return self.inner.apply_monolithic(layer, x, router_logits, input_ids)

    _method_cls = VllmNF4MoEMethod
    return _method_cls


def build_config_class():
    """Create the vLLM-facing config class (needs vLLM importable)."""
    from vllm.model_executor.layers.quantization.base_config import QuantizationConfig

    class B200NF4MoEVllmConfig(QuantizationConfig):
        """P/bnb_moe_4bit.py:22-56 as a real vLLM QuantizationConfig.  `operand_mode`: "fast" (default) or "exact"."""

        def __init__(self, top_k: int | None = None, operand_mode: str = "fast"):
            super().__init__()
            if operand_mode not in OPERAND_MODES:
                raise ValueError(f"operand_mode must be one of {sorted(OPERAND_MODES)}, got {operand_mode!r}")
            self.inner = B200NF4MoEConfig()
            self.top_k, self.operand_mode = top_k, operand_mode

        def get_name(self):
            return NAME

        def get_supported_act_dtypes(self):
            return [torch.float16, torch.bfloat16]

        @classmethod
        def get_min_capability(cls) -> int:
            return 100

        @staticmethod
        def get_config_filenames():
            return []

        @classmethod
        def from_config(cls, config):
            return cls(top_k=config.get("top_k"), operand_mode=config.get("operand_mode", "fast"))

        def get_quant_method(self, layer, prefix: str):
            if _is_moe_layer(layer) or self.inner.is_moe_weight(prefix):
                moe = getattr(layer, "moe_config", None)
                return build_method_class()(self.inner, moe, operand_mode=self.operand_mode,
                                            top_k=self.top_k or getattr(layer, "top_k", None))
            return None   # non-expert layers: unquantised / whatever the outer config decides

    return B200NF4MoEVllmConfig


def register() -> type:
    """Register with vLLM (idempotent).  Raises ImportError if vLLM is not installed."""
    global _registered
    if _registered is None:
        from vllm.model_executor.layers.quantization import register_quantization_config
        _registered = register_quantization_config(NAME)(build_config_class())
    return _registered
