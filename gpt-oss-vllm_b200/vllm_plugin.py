"""vLLM registration shim: makes `--quantization b200_nf4_moe` resolve to this package's method, replacing
the reference's string edits of an installed vLLM (P/patch_6_add_bnb_moe.py, P/patch_7_allow_moe_bnb.py,
P/patch_9_fusedmoe_unsloth.py) and its router monkey patch (P/monkey_patch_router_fp16.py) with a plugin that needs no
source edits.  vLLM is imported lazily: the rest of the package never depends on it.

    import nf4moe_b200.vllm_plugin as plug; plug.register()        # e.g. from a vllm.general_plugins entry point

What it registers is a `QuantizationConfig` whose `get_quant_method` hands FusedMoE / RoutedExperts layers a
`B200NF4MoEMethod` (V/model_executor/layers/quantization/bitsandbytes.py:159-168 does the same for its own method) and
leaves every other layer to vLLM's stock bitsandbytes linear method.
"""
from __future__ import annotations

import torch

from .moe_method import B200NF4MoEConfig, B200NF4MoEMethod

NAME = "b200_nf4_moe"
_registered = None


def _is_moe_layer(layer) -> bool:
    names = {c.__name__ for c in type(layer).__mro__}
    return bool(names & {"FusedMoE", "RoutedExperts", "SharedFusedMoE"})


def build_config_class():
    """Create the vLLM-facing config class (needs vLLM importable)."""
    from vllm.model_executor.layers.quantization.base_config import QuantizationConfig

    class B200NF4MoEVllmConfig(QuantizationConfig):
        """P/bnb_moe_4bit.py:22-56 as a real vLLM QuantizationConfig."""

        def __init__(self, top_k: int | None = None, operand_mode: str = "exact"):
            super().__init__()
            self.inner = B200NF4MoEConfig()
            self.top_k, self.operand_mode = top_k, operand_mode

        def get_name(self):
            return NAME

        def get_supported_act_dtypes(self):
            return [torch.float16]

        @classmethod
        def get_min_capability(cls) -> int:
            return 100

        @staticmethod
        def get_config_filenames():
            return []

        @classmethod
        def from_config(cls, config):
            return cls(top_k=config.get("top_k"), operand_mode=config.get("operand_mode", "exact"))

        def get_quant_method(self, layer, prefix: str):
            if _is_moe_layer(layer) or self.inner.is_moe_weight(prefix):
                moe = getattr(layer, "moe_config", None)
                return VllmNF4MoEMethod(self.inner, moe=moe, top_k=self.top_k or getattr(layer, "top_k", None))
            return None   # non-expert layers: unquantised / whatever the outer config decides

    return B200NF4MoEVllmConfig


class VllmNF4MoEMethod(B200NF4MoEMethod):
    """The method object vLLM's MoE runner calls (V/model_executor/layers/fused_moe/runner/moe_runner.py:498-539).  Adds the
    pass-through arguments of `fused_experts` that the gpt-oss path leaves at their defaults
    (V/model_executor/layers/quantization/bitsandbytes.py:493-505) and refuses the ones it does not implement."""

    def apply(self, layer, x, topk_weights, topk_ids, shared_experts=None, shared_experts_input=None):
        if getattr(layer, "expert_map", None) is not None:
            # vLLM's own EP remaps global -> local ids with -1 for foreign experts; the kernels drop ids outside [0, E)
            topk_ids = layer.expert_map[topk_ids.long()].to(torch.int32)
        out = super().apply(layer, x, topk_weights, topk_ids)
        if shared_experts is not None:
            raise NotImplementedError("gpt-oss has no shared experts")
        return out

    def get_fused_moe_quant_config(self, layer):   # vLLM asks; this method owns its kernels, nothing to configure
        return None


def register() -> type:
    """Register with vLLM (idempotent).  Raises ImportError if vLLM is not installed."""
    global _registered
    if _registered is None:
        from vllm.model_executor.layers.quantization import register_quantization_config
        _registered = register_quantization_config(NAME)(build_config_class())
    return _registered
