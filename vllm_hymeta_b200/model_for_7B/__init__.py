"""7B (inter-layer hybrid) side of the plugin.  ``register_7B_model`` keeps the reference's name
(vllm_hymeta/model_for_7B/__init__.py:4-13); see ``models/__init__.py`` for the scope note."""
from __future__ import annotations


def patch_reference() -> bool:
    try:
        import vllm_hymeta.model_for_7B.gla_attention as ref  # type: ignore
    except Exception:
        return False
    from ..models.fused_chunk_gla import fused_chunk_gla
    from .gla_cache import GLACacheManager, GLACacheParams
    from .my_fused_recurrent import my_fused_recurrent_gla
    ref.fused_chunk_gla = fused_chunk_gla
    ref.my_fused_recurrent_gla = my_fused_recurrent_gla
    try:
        import vllm_hymeta.model_for_7B.modeling_gla_swa as ref_m  # type: ignore
        ref_m.GLACacheManager, ref_m.GLACacheParams = GLACacheManager, GLACacheParams
    except Exception:
        pass
    return True


def register_7B_model():
    if not patch_reference():
        return
    from transformers import AutoConfig
    from vllm import ModelRegistry
    from vllm_hymeta.model_for_7B.configuration_gla_swa import GLAswaConfig  # type: ignore
    try:
        AutoConfig.register("gla_swa", GLAswaConfig)
    except ValueError:
        pass
    ModelRegistry.register_model("GLAswaForCausalLM",
                                 "vllm_hymeta.model_for_7B.modeling_gla_swa:GLAswaForCausalLM")
