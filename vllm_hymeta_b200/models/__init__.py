"""70B (intra-layer hybrid) side of the plugin.  ``register_model`` keeps the reference's name
(vllm_hymeta/models/__init__.py:4-13).  The model graph, MoE and weight loader are vLLM-side code
and out of scope: when the reference plugin is importable its model classes are
registered with their hot path re-pointed at the sm_100a kernels (see ``patch_reference``)."""
from __future__ import annotations


def patch_reference() -> bool:
    """Re-point the reference plugin's hot-path symbols at this package, in place.  Returns False
    when ``vllm_hymeta`` (or the vLLM it needs) is not importable."""
    try:
        import vllm_hymeta.models.modeling_hymeta as ref  # type: ignore
    except Exception:
        return False
    from ..attention import fa_kernel_metatoken as fk
    from ..attention import meta_attention as ma
    from .fused_chunk_gla import fused_chunk_gla
    from .hymeta_cache import HymetaCacheManager, HymetaCacheParams
    from .my_fused_recurrent_gla import my_fused_recurrent_gla
    ref.fused_chunk_gla = fused_chunk_gla
    ref.my_fused_recurrent_gla = my_fused_recurrent_gla
    ref.HymetaCacheManager, ref.HymetaCacheParams = HymetaCacheManager, HymetaCacheParams
    try:
        import vllm_hymeta.attention.meta_attention as ref_ma  # type: ignore
        ref_ma.metatoken_flash_attn_varlen_func = fk.metatoken_flash_attn_varlen_func
        ref_ma.metatoken_flash_attn_with_kvcache = fk.metatoken_flash_attn_with_kvcache
        ref_ma.MetaAttentionImpl.forward = ma.MetaAttentionImpl.forward
    except Exception:
        pass
    return True


def register_model():
    if not patch_reference():
        return
    from transformers import AutoConfig
    from vllm import ModelRegistry
    from vllm_hymeta.models.configuration_hymeta import HymetaConfig  # type: ignore
    try:
        AutoConfig.register("hybrid", HymetaConfig)
    except ValueError:
        pass                                   # already registered by the reference plugin
    ModelRegistry.register_model("HymetaForCausalLM",
                                 "vllm_hymeta.models.modeling_hymeta:HymetaForCausalLM")
