"""Platform plugin -- same class name and selection rule as the reference
vllm_hymeta/platform.py:10-30.  Imported by vLLM in every process BEFORE CUDA is initialised, so
this module must not touch CUDA or the kernel library."""
from __future__ import annotations

import logging

logger = logging.getLogger(__name__)

try:  # vLLM's NVML-based CUDA platform (does not initialise CUDA; reference platform.py:7)
    from vllm.platforms.cuda import NvmlCudaPlatform as _Base  # type: ignore
    from vllm.platforms import PlatformEnum  # type: ignore
except Exception:  # pragma: no cover - vLLM absent or incompatible
    _Base, PlatformEnum = object, None


def _is_xformers(selected_backend) -> bool:
    # vLLM 0.10.0 passes a ``_Backend`` enum member; newer versions an ``AttentionBackendEnum``
    # or a string -- compare by name so the rule survives the rename
    name = getattr(selected_backend, "name", selected_backend)
    return isinstance(name, str) and name.upper() == "XFORMERS"


class HymetaCudaPlatform(_Base):
    """Platform for the Hymeta caches on CUDA."""

    if PlatformEnum is not None:
        _enum = PlatformEnum.CUDA

    @classmethod
    def get_attn_backend_cls(cls, selected_backend, head_size=None, dtype=None, kv_cache_dtype=None,
                             block_size=None, use_v1=False, use_mla=False, *args, **kwargs) -> str:
        # vLLM >= 0.11 (V1 only) calls get_attn_backend_cls(selected_backend, attn_selector_config,
        # num_heads): the second positional argument is then a config object, not a head size
        if head_size is not None and not isinstance(head_size, int):
            return cls._get_attn_backend_cls_v1(selected_backend, head_size, *([dtype] if dtype is not None else []),
                                                **kwargs)
        # VLLM_ATTENTION_BACKEND="XFORMERS" selects the stock backend (7B / non-Hymeta models)
        if _is_xformers(selected_backend):
            logger.info("Using FlashAttentionBackend for 7B model.")
            return "vllm.attention.backends.flash_attn.FlashAttentionBackend"
        logger.info("Using MetaAttentionBackend (sm_100a kernels) for 70B model. "
                    "To use the 7B model or a non-hymeta model set VLLM_ATTENTION_BACKEND=\"XFORMERS\"")
        return "vllm_hymeta_b200.attention.meta_attention.MetaAttentionBackend"

    @classmethod
    def _get_attn_backend_cls_v1(cls, selected_backend, attn_selector_config, *args, **kwargs) -> str:
        """The installed vLLM's V1 selection protocol: an explicitly selected stock backend (any
        ``--attention-backend`` other than CUSTOM, the successor of the reference's
        ``VLLM_ATTENTION_BACKEND="XFORMERS"`` switch) goes to vLLM's own selector; otherwise the
        V1 adapter of the meta-token backend."""
        name = getattr(selected_backend, "name", selected_backend)
        if isinstance(name, str) and name.upper() != "CUSTOM":
            logger.info("Using vLLM's %s backend (stock path for 7B / non-hymeta models).", name)
            return super().get_attn_backend_cls(selected_backend, attn_selector_config, *args, **kwargs)
        if getattr(attn_selector_config, "use_mla", False):
            return super().get_attn_backend_cls(selected_backend, attn_selector_config, *args, **kwargs)
        logger.info("Using MetaAttentionV1Backend (sm_100a kernels).")
        return "vllm_hymeta_b200.attention.meta_attention_v1.MetaAttentionV1Backend"
