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
        # VLLM_ATTENTION_BACKEND="XFORMERS" selects the stock backend (7B / non-Hymeta models)
        if _is_xformers(selected_backend):
            logger.info("Using FlashAttentionBackend for 7B model.")
            return "vllm.attention.backends.flash_attn.FlashAttentionBackend"
        logger.info("Using MetaAttentionBackend (sm_100a kernels) for 70B model. "
                    "To use the 7B model or a non-hymeta model set VLLM_ATTENTION_BACKEND=\"XFORMERS\"")
        return "vllm_hymeta_b200.attention.meta_attention.MetaAttentionBackend"
