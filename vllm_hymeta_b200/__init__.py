"""vllm_hymeta_b200 -- B200 (sm_100a) native drop-in for the hybrid-attention hot path of the
vllm_hymeta plugin: gated linear attention (GLA) and sliding-window attention with the
meta-token prefix, behind the plugin's own entry points (reference vllm_hymeta/__init__.py:2-11).

Importing this package touches neither CUDA nor the kernel library (the plugin entry points run
in every vLLM process before CUDA init -- reference platform.py:7).
"""


def register():
    """vllm.platform_plugins entry point (reference vllm_hymeta/__init__.py:2-4)."""
    return "vllm_hymeta_b200.platform.HymetaCudaPlatform"


def register_model():
    """vllm.general_plugins entry point (reference vllm_hymeta/__init__.py:6-11).  Besides the two
    model registrations it puts the V1 adapter of the backend into the installed vLLM's
    ``AttentionBackendEnum.CUSTOM`` slot when that registry exists (vLLM >= 0.11)."""
    from .attention.meta_attention_v1 import register_v1_backend
    register_v1_backend()
    from .model_for_7B import register_7B_model
    register_7B_model()
    from .models import register_model as _register
    _register()
