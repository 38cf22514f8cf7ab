"""Drop-in for ``fla.ops.gla.fused_chunk_gla`` at the reference's call sites
(vllm_hymeta/models/modeling_hymeta.py:463,512; model_for_7B/gla_attention.py:206): head-first
``[1, H, T, D]`` tensors, optional ``initial_state [1, H, D, D]``, returns
``(o [1,H,T,D], final_state [1,H,D,D] fp32 | None)``.  The tcgen05 varlen kernel is called on
token-major VIEWS (no transpose copies).  Prefer ``ops.gla_prefill_varlen`` directly: it takes the
whole packed batch, slot states and un-repeated k/v/g in one launch."""
from __future__ import annotations

import torch

from .. import ops


def fused_chunk_gla(q, k, v, g, scale=None, initial_state=None, output_final_state=False):
    assert q.dim() == 4 and q.shape[0] == 1, "one request per call, as in the reference loop"
    _, H, T, D = q.shape
    Hk = k.shape[1]
    dev = q.device
    qt, kt, vt, gt = (x[0].transpose(0, 1) for x in (q, k, v, g))      # [T, H, D] views
    state = torch.zeros(1, Hk if Hk != H else H, D, D, dtype=torch.bfloat16, device=dev)
    mode = 0
    if initial_state is not None:
        state[0].copy_(initial_state[0][:: H // state.shape[1]])
        mode = 1
    i32 = lambda x: torch.tensor(x, dtype=torch.int32, device=dev)
    o = ops.gla_prefill_varlen(qt, kt, vt, gt, state, i32([0, T]), i32([0]), i32([mode]),
                               scale=scale, write_state=output_final_state, max_seqlen=T)
    final = None
    if output_final_state:
        final = state.float()
        if final.shape[1] != H:
            final = final.repeat_interleave(H // final.shape[1], dim=1)
    return o.transpose(0, 1).unsqueeze(0), final
