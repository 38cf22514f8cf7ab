"""CPU oracle for the Hymeta hybrid-attention hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: it may be
imported by ``tests/``, by ``__graft_entry__.smoke()`` and by the ``cpu_baseline``
/ ``--impl reference`` legs of ``bench.py`` -- as the checker or the timed CPU
baseline -- and by nothing else.  The product path (``vllm_hymeta_b200``) never
imports it and has no CPU fallback.

The oracle is a plain fp32 PyTorch/numpy restatement of the reference's torch
path (reference = FAndromedA/vllm-hymeta, pure Python; citations are
``file:line`` relative to the reference root):

* ``oracle.gla``  -- gated-linear-attention recurrence, gate pre-processing,
  slot-state prefill/decode semantics
  (vllm_hymeta/models/my_fused_recurrent.py:58-80,
  vllm_hymeta/models/modeling_hymeta.py:441-659,
  vllm_hymeta/model_for_7B/gla_attention.py:179-283).
* ``oracle.swa``  -- sliding-window attention with the meta-token prefix,
  LSE merge, paged-KV write and paged decode
  (vllm_hymeta/attention/fa_kernel_metatoken.py:110-122,124-285,377-410,
  vllm_hymeta/attention/meta_attention.py:703-982).

Third-party arithmetic the reference calls but does not vendor
(flash-linear-attention==0.1 ``fused_chunk_gla``; flash_attn==2.7.3
``_flash_attn_varlen_forward`` / ``fwd_kvcache``) is restated from its
published definition: the GLA step recurrence (identical to the installed
``fla/ops/gla/naive.py:15-44``) and masked softmax attention (identical to the
reference's own ``naive_metatoken_flash_attn``).

Parity pinning: the reference ships no golden vectors, so
the oracle is pinned against outputs of the reference's own Python run in the
authoring container: ``oracle/make_golden.py`` imports the unmodified
reference modules from /root/reference (plus fla's ``naive_recurrent_gla``),
and writes ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every
oracle function against them.
"""
