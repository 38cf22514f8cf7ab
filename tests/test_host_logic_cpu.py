"""Host-side mirror of the reference interface, on CPU: slot allocator (vLLM ConstantSizeCache
behaviour the reference cache managers inherit), cache managers, metadata builder (bit-exact
slot mapping / block tables), backend surface and error behaviour, plugin entry points."""
import pytest
import torch

from vllm_hymeta_b200.attention.meta_attention import (MetaAttentionBackend, MetaAttentionImpl,
                                                       MetaAttentionMetadataBuilder, SeqData)
from vllm_hymeta_b200.model_for_7B.gla_cache import GLACacheManager
from vllm_hymeta_b200.models.hymeta_cache import HymetaCacheManager, HymetaCacheParams
from vllm_hymeta_b200.slot_cache import PAD_SLOT_ID, SlotAllocator


def test_slot_allocator_lifecycle():
    a = SlotAllocator(4)
    copies = []
    cp = lambda s, d: copies.append((s, d))
    s0 = a.prepare({"r0": [0], "r1": [1]}, [], cp)
    assert len(set(s0)) == 2 and all(0 <= s < 4 for s in s0)
    # same requests keep their slots; a sibling sequence forks the state of its request
    s1 = a.prepare({"r0": [0, 7], "r1": [1]}, [], cp)
    assert s1[0] == s0[0] and s1[2] == s0[1] and s1[1] not in s0
    assert copies == [(s0[0], s1[1])]
    # finishing r0 frees both of its slots; a new request may reuse them
    s2 = a.prepare({"r1": [1], "r2": [2], "r3": [3]}, ["r0"], cp)
    assert s2[0] == s0[1] and set(s2[1:]) <= {s0[0], s1[1], 0, 1, 2, 3} - {s0[1]}
    assert len(a.free_cache_indices) == 1
    # a request finished before it ever ran gets the PAD slot
    assert a.prepare({"r9": [0]}, ["r9"], cp) == [PAD_SLOT_ID]
    a.prepare({"r4": [0]}, [], cp)
    with pytest.raises(RuntimeError, match="no free state slot"):
        a.prepare({"r5": [0]}, [], cp)
    assert a.cache_indices_mapping["r1"] == {1: s0[1]}


@pytest.fixture
def fake_state_ops(monkeypatch):
    """The cache managers have no CPU branch (the product has no CPU fallback): for the host-logic
    tests the two device kernels they launch are replaced by torch stand-ins with the same
    contract (state [L, slots, ...]; slots < 0 skipped)."""
    from vllm_hymeta_b200 import ops

    def copy(state, src, dst):
        state[:, dst].copy_(state[:, src])

    def zero(state, slots):
        keep = slots[slots >= 0].long()
        state[:, keep] = 0

    monkeypatch.setattr(ops, "state_copy_slot", copy)
    monkeypatch.setattr(ops, "state_zero_slots", zero)


def test_cache_managers_cpu_tensors(fake_state_ops):
    L, S, Hs, D, M, Hkv = 2, 3, 2, 8, 4, 1
    m = HymetaCacheManager(torch.bfloat16, (L, S, Hs, D, D), (L, Hs, D, D), (L, 2, M, Hkv, D), device="cpu")
    state, meta_lin, meta_fa, idx = m.current_run_tensors(request_ids_to_seq_ids={"a": [0], "b": [1]},
                                                          finished_requests_ids=[])
    assert state.shape == (L, S, Hs, D, D) and state.abs().max() == 0          # zero-initialised (Q7)
    assert meta_lin.shape == (L, Hs, D, D) and meta_fa.shape == (L, 2, M, Hkv, D)
    assert idx.dtype == torch.int32 and idx.numel() == 2
    state[:, idx[0].item()] = 1.0
    m._copy_cache(idx[0].item(), idx[1].item())
    assert torch.equal(state[:, idx[1].item()], state[:, idx[0].item()])
    m.clear_slots([idx[0].item()])
    assert state[:, idx[0].item()].abs().max() == 0 and state[:, idx[1].item()].min() == 1
    p = HymetaCacheParams(state, meta_lin, meta_fa, idx).at_layer_idx(1)
    assert p.hymeta_cache.shape == (S, Hs, D, D) and p.meta_fattn_cache.shape == (2, M, Hkv, D)
    # CUDA-graph protocol: capture buffers get PAD for rows beyond the live batch
    cap = m.get_seqlen_agnostic_capture_inputs(4)
    assert cap[1].tolist() == [PAD_SLOT_ID] * 4
    m.copy_inputs_before_cuda_graphs({"seqlen_agnostic_capture_inputs": cap},
                                     request_ids_to_seq_ids={"a": [0], "b": [1]}, finished_requests_ids=[])
    assert cap[1].tolist() == idx.tolist() + [PAD_SLOT_ID] * 2
    g = GLACacheManager(torch.bfloat16, (L, S, Hs, D, D), device="cpu")
    cache, gi = g.current_run_tensors(request_ids_to_seq_ids={"x": [0]}, finished_requests_ids=[])
    cache[:, gi[0].item()] = 2.0
    g._copy_cache(gi[0].item(), (gi[0].item() + 1) % S)           # copies SLOTS
    assert cache[:, (gi[0].item() + 1) % S].min() == 2


def test_metadata_builder_mixed_batch_bit_exact():
    bs = 16
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device="cpu")
    # two prompts (one chunk-continued with context 16) then two decodes
    b.add(SeqData(True, 20, 20, 0, [3, 9]))
    b.add(SeqData(True, 40, 24, 16, [5, 1, 7]))
    b.add(SeqData(False, 33, 1, 32, [2, 4, 6]))
    b.add(SeqData(False, 17, 1, 16, [8, 0]))
    md = b.build()
    assert (md.num_prefills, md.num_prefill_tokens, md.num_decode_tokens) == (2, 44, 2)
    exp = [3 * bs + i for i in range(16)] + [9 * bs + i for i in range(4)]
    exp += [1 * bs + i for i in range(16)] + [7 * bs + i for i in range(8)]
    exp += [6 * bs + 0, 0 * bs + 0]
    assert md.slot_mapping.tolist() == exp and md.slot_mapping.dtype == torch.int64
    assert md.query_start_loc.tolist() == [0, 20, 44, 45, 46]
    assert md.seq_start_loc.tolist() == [0, 20, 60, 93, 110]
    assert md.context_lens_tensor.tolist() == [0, 16, 32, 16]
    assert md.block_tables.dtype == torch.int32
    assert md.block_tables.tolist() == [[0, 0, 0], [0, 0, 0], [2, 4, 6], [8, 0, 0]]   # prompts: no cache read
    pm, dm = md.prefill_metadata, md.decode_metadata
    assert pm.seq_lens == [20, 40] and pm.slot_mapping.numel() == 44 and pm.max_prefill_seq_len == 40
    assert dm.seq_lens_tensor.tolist() == [33, 17] and dm.query_start_loc.tolist() == [0, 1, 2]
    assert dm.block_tables.tolist() == [[2, 4, 6], [8, 0, 0]] and dm.max_decode_seq_len == 33


def test_metadata_builder_window_graph_and_profile():
    bs = 16
    # sliding window 32: prompt tokens that fall out of the window are never cached
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=32, device="cpu")
    b.add(SeqData(True, 40, 40, 0, [1, 2, 3]))
    md = b.build()
    assert md.slot_mapping.tolist() == [PAD_SLOT_ID] * 8 + [1 * bs + i for i in range(8, 16)] + \
        [2 * bs + i for i in range(16)] + [3 * bs + i for i in range(8)]
    # decode with a windowed block table keeps the last blocks only
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=32, device="cpu")
    b.add(SeqData(False, 32, 1, 49, [1, 2, 3, 4], curr_sliding_window_block=2, orig_seq_len=50))
    assert b.block_tables == [[3, 4]]
    # CUDA-graph build pads slots with PAD and uses the runner's fixed-width table
    import numpy as np
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device="cpu",
                                     graph_block_tables=np.zeros((4, 5), dtype=np.int32))
    b.add(SeqData(False, 18, 1, 17, [7, 8]))
    md = b.build(cuda_graph_pad_size=3, batch_size=4)
    assert md.use_cuda_graph and md.slot_mapping.tolist() == [8 * bs + 1] + [PAD_SLOT_ID] * 3
    assert md.block_tables.shape == (1, 5) and md.block_tables[0].tolist() == [7, 8, 0, 0, 0]
    assert md.num_decode_tokens == 4
    # profiling run: no block table -> every slot is PAD
    b = MetaAttentionMetadataBuilder(block_size=bs, sliding_window=None, device="cpu")
    b.add(SeqData(True, 6, 6, 0, None))
    assert b.build().slot_mapping.tolist() == [PAD_SLOT_ID] * 6


def test_backend_surface_and_errors():
    B = MetaAttentionBackend
    assert B.get_name() == "META_FLASH_ATTN" and B.accept_output_buffer
    assert B.get_kv_cache_shape(10, 256, 4, 128) == (2, 10, 256, 4, 128)
    for bad in (8, 24, 250):
        with pytest.raises(ValueError):
            B.get_kv_cache_shape(10, bad, 4, 128)
    assert B.get_impl_cls() is MetaAttentionImpl
    kw = dict(num_heads=28, head_size=128, scale=128 ** -0.5, num_kv_heads=4, alibi_slopes=None,
              sliding_window=4096, kv_cache_dtype="auto")
    impl = MetaAttentionImpl(**kw)
    assert impl.sliding_window == (4095, 0) and impl.num_queries_per_kv == 7 and impl.num_meta_tokens == 128
    assert MetaAttentionImpl(**{**kw, "sliding_window": None}).sliding_window == (-1, -1)
    with pytest.raises(ValueError):
        MetaAttentionImpl(**{**kw, "head_size": 80})
    with pytest.raises(ValueError):
        MetaAttentionImpl(**kw, blocksparse_params={"x": 1})
    with pytest.raises(NotImplementedError):
        MetaAttentionImpl(**kw, kv_sharing_target_layer_name="l0")
    with pytest.raises(NotImplementedError):
        MetaAttentionImpl(**{**kw, "kv_cache_dtype": "fp8"})
    # swap / copy blocks are plain tensor moves
    kv = torch.arange(2 * 4 * 2 * 1 * 2, dtype=torch.float32).view(2, 4, 2, 1, 2)
    dst = torch.zeros_like(kv)
    B.swap_blocks(kv, dst, torch.tensor([[1, 3]]))
    assert torch.equal(dst[:, 3], kv[:, 1]) and dst[:, 1].abs().max() == 0
    B.copy_blocks([kv], torch.tensor([[0, 2]]))
    assert torch.equal(kv[:, 2], kv[:, 0])


def test_plugin_entry_points_do_not_touch_cuda():
    import vllm_hymeta_b200 as pkg
    assert pkg.register() == "vllm_hymeta_b200.platform.HymetaCudaPlatform"
    pkg.register_model()                      # no reference plugin importable here: a no-op
    from vllm_hymeta_b200.platform import HymetaCudaPlatform as P
    assert P.get_attn_backend_cls("XFORMERS", 128, None, None, 256, False, False).endswith("FlashAttentionBackend")
    assert P.get_attn_backend_cls(None, 128, None, None, 256, False, False) == \
        "vllm_hymeta_b200.attention.meta_attention.MetaAttentionBackend"
    assert not torch.cuda.is_initialized()


# ------------------------------------------------------------------------------------- f4 (native slot manager)
def test_native_slot_manager_matches_python_spec_on_random_traces():
    """The C++ manager behind the C ABI (csrc/slot_mgr.cu) against the Python restatement of
    ConstantSizeCache on random admit / fork / finish traces: same slots, same fork copies in the
    same order, same free list, same live map -- bit-exact."""
    import random
    from vllm_hymeta_b200.slot_cache import NativeSlotAllocator
    rng = random.Random(0)
    for trial in range(20):
        cap = rng.choice([1, 4, 17, 64])
        py, nat = SlotAllocator(cap), NativeSlotAllocator(cap)
        live, next_id = {}, 0
        for step in range(60):
            finished = [r for r in list(live) if rng.random() < 0.2]
            for r in finished:
                live.pop(r)
            if rng.random() < 0.15:
                finished.append(f"ghost-{step}")              # finished but never seen
            batch = {r: list(s) for r, s in live.items() if rng.random() < 0.8}
            for _ in range(rng.randint(0, 3)):                # admit
                batch[f"req-{next_id}"] = [next_id * 10]
                next_id += 1
            for r in list(batch):                             # fork siblings
                if rng.random() < 0.15:
                    batch[r].append(batch[r][-1] + 1)
            if rng.random() < 0.1 and batch:                  # finished before it ever ran
                finished.append(next(iter(batch)))
            pc, nc_ = [], []
            err_py = err_nat = None
            try:
                sp = py.prepare(batch, finished, lambda a, b: pc.append((a, b)))
            except RuntimeError as e:
                err_py = str(e)
            try:
                sn = nat.prepare(batch, finished, lambda a, b: nc_.append((a, b)))
            except RuntimeError as e:
                err_nat = str(e)
            assert (err_py is None) == (err_nat is None), (trial, step, err_py, err_nat)
            assert pc == nc_
            assert py.free_cache_indices == nat.free_cache_indices
            assert py.cache_indices_mapping == nat.cache_indices_mapping
            assert sorted(py.live()) == sorted(nat.live())
            if err_py is not None:
                break
            assert sp == sn
            for r, seqs in batch.items():
                if r not in finished:
                    live[r] = seqs
            live = {r: s for r, s in live.items() if r in py.cache_indices_mapping}


def test_native_slot_manager_batch_outputs_and_errors():
    from vllm_hymeta_b200.slot_cache import NativeSlotAllocator
    a = NativeSlotAllocator(3)
    slots, copies, new = a.prepare_batch({"a": [0], "b": [1]}, [])
    assert slots == [2, 1] and copies == [] and new == [2, 1]
    slots, copies, new = a.prepare_batch({"a": [0, 5], "b": [1]}, [])
    assert slots == [2, 0, 1] and copies == [(2, 0)] and new == []
    assert a.slot_of("a", 5) == 0 and a.free_cache_indices == []
    with pytest.raises(RuntimeError):
        a.prepare_batch({"c": [9]}, [])
    with pytest.raises(KeyError):
        a.slot_of("zz", 0)
    slots, _, new = a.prepare_batch({"c": [9]}, ["a", "c2"])
    assert slots == [0] and new == [0] and a.free_cache_indices == [2]   # LIFO: a's 2 then 0 freed
    a.release(["b", "c"])
    assert sorted(a.free_cache_indices) == [0, 1, 2] and a.live() == []
    assert a.prepare_batch({}, []) == ([], [], [])
    assert a.prepare_batch({"x": [1]}, ["x"])[0] == [PAD_SLOT_ID]


# ------------------------------------------------------------------------------------- b (packaging)
def test_vllm_plugin_loader_discovers_the_package(tmp_path):
    """pyproject.toml declares the reference's two entry-point groups (setup.py:69-76).  The package
    is pip-installed from the source tree into a scratch directory and the INSTALLED vLLM's own
    plugin loader (vllm/plugins/__init__.py) must find and load both entry points."""
    import os
    import subprocess
    import sys
    pytest.importorskip("vllm")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    target = tmp_path / "site"
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
                        "--no-deps", "--quiet", "--target", str(target), root],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    code = (
        "from vllm.plugins import load_plugins_by_group\n"
        "p = load_plugins_by_group('vllm.platform_plugins')\n"
        "g = load_plugins_by_group('vllm.general_plugins')\n"
        "assert p['hymeta_b200']() == 'vllm_hymeta_b200.platform.HymetaCudaPlatform', p\n"
        "g['hymeta_b200_model']()\n"
        "from vllm.v1.attention.backends.registry import AttentionBackendEnum as E\n"
        "print('BACKEND', E.CUSTOM.get_path())\n"
        "import vllm_hymeta_b200.platform as pl\n"
        "print('PLATFORM', pl.HymetaCudaPlatform.__mro__[1].__name__)\n")
    env = dict(os.environ, PYTHONPATH=str(target), VLLM_LOGGING_LEVEL="ERROR")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env,
                       cwd=str(tmp_path))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-3000:]
    assert "BACKEND vllm_hymeta_b200.attention.meta_attention_v1.MetaAttentionV1Backend" in r.stdout
    # binds to vLLM's NVML platform (reference platform.py:7) where vllm.platforms.cuda imports -- it
    # needs libcuda, so on a GPU-less host the class is a plain object (checked on the GPU box by
    # tests/test_gpu_modules.py::test_platform_binds_to_vllm_nvml_platform)
    assert "PLATFORM NvmlCudaPlatform" in r.stdout or "PLATFORM object" in r.stdout


def test_platform_selection_rules():
    from types import SimpleNamespace
    from vllm_hymeta_b200.platform import HymetaCudaPlatform as P
    # V0 protocol (reference platform.py:20-30)
    assert P.get_attn_backend_cls(None, 128, torch.bfloat16, "auto", 256, False, False).endswith(
        "meta_attention.MetaAttentionBackend")
    assert P.get_attn_backend_cls("XFORMERS", 128, torch.bfloat16, "auto", 256, False, False).endswith(
        "flash_attn.FlashAttentionBackend")
    # V1 protocol of the installed vLLM: (selected_backend, attn_selector_config, num_heads)
    cfg = SimpleNamespace(use_mla=False, head_size=128)
    assert P.get_attn_backend_cls(None, cfg, 28).endswith("meta_attention_v1.MetaAttentionV1Backend")


def test_unified_meta_attention_custom_ops_route_through_forward_context():
    """``torch.ops.vllm.unified_meta_attention_with_output`` (reference models/attention.py:361-409):
    looks the layer up by name in the forward context and calls ``impl.forward`` with the context's
    metadata, the layer's kv cache and the query2/key2/value2 tensors; the fakes accept the same
    arguments."""
    from vllm_hymeta_b200.models import attention as A
    layer = A.MetaAttention(4, 128, 128 ** -0.5, 2, per_layer_sliding_window=64, num_meta_tokens=8,
                            prefix="test.layers.0.attn", use_direct_call=False)
    with pytest.raises(ValueError, match="Duplicate layer name"):
        A.MetaAttention(4, 128, 128 ** -0.5, 2, prefix="test.layers.0.attn")
    seen = {}

    def fake_forward(lyr, q, k, v, q2, k2, v2, kv_cache, md, output=None):
        seen.update(layer=lyr, md=md, kv=kv_cache, shapes=(q.shape, k.shape, q2.shape, k2.shape))
        output.fill_(3.0)

    layer.impl.forward = fake_forward
    layer.bind_kv_cache([torch.zeros(1), torch.ones(1)])         # one cache per virtual engine (:170-176)
    T, M = 5, 8
    q, k, v = torch.zeros(T, 4 * 128), torch.zeros(T, 2 * 128), torch.zeros(T, 2 * 128)
    q2, k2, v2 = torch.zeros(M, 4 * 128), torch.zeros(M, 2, 128), torch.zeros(M, 2, 128)
    with A.set_forward_context("MD", virtual_engine=1):
        out = layer.forward(q, k, v, query2=q2, key2=k2, value2=v2)
    assert out.shape == (T + M, 4 * 128) and out.min() == 3.0
    assert seen["layer"] is layer and seen["md"] == "MD" and seen["kv"].item() == 1.0
    assert seen["shapes"] == ((T, 4, 128), (T, 2, 128), (M, 4, 128), (M, 2, 128))
    # fake (meta-device) evaluation with the meta-token arguments present
    mq, mk = torch.empty(T, 4, 128, device="meta"), torch.empty(T, 2, 128, device="meta")
    mq2, mk2 = torch.empty(M, 4, 128, device="meta"), torch.empty(M, 2, 128, device="meta")
    o = torch.ops.vllm.unified_meta_attention(mq, mk, mk, "test.layers.0.attn", mq2, mk2, mk2)
    assert o.shape == (T + M, 4, 128)
    torch.ops.vllm.unified_meta_attention_with_output(mq, mk, mk, torch.empty(T + M, 4, 128, device="meta"),
                                                     "test.layers.0.attn", mq2, mk2, mk2)
    del A._LAYERS["test.layers.0.attn"]


# ------------------------------------------------------------------------------------- f4 (V1 adapter)
def test_v1_metadata_builder_bit_exact():
    """``CommonAttentionMetadata`` -> ``MetaAttentionV1Metadata``: the decode / prefill split and the
    rebased offsets are exact integer work."""
    from types import SimpleNamespace
    from vllm_hymeta_b200.attention.meta_attention_v1 import (MetaAttentionV1Backend,
                                                              MetaAttentionV1Impl,
                                                              MetaAttentionV1MetadataBuilder)
    qsl = torch.tensor([0, 1, 2, 3, 13, 41, 42], dtype=torch.int32)      # 3 decodes, 10, 28, then a 1-token chunk
    seq = torch.tensor([5, 9, 300, 10, 60, 7], dtype=torch.int32)
    cm = SimpleNamespace(query_start_loc=qsl, query_start_loc_cpu=qsl, seq_lens=seq, num_reqs=6,
                         num_actual_tokens=42, max_query_len=28, max_seq_len=300,
                         block_table_tensor=torch.arange(6 * 20, dtype=torch.int32).view(6, 20),
                         slot_mapping=torch.arange(50), causal=True)            # slot_mapping padded
    b = MetaAttentionV1MetadataBuilder(SimpleNamespace(block_size=16), ["l"], None, "cpu")
    md = b.build(0, cm)
    assert (md.num_decodes, md.num_prefills, md.num_prefill_tokens, md.num_actual_tokens) == (3, 3, 39, 42)
    assert md.prefill_query_start_loc.tolist() == [0, 10, 38, 39]
    assert md.prefill_seq_start_loc.tolist() == [0, 10, 70, 77]
    assert md.max_prefill_query_len == 28 and md.slot_mapping.numel() == 42
    assert md.block_table.shape == (6, 20) and md.seq_lens.dtype == torch.int32
    cm.query_start_loc = cm.query_start_loc_cpu = torch.tensor([0, 1, 2], dtype=torch.int32)
    cm.seq_lens, cm.num_reqs, cm.num_actual_tokens, cm.max_query_len = seq[:2], 2, 2, 1
    md = b.build(0, cm)
    assert md.num_decodes == 2 and md.num_prefills == 0 and md.prefill_query_start_loc is None
    assert MetaAttentionV1Backend.get_kv_cache_shape(7, 256, 4, 128) == (2, 7, 256, 4, 128)
    with pytest.raises(ValueError):
        MetaAttentionV1Backend.get_kv_cache_shape(7, 24, 4, 128)
    with pytest.raises(NotImplementedError):
        MetaAttentionV1Impl(28, 128, 0.1, 4, alibi_slopes=[1.0])
    with pytest.raises(ValueError):
        MetaAttentionV1Impl(28, 64, 0.1, 4)
    assert MetaAttentionV1Impl(28, 128, 0.1, 4, sliding_window=4096).window_left == 4095


def test_gla_prefill_scheduling_policy_is_pure_host_logic():
    """Which GLA prefill form a launch takes is decided on the host from (chains, longest sequence,
    SM count) alone -- no device query, no sync: balanced work lists only where a partial last wave
    costs more than the per-piece pipeline restart."""
    from vllm_hymeta_b200 import ops
    assert ops.gla_prefill_balanced_pays(256, 8192, 148)        # BASELINE config 2: 1.73 waves of 512 steps
    assert ops.gla_prefill_balanced_pays(160, 8192, 148)
    assert ops.gla_prefill_balanced_pays(400, 4096, 148)
    assert not ops.gla_prefill_balanced_pays(148, 8192, 148)    # exactly one wave
    assert not ops.gla_prefill_balanced_pays(8, 32768, 148)     # few long chains: the segmented form's case
    assert not ops.gla_prefill_balanced_pays(1024, 1024, 148)   # 6.9 waves of short chains
    assert not ops.gla_prefill_balanced_pays(4000, 256, 148)
    assert not ops.gla_prefill_balanced_pays(256, 0, 148)
