"""Generate tests/golden/*.npz from the REFERENCE's own Python.  TEST INFRASTRUCTURE.

Runs only in the authoring container (needs /root/reference, read-only, and the
installed flash-linear-attention for its ``naive_recurrent_gla``).  The GPU box
never runs this; it only reads the committed fixtures.

    python oracle/make_golden.py            # rewrites tests/golden/

Reference code is imported unmodified; only third-party GPU-only symbols are
monkeypatched in the importing module's namespace:
  * chunk_gla / fused_chunk_gla / fused_recurrent_gla -> fla's own
    ``fla.ops.gla.naive.naive_recurrent_gla`` (with the head-first <-> seq-first
    transpose the reference's fla==0.1 API implies),
  * fla.modules.RMSNorm (Triton) -> the reference's torch hf_7B_model/norms.py,
  * ``Tensor.cuda`` -> identity, so ``naive_metatoken_flash_attn`` runs on CPU.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

REF = os.environ.get("HYMETA_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def f32(x):
    return x.detach().float().numpy()


def naive_gla_headfirst(q, k, v, gk, initial_state=None, output_final_state=False, **kw):
    from fla.ops.gla.naive import naive_recurrent_gla
    tr = lambda x: x.transpose(1, 2)
    o, h = naive_recurrent_gla(tr(q), tr(k), tr(v), tr(gk), initial_state=initial_state,
                               output_final_state=output_final_state)
    return tr(o), h


def golden_gla_naive():
    """fla's naive recurrence on both gate parameterisations, with initial state."""
    from fla.ops.gla.naive import naive_recurrent_gla
    from oracle.gla import synth_gla_inputs
    out = {}
    for tag, model in (("7b", "7b"), ("70b", "70b")):
        T, H, Hkv, D = 96, 4, 2, 128
        q, k, v, g = synth_gla_inputs(T, H, Hkv, D, model, seed=11)
        gen = torch.Generator().manual_seed(5)
        S0 = (torch.randn(H, D, D, generator=gen) * 0.1)
        rep = lambda x: x.repeat_interleave(H // Hkv, dim=1)
        o, S = naive_recurrent_gla(q[None].float(), rep(k)[None].float(), rep(v)[None].float(),
                                   rep(g)[None].float(), initial_state=S0[None],
                                   output_final_state=True)
        out.update({f"{tag}_q": f32(q), f"{tag}_k": f32(k), f"{tag}_v": f32(v), f"{tag}_g": f32(g),
                    f"{tag}_S0": f32(S0), f"{tag}_o": f32(o[0]), f"{tag}_S": f32(S[0])})
    np.savez_compressed(os.path.join(OUT, "gla_naive.npz"), **out)


def golden_swa_naive():
    """The reference's dense definition ``naive_metatoken_flash_attn``."""
    import vllm_hymeta.attention.fa_kernel_metatoken as fk
    orig_cuda = torch.Tensor.cuda
    torch.Tensor.cuda = lambda self, *a, **k: self
    try:
        out = {}
        gen = torch.Generator().manual_seed(3)
        H, D, M = 4, 64, 16
        cases = {"prefill": (40, 40, 12, True), "suffix": (9, 40, 12, False),
                 "decode": (1, 33, 8, False), "full": (24, 24, 1 << 30, True)}
        for tag, (Lq, Lk, W, with_q2) in cases.items():
            q1 = torch.randn(1, Lq, H, D, generator=gen)
            k1 = torch.randn(1, Lk, H, D, generator=gen)
            v1 = torch.randn(1, Lk, H, D, generator=gen)
            q2 = torch.randn(1, M, H, D, generator=gen) if with_q2 else None
            k2 = torch.randn(1, M, H, D, generator=gen)
            v2 = torch.randn(1, M, H, D, generator=gen)
            o = fk.naive_metatoken_flash_attn(q1, q2, k1, k2, v1, v2, softmax_scale=D ** -0.5,
                                              num_meta_tokens=M, sliding_window=(W, 0))
            out.update({f"{tag}_q1": f32(q1[0]), f"{tag}_k1": f32(k1[0]), f"{tag}_v1": f32(v1[0]),
                        f"{tag}_k2": f32(k2[0]), f"{tag}_v2": f32(v2[0]), f"{tag}_o": f32(o[0]),
                        f"{tag}_W": np.int64(W)})
            if with_q2:
                out[f"{tag}_q2"] = f32(q2[0])
        np.savez_compressed(os.path.join(OUT, "swa_naive.npz"), **out)
    finally:
        torch.Tensor.cuda = orig_cuda


def golden_merge():
    """The reference's TorchScript ``_update_out_and_lse``."""
    import vllm_hymeta.attention.fa_kernel_metatoken as fk
    gen = torch.Generator().manual_seed(9)
    B, L, H, D = 2, 7, 3, 32
    out = torch.randn(B, L, H, D, generator=gen)
    bout = torch.randn(B, L, H, D, generator=gen)
    lse = torch.randn(B, H, L, generator=gen) * 3
    blse = torch.randn(B, H, L, generator=gen) * 3
    o, l = fk._update_out_and_lse(out, lse, bout, blse)
    np.savez_compressed(os.path.join(OUT, "merge.npz"), out=f32(out), bout=f32(bout),
                        lse=f32(lse), blse=f32(blse), o=f32(o), l=f32(l))


def golden_hf7b_gla_module():
    """hf_7B_model.GatedLinearAttention (unmodified) with shimmed kernels: pins the
    7B gate pre-processing, repeat_kv and the post-kernel per-head g_norm."""
    import hf_7B_model.gla_attention as m
    import hf_7B_model.norms as norms
    cap = {}

    def shim(q, k, v, gk, initial_state=None, output_final_state=False, **kw):
        cap.update(q=q, k=k, v=v, gk=gk)
        o, h = naive_gla_headfirst(q, k, v, gk, initial_state, output_final_state)
        cap.update(o=o, h=h)
        return o, h

    m.chunk_gla = m.fused_chunk_gla = m.fused_recurrent_gla = shim
    m.RMSNorm = norms.RMSNorm
    torch.manual_seed(21)
    mod = m.GatedLinearAttention(mode="chunk", hidden_size=512, num_heads=4,
                                 num_key_value_heads=2, layer_idx=0, norm_eps=1e-5).eval()
    with torch.no_grad():
        mod.g_norm.weight.copy_(torch.rand(128) + 0.5)
        for lin in (mod.q_proj, mod.k_proj, mod.v_proj, mod.gk_proj[1]):
            lin.bias.normal_(0, 0.5)
        mod.gk_proj[0].weight.normal_(0, 0.3)
        mod.gk_proj[1].weight.normal_(0, 0.3)
        x = torch.randn(1, 40, 512) * 2
        y, _, _ = mod(x, use_cache=True)
        out = dict(q_raw=f32(mod.q_proj(x)[0]), k_raw=f32(mod.k_proj(x)[0]),
                   v_raw=f32(mod.v_proj(x)[0]), gk_raw=f32(mod.gk_proj(x)[0]),
                   q=f32(cap["q"][0]), k=f32(cap["k"][0]), v=f32(cap["v"][0]), gk=f32(cap["gk"][0]),
                   o=f32(cap["o"][0]), h=f32(cap["h"][0]), g_norm_w=f32(mod.g_norm.weight),
                   o_proj_w=f32(mod.o_proj.weight), y=f32(y[0]))
    np.savez_compressed(os.path.join(OUT, "hf7b_gla_module.npz"), **out)


def golden_hf70b_linear_module():
    """hf_model.HLinearAttention (unmodified) with shimmed kernels: pins the 70B
    feature map, clamp and lower-bound gate."""
    import hf_model.linear_attention as m
    out = {}
    for tag, layer_idx, use_lb in (("first", 0, True), ("lb", 3, True), ("nolb", 2, False)):
        cap = {}

        def shim(q, k, v, g, initial_state=None, output_final_state=False, **kw):
            cap.update(q=q, k=k, v=v, g=g)
            o, h = naive_gla_headfirst(q, k, v, g, initial_state, output_final_state)
            cap.update(o=o, h=h)
            return o, h

        m.chunk_gla = m.fused_chunk_gla = m.fused_recurrent_gla = shim
        torch.manual_seed(31 + layer_idx)
        import inspect
        kw = dict(hidden_size=512, num_heads=4, num_key_value_heads=2, layer_idx=layer_idx)
        sig = inspect.signature(m.HLinearAttention.__init__).parameters
        if "mode" in sig:
            kw["mode"] = "chunk"
        if "num_meta_tokens" in sig:
            kw["num_meta_tokens"] = 0
        mod = m.HLinearAttention(**kw).eval()
        with torch.no_grad():
            for lin in (mod.q_proj, mod.k_proj, mod.v_proj):
                if lin.bias is not None:
                    lin.bias.normal_(0, 0.5)
            x = torch.randn(1, 24, 512) * 3
            lb = None
            if use_lb:
                p = torch.randn(6, 512)
                lbs = p.softmax(0)
                lbs = lbs.cumsum(0) - lbs[0]
                lb = lbs[layer_idx]
            y, _, _ = mod(x, lower_bound=lb, use_cache=True)
            out.update({f"{tag}_q_raw": f32(mod.q_proj(x)[0]), f"{tag}_k_raw": f32(mod.k_proj(x)[0]),
                        f"{tag}_v_raw": f32(mod.v_proj(x)[0]),
                        f"{tag}_q": f32(cap["q"][0]), f"{tag}_k": f32(cap["k"][0]),
                        f"{tag}_g": f32(cap["g"][0]), f"{tag}_o": f32(cap["o"][0]),
                        f"{tag}_layer": np.int64(layer_idx)})
            if lb is not None:
                out[f"{tag}_lb"] = f32(lb)
    np.savez_compressed(os.path.join(OUT, "hf70b_linear_module.npz"), **out)


def golden_rope():
    """hf_7B_model/window_attention.py rotary embedding (unmodified)."""
    import hf_7B_model.window_attention as w
    torch.manual_seed(2)
    rot = w.RotaryEmbedding(128, max_position_embeddings=131072, base=1e6)
    q = torch.randn(1, 4, 20, 128)
    k = torch.randn(1, 2, 20, 128)
    pos = (torch.arange(20) * 997 + 128)[None]
    cos, sin = rot(q, pos)
    qe, ke = w.apply_rotary_pos_emb(q, k, cos, sin)
    np.savez_compressed(os.path.join(OUT, "rope.npz"), q=f32(q[0]), k=f32(k[0]),
                        pos=pos[0].numpy(), qe=f32(qe[0]), ke=f32(ke[0]))


def _import_reference_model_module():
    """``vllm_hymeta.models.modeling_hymeta`` imported UNMODIFIED.  It is written against vLLM
    0.10.0; the vLLM installed here (0.22) has moved or dropped most of what it imports, so every
    ``vllm.*`` / ``fla.*`` name it asks for that does not resolve is satisfied by a stub (third-party
    symbols only: ``CustomOp`` becomes ``torch.nn.Module``, everything else an inert placeholder).
    The reference's own classes and arithmetic run as written."""
    import importlib
    import importlib.abc
    import importlib.machinery
    import types

    class _AnyMeta(type):
        def __getattr__(cls, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return name

    class _AnyObj:
        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _AnyObj()

        def __call__(self, *a, **k):
            return None

    class _Stub(types.ModuleType):
        __path__ = []

        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            if name == "CustomOp":
                obj = torch.nn.Module
            elif name[:1].islower():      # functions / singletons (direct_register_custom_op, current_platform): inert
                obj = _AnyObj()
            else:                         # classes / enums: any class attribute resolves, any call constructs
                obj = _AnyMeta(name, (), {"__init__": lambda self, *a, **k: None})
            setattr(self, name, obj)
            return obj

    class _Finder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
        def find_spec(self, fullname, path=None, target=None):
            if fullname.split(".")[0] in ("vllm", "fla", "flash_attn"):
                return importlib.machinery.ModuleSpec(fullname, self, is_package=True)
            return None

        def create_module(self, spec):
            return _Stub(spec.name)

        def exec_module(self, module):
            pass

    saved = {k: v for k, v in sys.modules.items() if k.split(".")[0] in ("vllm", "fla", "flash_attn")}
    for k in saved:
        del sys.modules[k]
    finder = _Finder()
    sys.meta_path.insert(0, finder)     # ahead of the real vllm: the module only needs names to exist
    try:
        return importlib.import_module("vllm_hymeta.models.modeling_hymeta")
    finally:
        sys.meta_path.remove(finder)
        for k in [k for k in sys.modules if k.split(".")[0] in ("vllm", "fla", "flash_attn")]:
            del sys.modules[k]
        sys.modules.update(saved)


def golden_rmsnorm_tp():
    """``HymetaRMSNormTP`` (modeling_hymeta.py:120-180, unmodified) at tensor-parallel world sizes
    1, 2 and 4 -- the ranks run as threads and ``tensor_model_parallel_all_reduce`` is a barrier
    that sums what the ranks deposit -- followed by the decoder layer's merge of the two branches
    ``((norm1(out_attn) + norm2(out_linear)) / 2).to(bfloat16)`` (:909-917, restated here in one
    line because it sits in the middle of the layer's forward).  Pins oracle.gla.rms_norm_tp /
    branch_norm_merge, the checker of f1."""
    import threading
    mh = _import_reference_model_module()
    gen = torch.Generator().manual_seed(17)
    T, hidden, eps = 37, 512, 1e-6
    x1 = (torch.randn(T, hidden, generator=gen) * 1.7).bfloat16()
    x2 = (torch.randn(T, hidden, generator=gen) * 0.6).bfloat16()
    w1 = 1 + 0.2 * torch.randn(hidden, generator=gen)
    w2 = 1 + 0.2 * torch.randn(hidden, generator=gen)
    out = {"x_attn": f32(x1), "x_linear": f32(x2), "w1": f32(w1), "w2": f32(w2), "eps": np.float32(eps)}
    for W in (1, 2, 4):
        barrier = threading.Barrier(W)
        box, lock = {}, threading.Lock()

        def all_reduce(v):
            tag = all_reduce.calls[threading.get_ident()] = all_reduce.calls.get(threading.get_ident(), 0) + 1
            with lock:
                box.setdefault(tag, []).append(v)
            barrier.wait()
            tot = torch.stack(box[tag]).sum(0)
            barrier.wait()
            return tot
        all_reduce.calls = {}
        mh.tensor_model_parallel_all_reduce = all_reduce
        mh.get_tensor_model_parallel_world_size = lambda: W
        res = [None] * W

        def run(rank):
            mh.get_tensor_model_parallel_rank = lambda: rank
            n1, n2 = mh.HymetaRMSNormTP(hidden, eps), mh.HymetaRMSNormTP(hidden, eps)
            mh.HymetaRMSNormTP.weight_loader(n1.weight, w1)      # slices by the patched rank getter
            mh.HymetaRMSNormTP.weight_loader(n2.weight, w2)
            sl = slice(rank * hidden // W, (rank + 1) * hidden // W)
            res[rank] = (n1, n2, sl)
        for r in range(W):          # construction reads the rank getter: one rank at a time
            run(r)

        def fwd(rank):
            n1, n2, sl = res[rank]
            a = n1._forward(x1[:, sl])
            b = n2._forward(x2[:, sl])
            res[rank] = ((a + b) / 2).to(torch.bfloat16)
        th = [threading.Thread(target=fwd, args=(r,)) for r in range(W)]
        [t.start() for t in th]
        [t.join() for t in th]
        out[f"merged_tp{W}"] = f32(torch.cat(res, dim=1))
    np.savez_compressed(os.path.join(OUT, "hymeta_rmsnorm_tp.npz"), **out)


def main():
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    os.makedirs(OUT, exist_ok=True)
    torch.set_grad_enabled(False)
    golden_gla_naive()
    golden_swa_naive()
    golden_merge()
    golden_hf7b_gla_module()
    golden_hf70b_linear_module()
    golden_rope()
    golden_rmsnorm_tp()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
