"""Per-kernel micro-benchmarks at the BASELINE config shapes (CUDA events, L2 flushed between
iterations).  Development aid; bench.py is the contract benchmark."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_hymeta_b200 import ops  # noqa: E402

PEAK, TPEAK = 6524.6, 1401.9
try:
    _pk = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
    PEAK, TPEAK = _pk["hbm_gbs"], _pk["bf16_tflops_sustained"]
except Exception:
    pass


def timeit(fn, iters=20, warmup=3, flush=None):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


def time_graph(fns, reps=5):
    """Launch-overhead-free timing of short kernels: the calls in `fns` (each on its OWN buffers,
    together larger than L2 so no call finds its data cached) are captured into one CUDA graph;
    returns the median replay time divided by len(fns)."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for f in fns:
                f()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps + 2):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        g.replay()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts = sorted(ts[2:])
    return ts[len(ts) // 2] * 1e-3 / len(fns)


def bench_gla_decode(B, H=28, Hkv=4, D=128, layout="compact", gdt=torch.bfloat16, flush=None, mode="plain"):
    """mode: plain | gn (g_norm epilogue) | gate (g_norm + the 7B feature map / gate on raw rows)"""
    dev = "cuda"
    Hs = Hkv if layout == "compact" else H
    Hk = Hkv if layout == "compact" else H
    gb = torch.empty(0, dtype=gdt).element_size()
    alg = B * (Hs * 2 * D * D * 2 + H * 2 * D * 2 + Hk * D * (4 + gb))
    nrot = max(2, min(64, int(600e6 // (alg / 2)) + 1))      # rotate over > 600 MB of state
    fns = []
    for _ in range(nrot):
        q = torch.randn(B, H, D, device=dev).bfloat16().relu()
        k = torch.randn(B, Hk, D, device=dev).bfloat16().relu()
        v = torch.randn(B, Hk, D, device=dev).bfloat16()
        g = (-torch.rand(B, Hk, D, device=dev) * 0.1).to(gdt)
        state = torch.zeros(B, Hs, D, D, device=dev, dtype=torch.bfloat16)
        slots = torch.randperm(B, device=dev).to(torch.int32)
        out = torch.empty(B, H, D, device=dev, dtype=torch.bfloat16)
        gw = torch.ones(D, device=dev, dtype=torch.bfloat16) if mode != "plain" else None
        gate = 16.0 if mode == "gate" else None
        fns.append(lambda q=q, k=k, v=v, g=g, state=state, slots=slots, out=out:
                   ops.gla_decode(q, k, v, g, state, slots, out=out, gnorm_weight=gw, gate_normalizer=gate))
    t = time_graph(fns)
    print(f"gla_decode[{mode}] B={B} layout={layout}: {t*1e6:8.1f} us  {alg/t/1e9:8.1f} GB/s "
          f"({alg/t/1e9/PEAK:.2%} of measured {PEAK:.0f})  alg={alg/1e6:.1f} MB  [graph of {nrot} rotating buffers]")


def bench_swa_decode(B, ctx, H=28, Hkv=4, D=128, bs=256, wl=4096, M=0, flush=None, ragged=False):
    """ragged: lengths ~ U[1024, ctx] (config 4: 1000 seqs, <= 9K), else all == ctx."""
    dev = "cuda"
    nblk = (ctx + bs - 1) // bs
    nb = B * nblk + 1
    kv = torch.randn(2, nb, bs, Hkv, D, device=dev).bfloat16()
    tables = torch.randperm(nb - 1, device=dev).to(torch.int32).view(B, nblk)
    lens = torch.full((B,), ctx, dtype=torch.int32, device=dev)
    if ragged:
        lens = torch.randint(1024, ctx + 1, (B,), generator=torch.Generator().manual_seed(0)).to(torch.int32).to(dev)
    q = torch.randn(B, H, D, device=dev).bfloat16()
    mk = mv = None
    if M:
        mk = torch.randn(M, Hkv, D, device=dev).bfloat16()
        mv = torch.randn(M, Hkv, D, device=dev).bfloat16()
    out = torch.empty(B, H, D, device=dev, dtype=torch.bfloat16)
    t = timeit(lambda: ops.swa_meta_decode_paged(q, kv[0], kv[1], tables, lens, mk, mv, wl, out=out),
               flush=flush)
    nk = (lens.clamp(max=wl + 1) if wl >= 0 else lens).sum().item()
    alg = Hkv * 2 * D * 2 * nk + B * H * 2 * D * 2 + Hkv * 2 * D * 2 * M
    print(f"swa_decode B={B} ctx={'<=' if ragged else ''}{ctx} wl={wl} M={M}: {t*1e6:8.1f} us  {alg/t/1e9:8.1f} GB/s "
          f"({alg/t/1e9/PEAK:.2%})  alg={alg/1e6:.1f} MB")


def bench_kv_write(T, Hkv=4, D=128, bs=256, flush=None):
    dev = "cuda"
    nb = (T + bs - 1) // bs
    kv = torch.zeros(2, nb, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
    key = torch.randn(T, Hkv, D, device=dev).bfloat16()
    val = torch.randn(T, Hkv, D, device=dev).bfloat16()
    slots = torch.randperm(nb * bs, device=dev)[:T].to(torch.int64)
    t = timeit(lambda: ops.kv_write_paged(key, val, kv[0], kv[1], slots), flush=flush)
    alg = T * (2 * Hkv * D * 2 * 2 + 8)
    print(f"kv_write T={T}: {t*1e6:8.1f} us  {alg/t/1e9:8.1f} GB/s ({alg/t/1e9/PEAK:.2%})")


def bench_swa_prefill(N, T, H=28, Hkv=4, D=128, wl=4096, M=0, iters=5):
    dev = "cuda"
    q = torch.randn(N * T, H, D, device=dev).bfloat16()
    k = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    mk = mv = None
    if M:
        mk = torch.randn(M, Hkv, D, device=dev).bfloat16()
        mv = torch.randn(M, Hkv, D, device=dev).bfloat16()
    cu = (torch.arange(N + 1, device=dev) * T).to(torch.int32)
    out = torch.empty(N * T, H, D, device=dev, dtype=torch.bfloat16)
    t = timeit(lambda: ops.swa_meta_prefill_varlen(q, k, v, None, mk, mv, cu, cu, T, wl, out=out),
               iters=iters, warmup=2)
    i = torch.arange(T, dtype=torch.float64)
    nkeys = (M + (torch.clamp(i + 1, max=wl + 1) if wl >= 0 else i + 1)).sum().item()
    fl = N * H * 4 * D * nkeys
    print(f"swa_prefill N={N} T={T} H={H}/{Hkv} wl={wl} M={M}: {t*1e3:8.3f} ms  {fl/t/1e12:7.1f} TFLOP/s "
          f"({fl/t/1e12/TPEAK:.2%} of sustained {TPEAK:.0f})  {N*T/t/1e6:.2f} Mtok/s")


def bench_rope_kv(T, H=28, Hkv=4, D=128, bs=256, flush=None):
    """Fused SWA prologue (f3): RoPE(q,k) in place + paged write of rotated k and v, against the
    three launches it replaces here (rotate-only launch + kv_write_paged)."""
    dev = "cuda"
    nb = (T + bs - 1) // bs
    kv = torch.zeros(2, nb, bs, Hkv, D, device=dev, dtype=torch.bfloat16)
    qkv = torch.randn(T, (H + 2 * Hkv) * D, device=dev).bfloat16()
    q, k, v = (x.view(T, -1, D) for x in qkv.split([H * D, Hkv * D, Hkv * D], dim=-1))
    table = ops.rope_cos_sin_cache(D, 16384, 1e6, device=dev)
    pos = torch.randint(0, 16000, (T,), device=dev)
    slots = torch.randperm(nb * bs, device=dev)[:T].to(torch.int64)
    t = timeit(lambda: ops.rope_kv_write_paged(q, k, v, table, pos, 128, kv[0], kv[1], slots), flush=flush)
    def unfused():
        ops.rope_kv_write_paged(q, k, None, table, pos, 128)
        ops.kv_write_paged(k, v, kv[0], kv[1], slots)
    t2 = timeit(unfused, flush=flush)
    # q, k read + written, v read, k and v written to pages, cos|sin row + position + slot per token
    alg = T * ((H + Hkv) * D * 2 * 2 + Hkv * D * 2 + 2 * Hkv * D * 2 + D * 2 + 16)
    print(f"rope_kv_write T={T}: {t*1e6:8.1f} us  {alg/t/1e9:8.1f} GB/s ({alg/t/1e9/PEAK:.2%})   "
          f"[rotate + kv_write as two launches: {t2*1e6:8.1f} us]")


def bench_branch_merge(T, hidden=3584, flush=None):
    """f1: norm1 + norm2 + average + bf16 cast in one pass vs the reference's torch op sequence."""
    dev = "cuda"
    xa = torch.randn(T, hidden, device=dev).bfloat16()
    xl = torch.randn(T, hidden, device=dev).bfloat16()
    w1 = torch.ones(hidden, device=dev)
    w2 = torch.ones(hidden, device=dev)
    out = torch.empty_like(xa)
    t = timeit(lambda: ops.branch_norm_merge(xa, xl, w1, w2, 1e-6, out=out), flush=flush, iters=10)

    def ref_norm(x, w):
        xf = x.float()
        xf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6)
        return xf.to(torch.bfloat16) * w
    t2 = timeit(lambda: ((ref_norm(xa, w1) + ref_norm(xl, w2)) / 2).to(torch.bfloat16), flush=flush, iters=5)
    alg = T * hidden * 2 * 3
    print(f"branch_norm_merge T={T} hidden={hidden}: {t*1e6:8.1f} us  {alg/t/1e9:8.1f} GB/s ({alg/t/1e9/PEAK:.2%})   "
          f"[torch op sequence: {t2*1e6:8.1f} us]")


def bench_gla_gnorm(N, T, B, H=28, Hkv=4, D=128):
    """g_norm epilogue (f2): fused vs kernel + separate torch RMSNorm pass."""
    dev = "cuda"
    w = torch.ones(D, device=dev, dtype=torch.bfloat16)
    q = torch.randn(N * T, H, D, device=dev).bfloat16().relu()
    k = torch.randn(N * T, Hkv, D, device=dev).bfloat16().relu()
    v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(N * T, Hkv, D, device=dev)) / 16).bfloat16()
    state = torch.zeros(max(N, B), Hkv, D, D, device=dev, dtype=torch.bfloat16)
    cu = (torch.arange(N + 1, device=dev) * T).to(torch.int32)
    slots = torch.arange(max(N, B), device=dev, dtype=torch.int32)
    init = torch.zeros(N, device=dev, dtype=torch.int32)
    out = torch.empty(N * T, H, D, device=dev, dtype=torch.bfloat16)

    def norm(x):
        xf = x.float()
        return w * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6)).to(x.dtype)
    run = lambda **kw: ops.gla_prefill_varlen(q, k, v, g, state, cu, slots[:N], init, max_seqlen=T, out=out, **kw)
    t0 = timeit(lambda: run(), iters=5, warmup=2)
    t1 = timeit(lambda: run(gnorm_weight=w), iters=5, warmup=2)
    t2 = timeit(lambda: norm(run()), iters=5, warmup=2)
    print(f"gla_prefill N={N} T={T} g_norm: plain {t0*1e3:.3f} ms | fused epilogue {t1*1e3:.3f} ms | "
          f"kernel + torch RMSNorm {t2*1e3:.3f} ms")
    dq, dk, dv, dg, do = q[:B], k[:B], v[:B], g[:B], out[:B]
    drun = lambda **kw: ops.gla_decode(dq, dk, dv, dg, state, slots[:B], out=do, **kw)
    t0 = timeit(lambda: drun(), iters=50)
    t1 = timeit(lambda: drun(gnorm_weight=w), iters=50)
    t2 = timeit(lambda: norm(drun()), iters=50)
    print(f"gla_decode B={B} g_norm (eager launches): plain {t0*1e6:.1f} us | fused epilogue {t1*1e6:.1f} us | "
          f"kernel + torch RMSNorm {t2*1e6:.1f} us")


def bench_gla_prefill(N, T, H=28, Hkv=4, D=128, gdt=torch.bfloat16, iters=5, gn=False):
    dev = "cuda"
    q = torch.randn(N * T, H, D, device=dev).bfloat16().relu()
    k = torch.randn(N * T, Hkv, D, device=dev).bfloat16().relu()
    v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(N * T, Hkv, D, device=dev)) / 16).to(gdt)
    state = torch.zeros(N, Hkv, D, D, device=dev, dtype=torch.bfloat16)
    cu = (torch.arange(N + 1, device=dev) * T).to(torch.int32)
    slots = torch.arange(N, device=dev, dtype=torch.int32)
    init = torch.zeros(N, device=dev, dtype=torch.int32)
    out = torch.empty(N * T, H, D, device=dev, dtype=torch.bfloat16)
    gw = torch.ones(D, device=dev, dtype=torch.bfloat16) if gn else None
    t = timeit(lambda: ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, init, max_seqlen=T, out=out,
                                              gnorm_weight=gw, relu_qk=gn), iters=iters, warmup=2)
    by = N * T * (H * 2 * D * 2 + Hkv * D * (4 + g.element_size()))
    fl = N * T * H * (4 * D * D + 2 * 64 * 2 * D)
    print(f"gla_prefill{' +relu +g_norm' if gn else ''} N={N} T={T} H={H}/{Hkv}: {t*1e3:8.3f} ms  {by/t/1e9:7.1f} GB/s ({by/t/1e9/PEAK:.2%})  "
          f"{fl/t/1e12:6.1f} TFLOP/s ({fl/t/1e12/TPEAK:.2%})  {N*T/t/1e6:.2f} Mtok/s")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", default="all")
    a = ap.parse_args()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    if a.which in ("all", "gla_decode"):
        for B in (64, 256, 1000):
            for mode in ("plain", "gn", "gate"):
                bench_gla_decode(B, flush=flush, mode=mode)
        bench_gla_decode(64, layout="reference", flush=flush)
        bench_gla_decode(1000, layout="reference", gdt=torch.float32, flush=flush)
    if a.which in ("all", "swa_decode"):
        bench_swa_decode(64, 8192, flush=flush)
        bench_swa_decode(64, 8704, M=128, wl=4095, flush=flush)
        bench_swa_decode(1000, 5000, flush=flush)
        bench_swa_decode(1000, 9216, flush=flush, ragged=True)     # config 4
        bench_swa_decode(8, 8192, flush=flush)
        bench_swa_decode(1, 8192, flush=flush)
    if a.which in ("all", "swa_prefill"):
        bench_swa_prefill(64, 8192)
        bench_swa_prefill(1, 32768, H=7, Hkv=1, wl=4095, M=128)
        bench_swa_prefill(1, 32768, H=7, Hkv=1, wl=-1, M=128)
    if a.which in ("all", "gla_prefill"):
        bench_gla_prefill(64, 8192)
        bench_gla_prefill(64, 8192, gn=True)
        bench_gla_prefill(8, 32768)
        bench_gla_prefill(1, 32768, H=7, Hkv=1, gdt=torch.float32)
    if a.which in ("all", "branch_merge"):
        bench_branch_merge(64, flush=flush)
        bench_branch_merge(65536, flush=flush)
    if a.which in ("all", "gnorm"):
        bench_gla_gnorm(64, 8192, 64)
        bench_gla_gnorm(8, 1024, 1000)
    if a.which in ("all", "rope_kv"):
        bench_rope_kv(64, flush=flush)
        bench_rope_kv(524288, flush=flush)
    if a.which in ("all", "kv_write"):
        bench_kv_write(64)
        bench_kv_write(524288, flush=flush)
