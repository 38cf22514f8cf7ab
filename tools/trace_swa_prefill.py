#!/usr/bin/env python3
"""Pipeline timeline of swa_prefill_kernel: builds a -DHYMETA_TRACE copy of the kernel library
(clock64 stamps of CTA (0,0,0) = head 0, last query pair of request 0, at every hand-off), runs the
config-2 shape and prints, per key step, when each role passed each event.

  python tools/trace_swa_prefill.py --build-only [--defs "-DPF_X=1 ..."] [--tag _x]   # here (no GPU)
  python tools/trace_swa_prefill.py [--tag _x] [--N 16] [--T 8192] [--first 10 --count 8]   # GPU box
"""
import argparse
import ctypes as C
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CSRC = os.path.join(ROOT, "vllm_hymeta_b200", "csrc")
EV = ["s0_S", "s0_ld", "s0_max", "s0_P", "s1_S", "s1_ld", "s1_max", "s1_P", "m_p0", "m_is0", "m_p1", "m_is1",
      "ld_K", "ld_V", "-", "-"]


def lib_path(tag):
    return os.path.join(CSRC, "build", f"libhymeta_b200_trace{tag}.so")


def build(defs, tag):
    out = lib_path(tag)
    os.makedirs(os.path.dirname(out), exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "--use_fast_math", "-Xcompiler", "-fPIC", "-DHYMETA_TRACE", *defs.split(), "-shared", "-cudart", "static",
           "-o", out] + srcs
    subprocess.run(cmd, check=True)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build-only", action="store_true")
    ap.add_argument("--defs", default="")
    ap.add_argument("--tag", default="")
    ap.add_argument("--T", type=int, default=8192)
    ap.add_argument("--N", type=int, default=16)
    ap.add_argument("--H", type=int, default=28)
    ap.add_argument("--Hkv", type=int, default=4)
    ap.add_argument("--wl", type=int, default=4096)
    ap.add_argument("--first", type=int, default=10)
    ap.add_argument("--count", type=int, default=8)
    a = ap.parse_args()
    path = lib_path(a.tag)
    if a.build_only or not os.path.exists(path):
        build(a.defs, a.tag)
        if a.build_only:
            return
    import torch
    from vllm_hymeta_b200 import _lib
    _lib.LIB_PATH = path
    from vllm_hymeta_b200 import ops
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    T, N, H, Hkv, D = a.T, a.N, a.H, a.Hkv, 128
    q = torch.randn(N * T, H, D, device=dev).bfloat16()
    k = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    cu = (torch.arange(N + 1, device=dev) * T).to(torch.int32)
    out = torch.empty(N * T, H, D, device=dev, dtype=torch.bfloat16)
    trace = torch.zeros(256 * 16, device=dev, dtype=torch.int64)
    h = _lib.lib()
    h.hymeta_debug_set_swa_trace.restype = C.c_int
    h.hymeta_debug_set_swa_trace.argtypes = [C.c_void_p]
    run = lambda: ops.swa_meta_prefill_varlen(q, k, v, None, None, None, cu, cu, T, a.wl, out=out)
    for _ in range(2):
        run()
    torch.cuda.synchronize()
    assert h.hymeta_debug_set_swa_trace(trace.data_ptr()) == 0
    run()
    torch.cuda.synchronize()
    h.hymeta_debug_set_swa_trace(None)
    tr = trace.cpu().view(256, 16)
    steps = int((tr[:, 0] > 0).sum())
    tr = tr[:steps]
    t0 = int(tr[tr > 0].min())
    total = int(tr.max()) - t0
    print(f"[{a.tag}] N={N} T={T} steps={steps} span={total} cycles -> {total / max(steps, 1):.0f} cycles/step")
    print("step " + " ".join(f"{e:>7s}" for e in EV[:14]))
    for i in range(a.first, min(steps, a.first + a.count)):
        base = int(tr[a.first][0])
        print(f"{i:4d} " + " ".join(f"{(int(x) - base) if x > 0 else -1:7d}" for x in tr[i][:14]))
    lo, hi = 6, steps - 4
    d = lambda i, e: int(tr[i][e])

    def avg(f):
        vals = [f(i) for i in range(lo, hi)]
        return sum(vals) / max(len(vals), 1)
    print("avg cycles over the unmasked middle steps:")
    print(f"  step period  (s0_S(t+1) - s0_S(t))            {avg(lambda i: d(i + 1, 0) - d(i, 0)):8.0f}")
    for i, nm in ((0, "tile 0"), (1, "tile 1")):
        o = 4 * i
        print(f"  {nm}: S landed -> LDTM done                  {avg(lambda t: d(t, o + 1) - d(t, o)):8.0f}")
        print(f"  {nm}: LDTM done -> max/rescale done          {avg(lambda t: d(t, o + 2) - d(t, o + 1)):8.0f}")
        print(f"  {nm}: exp + P store + arrive                 {avg(lambda t: d(t, o + 3) - d(t, o + 2)):8.0f}")
        print(f"  {nm}: P arrive -> MMA warp saw first group   {avg(lambda t: d(t, 8 + 2 * i) - d(t, o + 3)):8.0f}")
        print(f"  {nm}: MMA saw P -> PV + next S issued        {avg(lambda t: d(t, 9 + 2 * i) - d(t, 8 + 2 * i)):8.0f}")
        print(f"  {nm}: issue done -> S(t+1) landed (softmax)  {avg(lambda t: d(t + 1, o) - d(t, 9 + 2 * i)):8.0f}")
    print(f"  producer lead: K(t) issued before s0_S(t)      {avg(lambda t: d(t, 0) - d(t, 12)):8.0f}")


if __name__ == "__main__":
    main()
