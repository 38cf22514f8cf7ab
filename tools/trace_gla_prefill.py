#!/usr/bin/env python3
"""Pipeline timeline of gla_prefill_kernel: builds a -DHYMETA_TRACE copy of the kernel library
(clock64 stamps of CTA (0,0) at every mbarrier hand-off), runs one chain and prints, per step,
when each role passed each event (SM cycles relative to the step's first event).

  python tools/trace_gla_prefill.py --build-only      # here (no GPU)
  python tools/trace_gla_prefill.py [--T 1024] [--N 1] [--first 40 --count 6]   # on the GPU box
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
TRACE_LIB = os.path.join(CSRC, "build", "libhymeta_b200_trace.so")
EV = ["gate_kf", "v_kf", "g_cum", "st_x", "q_top", "q_qf", "q_a", "q_p", "q_o", "is_top", "is_x",
      "is_wp", "is_p", "is_end", "g_land", "g_top"]


def build(exp=0):
    os.makedirs(os.path.dirname(TRACE_LIB), exist_ok=True)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "--use_fast_math", "-Xcompiler", "-fPIC", "-DHYMETA_TRACE", f"-DGP_EXP={exp}", "-shared", "-cudart", "static",
           "-o", TRACE_LIB] + srcs
    subprocess.run(cmd, check=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build-only", action="store_true")
    ap.add_argument("--T", type=int, default=1024)
    ap.add_argument("--N", type=int, default=1)
    ap.add_argument("--H", type=int, default=28)
    ap.add_argument("--Hkv", type=int, default=4)
    ap.add_argument("--first", type=int, default=40)
    ap.add_argument("--count", type=int, default=6)
    ap.add_argument("--fused", type=int, default=0,
                    help="1: raw bf16 gate rows through the 7B gate table; 2: + relu + g_norm epilogue")
    ap.add_argument("--exp", type=int, default=0, help="GP_EXP timing-experiment mask (results are wrong)")
    a = ap.parse_args()
    global TRACE_LIB
    if a.exp:
        TRACE_LIB = TRACE_LIB.replace(".so", f"_exp{a.exp}.so")
    if a.build_only or not os.path.exists(TRACE_LIB):
        build(a.exp)
        if a.build_only:
            return
    import torch
    from vllm_hymeta_b200 import _lib
    _lib.LIB_PATH = TRACE_LIB
    from vllm_hymeta_b200 import ops
    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    T, N, H, Hkv, D = a.T, a.N, a.H, a.Hkv, 128
    q = torch.randn(N * T, H, D, device=dev).relu().bfloat16()
    k = torch.randn(N * T, Hkv, D, device=dev).relu().bfloat16()
    v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(N * T, Hkv, D, device=dev)) / 16).bfloat16()
    state = torch.zeros(N, Hkv, D, D, device=dev, dtype=torch.bfloat16)
    cu = torch.arange(0, (N + 1) * T, T, device=dev, dtype=torch.int32)
    slots = torch.arange(N, device=dev, dtype=torch.int32)
    modes = torch.zeros(N, device=dev, dtype=torch.int32)
    steps = (T + 15) // 16
    trace = torch.zeros(4096 * 16, device=dev, dtype=torch.int64)
    h = _lib.lib()
    h.hymeta_debug_set_gla_trace.restype = C.c_int
    h.hymeta_debug_set_gla_trace.argtypes = [C.c_void_p]
    kw = {}
    if a.fused >= 1:
        g = torch.randn(N * T, Hkv, D, device=dev).bfloat16()          # raw gk projection rows
        kw["gate_table"] = ops.gla_gate_7b_table(16.0, q.device)
    if a.fused >= 2:
        kw.update(relu_qk=True, gnorm_weight=torch.ones(D, device=dev, dtype=torch.bfloat16))
    for _ in range(2):
        ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, modes, max_seqlen=T, **kw)
    torch.cuda.synchronize()
    assert h.hymeta_debug_set_gla_trace(trace.data_ptr()) == 0
    ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, modes, max_seqlen=T, **kw)
    torch.cuda.synchronize()
    h.hymeta_debug_set_gla_trace(None)
    tr = trace.cpu().view(4096, 16)[:steps]
    t0 = int(tr[tr > 0].min())
    total = int(tr.max()) - t0
    print(f"T={T} N={N} steps={steps} total={total} cycles -> {total / steps:.0f} cycles/step")
    print("step " + " ".join(f"{e:>8s}" for e in EV))
    for i in range(a.first, min(steps, a.first + a.count)):
        base = int(tr[i][tr[i] > 0].min())
        print(f"{i:4d} " + " ".join(f"{(int(x) - base) if x > 0 else -1:8d}" for x in tr[i]) +
              f"   (base {base - t0})")
    # average deltas along the two critical chains
    import itertools
    lo, hi = 8, steps - 4
    def avg(f):
        vals = [f(i) for i in range(lo, hi)]
        return sum(vals) / len(vals)
    d = lambda i, e: int(tr[i][e])
    print("avg cycles:")
    print(f"  step period (is_x(i+1)-is_x(i))      {avg(lambda i: d(i + 1, 10) - d(i, 10)):8.0f}")
    print(f"  gates: top -> raw landed             {avg(lambda i: d(i, 14) - d(i, 15)):8.0f}")
    print(f"  gates: landed -> cumsum done          {avg(lambda i: d(i, 2) - d(i, 14)):8.0f}")
    print(f"  gates: cumsum -> kf arrive (ke wait + writes) {avg(lambda i: d(i, 0) - d(i, 2)):8.0f}")
    print(f"  st_x(i) -> is_x(i+1)                  {avg(lambda i: d(i + 1, 10) - d(i, 3)):8.0f}")
    print(f"  is_top -> is_x wait                   {avg(lambda i: d(i, 10) - d(i, 9)):8.0f}")
    print(f"  is_wp -> is_p wait                    {avg(lambda i: d(i, 12) - d(i, 11)):8.0f}")
    print(f"  q_top -> q_qf (build)                 {avg(lambda i: d(i, 5) - d(i, 4)):8.0f}")
    print(f"  q_qf -> q_a                           {avg(lambda i: d(i, 6) - d(i, 5)):8.0f}")
    print(f"  q_a -> q_p (P conversion)             {avg(lambda i: d(i, 7) - d(i, 6)):8.0f}")
    print(f"  q_p -> q_o                            {avg(lambda i: d(i, 8) - d(i, 7)):8.0f}")
    print(f"  q_o(i) -> q_top(i+2) (epilogue)       {avg(lambda i: d(i + 2, 4) - d(i, 8)):8.0f}")
    print(f"  gate period                           {avg(lambda i: d(i + 1, 0) - d(i, 0)):8.0f}")
    print(f"  gate lead over issuer (is_x - gate_kf){avg(lambda i: d(i, 10) - d(i, 0)):8.0f}")


if __name__ == "__main__":
    main()
