#!/usr/bin/env python3
"""A/B harness for compile-time variants of one kernel source (default swa_prefill.cu).

  python tools/ab_swa_prefill.py --build "base:-DPF_OPT_REGS=0 -DPF_OPT_PSPLIT=1" "regs:-DPF_OPT_PSPLIT=1" ...   # here
  python tools/ab_swa_prefill.py --run base regs ...                                                        # GPU box

Each variant is the source recompiled with its -D flags and linked with the other objects of the
product build (vllm_hymeta_b200/csrc/build/*.o) into build/libhymeta_ab_<tag>.so.  --run times every
variant in its own process at the config-2 shape (64 x 8192, 28/4 heads, window 4096), the config-3
shape (2 x 32768 + 128 meta tokens, window 4095) and a short-prompt shape, and prints an output
checksum and the max |difference| to the first variant so a variant that changes results is caught.
"""
import argparse
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CSRC = os.path.join(ROOT, "vllm_hymeta_b200", "csrc")
BUILD = os.path.join(CSRC, "build")


def lib_path(tag):
    return os.path.join(BUILD, f"libhymeta_ab_{tag}.so")


def build(tag, defs, src="swa_prefill"):
    obj = os.path.join(BUILD, f"ab_{tag}.o")
    flags = defs.split()
    fast = [] if "-PRECISE" in flags else ["--use_fast_math"]     # "-PRECISE": as __graft_entry__.PRECISE_SOURCES
    flags = [f for f in flags if f != "-PRECISE"]
    srcfile = os.path.join(CSRC, src + ".cu")
    for f in list(flags):                                          # "-SRC=path": compile another copy of the source
        if f.startswith("-SRC="):
            srcfile = f[5:]
            flags.remove(f)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           *fast, "-Xcompiler", "-fPIC", "-Xptxas", "-v", "-I", CSRC, *flags, "-c", srcfile, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode:
        sys.exit(r.stderr)
    info = [l.strip() for l in r.stderr.splitlines() if "spill" in l or "Used" in l]
    print(tag, defs, "|", " ".join(info[:2]))
    others = [o for o in glob.glob(os.path.join(BUILD, "*.o"))
              if os.path.basename(o) != src + ".o" and not os.path.basename(o).startswith("ab_")]
    subprocess.run(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib_path(tag), obj,
                    *others, "-cudart", "static"], check=True)


SHAPES = {"cfg2": (64, 8192, 28, 4, 4096, 0), "cfg3": (2, 32768, 28, 4, 4095, 128),
          "short": (256, 1024, 28, 4, 4096, 0)}


def run_one(tag, iters):
    import torch
    from vllm_hymeta_b200 import _lib
    _lib.LIB_PATH = lib_path(tag)
    from vllm_hymeta_b200 import ops
    dev = "cuda"
    res = {}
    for name, (N, T, H, Hkv, wl, M) in SHAPES.items():
        g = torch.Generator(device=dev).manual_seed(1)
        q = torch.randn(N * T, H, 128, device=dev, generator=g).bfloat16()
        k = torch.randn(N * T, Hkv, 128, device=dev, generator=g).bfloat16()
        v = torch.randn(N * T, Hkv, 128, device=dev, generator=g).bfloat16()
        mk = mv = None
        if M:
            mk = torch.randn(M, Hkv, 128, device=dev, generator=g).bfloat16()
            mv = torch.randn(M, Hkv, 128, device=dev, generator=g).bfloat16()
        cu = (torch.arange(N + 1, device=dev) * T).to(torch.int32)
        out = torch.empty(N * T, H, 128, device=dev, dtype=torch.bfloat16)
        fn = lambda: ops.swa_meta_prefill_varlen(q, k, v, None, mk, mv, cu, cu, T, wl, out=out)
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(iters):
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record(); fn(); e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        t = ts[len(ts) // 2] * 1e-3
        i = torch.arange(T, dtype=torch.float64)
        nkeys = (M + (torch.clamp(i + 1, max=wl + 1) if wl >= 0 else i + 1)).sum().item()
        fl = N * H * 4 * 128 * nkeys
        samp = out[:: max(1, N * T // 4096)].float().cpu()
        res[name] = samp
        print(f"{tag:>12s} {name}: {t*1e3:8.3f} ms  {fl/t/1e12:7.1f} TFLOP/s  checksum {samp.double().sum().item():.6f}",
              flush=True)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--build", nargs="*")
    ap.add_argument("--src", default="swa_prefill")
    ap.add_argument("--run", nargs="*")
    ap.add_argument("--one")
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    if a.build:
        for spec in a.build:
            tag, _, defs = spec.partition(":")
            build(tag, defs, a.src)
        return
    import torch
    if a.one:
        torch.save(run_one(a.one, a.iters), f"/tmp/ab_{a.one}.pt")
        return
    ref = None
    for tag in a.run:
        subprocess.run([sys.executable, __file__, "--one", tag, "--iters", str(a.iters)], check=False)
        try:
            cur = torch.load(f"/tmp/ab_{tag}.pt")
        except Exception:
            continue
        if ref is None:
            ref = cur
        else:
            print(f"{tag:>12s} max|d| vs {a.run[0]}: " +
                  " ".join(f"{k}={(cur[k] - ref[k]).abs().max().item():.3e}" for k in cur), flush=True)


if __name__ == "__main__":
    main()
