import sys, time, torch
sys.path.insert(0, "/root/repo")
from vllm_hymeta_b200 import ops
dev = torch.device("cuda:0")
def run(N, T, H=28, Hkv=4, reps=3):
    D = 128
    q = torch.randn(N * T, H, D, device=dev).relu().bfloat16()
    k = torch.randn(N * T, Hkv, D, device=dev).relu().bfloat16()
    v = torch.randn(N * T, Hkv, D, device=dev).bfloat16()
    g = (torch.nn.functional.logsigmoid(torch.randn(N * T, Hkv, D, device=dev)) / 16).bfloat16()
    state = torch.zeros(N, Hkv, D, D, device=dev, dtype=torch.bfloat16)
    cu = torch.arange(0, (N + 1) * T, T, device=dev, dtype=torch.int32)
    slots = torch.arange(N, device=dev, dtype=torch.int32)
    modes = torch.zeros(N, device=dev, dtype=torch.int32)
    o = torch.empty_like(q)
    for _ in range(2):
        ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, modes, max_seqlen=T, out=o)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        ops.gla_prefill_varlen(q, k, v, g, state, cu, slots, modes, max_seqlen=T, out=o)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"N={N} T={T} H={H}/{Hkv}: {ms:.3f} ms  {N*T*17408*(Hkv/4)/ms/1e6:.0f} GB/s", flush=True)
for N, T in [(1, 2048), (4, 2048), (16, 2048), (37, 2048), (64, 2048), (64, 8192)]:
    run(N, T)
print("long-context / few chains (auto segmentation)")
for N, T, H, Hkv in [(1, 32768, 7, 1), (8, 32768, 28, 4), (8, 131072, 28, 4), (1, 131072, 7, 1), (2, 8192, 28, 4)]:
    run(N, T, H, Hkv)
