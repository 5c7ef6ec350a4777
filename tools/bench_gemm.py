"""Micro-benchmark: tcgen05 GEMM vs torch.matmul (cuBLAS) on the GPT-NeoX-1.3B layer shapes (CUDA events, L2 flushed)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import kernels as K

def timeit(fn, iters=20, warm=5):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]

res = []
dt = torch.float16
for (M, N, Kk, ta, tb, tag) in [
    (8192, 6144, 2048, False, True, "qkv fwd"), (8192, 2048, 2048, False, True, "proj fwd"),
    (8192, 8192, 2048, False, True, "ff1 fwd"), (8192, 2048, 8192, False, True, "ff2 fwd"),
    (8192, 2048, 8192, False, False, "ff1 dgrad"), (8192, 2048, 8192, True, False, "ff1 wgrad(M=8192out)"),
    (2048, 8192, 8192, True, False, "ff2 wgrad"), (8192, 50432, 2048, False, True, "lm head"),
    (4096, 4096, 4096, False, True, "square4k"), (8192, 8192, 8192, False, True, "square8k")]:
    a = torch.randn((Kk, M) if ta else (M, Kk), device="cuda", dtype=dt)
    b = torch.randn((N, Kk) if tb else (Kk, N), device="cuda", dtype=dt)
    out = torch.empty(M, N, device="cuda", dtype=dt)
    K.GEMM_KERNEL = "single"
    t_ours = timeit(lambda: K.gemm(a, b, trans_a=ta, trans_b=tb, out=out))
    K.GEMM_KERNEL = "pair"
    t_pair = timeit(lambda: K.gemm(a, b, trans_a=ta, trans_b=tb, out=out))
    at = a.t() if ta else a
    bt = b.t() if tb else b
    t_ref = timeit(lambda: torch.matmul(at, bt, out=out))
    fl = 2.0 * M * N * Kk
    res.append(dict(tag=tag, M=M, N=N, K=Kk, ours_ms=round(t_ours, 4), pair_ms=round(t_pair, 4), cublas_ms=round(t_ref, 4),
                    ours_tflops=round(fl / t_ours / 1e9, 1), pair_tflops=round(fl / t_pair / 1e9, 1), cublas_tflops=round(fl / t_ref / 1e9, 1)))
    print(res[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/bench_gemm.json", "w"), indent=1)

# ---- sustained mode: 150 back-to-back launches per kernel (power-limited regime, how a training step runs them)
def sustained(fn, iters=150):
    for _ in range(10):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters

sus = []
for (M, N, Kk, tag) in [(8192, 6144, 2048, "qkv fwd"), (8192, 8192, 2048, "ff1 fwd"), (8192, 2048, 8192, "ff2 fwd"),
                        (8192, 50432, 2048, "lm head")]:
    a = torch.randn(M, Kk, device="cuda", dtype=dt)
    b = torch.randn(N, Kk, device="cuda", dtype=dt)
    out = torch.empty(M, N, device="cuda", dtype=dt)
    r = dict(tag=tag, M=M, N=N, K=Kk)
    fl = 2.0 * M * N * Kk
    for rep in range(2):
        for name in ("single", "pair", "cublas"):
            if name == "cublas":
                t = sustained(lambda: torch.matmul(a, b.t(), out=out))
            else:
                K.GEMM_KERNEL = name
                t = sustained(lambda: K.gemm(a, b, out=out))
            r[f"{name}_tflops_{rep}"] = round(fl / t / 1e9, 1)
    sus.append(r)
    print("sustained", r, flush=True)
json.dump(dict(flushed=res, sustained=sus), open("gpurun_out/bench_gemm.json", "w"), indent=1)
