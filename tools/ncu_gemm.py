"""Tiny driver for an ncu capture of the tcgen05 GEMMs (single-CTA, CTA-pair) next to cuBLAS on two GPT-NeoX-1.3B shapes.

Run plainly first, then under `ncu --set full --clock-control none --import-source on --profile-from-start off`.
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import kernels as K

shapes = [(8192, 8192, 2048), (8192, 2048, 8192)]
bufs = []
for M, N, Kk in shapes:
    a = torch.randn(M, Kk, device="cuda", dtype=torch.float16)
    b = torch.randn(N, Kk, device="cuda", dtype=torch.float16)
    out = torch.empty(M, N, device="cuda", dtype=torch.float16)
    bufs.append((a, b, out))
    for mode in ("single", "pair"):
        K.GEMM_KERNEL = mode
        for _ in range(3):
            K.gemm(a, b, out=out)
    for _ in range(3):
        torch.matmul(a, b.t(), out=out)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for a, b, out in bufs:
    for mode in ("single", "pair"):
        K.GEMM_KERNEL = mode
        K.gemm(a, b, out=out)
    torch.matmul(a, b.t(), out=out)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
