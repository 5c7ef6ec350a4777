"""Tiny driver for an ncu capture of the tcgen05 GEMMs (single-CTA, CTA-pair) next to cuBLAS on GPT-NeoX-1.3B shapes.

Run plainly first, then under `ncu --set full --clock-control none --import-source on --profile-from-start off`.
Shapes: ff1 fwd (K-major operands), ff2 fwd (long K), ff1 wgrad (both operands MN-major), ff1 dgrad (B MN-major).
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import kernels as K

which = sys.argv[1] if len(sys.argv) > 1 else "all"
shapes = {"ff1_fwd": (8192, 8192, 2048, False, True), "ff2_fwd": (8192, 2048, 8192, False, True),
          "ff1_wgrad": (8192, 2048, 8192, True, False), "ff1_dgrad": (8192, 2048, 8192, False, False)}
if which != "all":
    shapes = {k: v for k, v in shapes.items() if k in which.split(",")}
modes = (os.environ.get("NCU_MODES") or "single,pair,cublas").split(",")
bufs = []
for name, (M, N, Kk, ta, tb) in shapes.items():
    a = torch.randn((Kk, M) if ta else (M, Kk), device="cuda", dtype=torch.float16)
    b = torch.randn((N, Kk) if tb else (Kk, N), device="cuda", dtype=torch.float16)
    out = torch.empty(M, N, device="cuda", dtype=torch.float16)
    bufs.append((a, b, out, ta, tb))


def run(a, b, out, ta, tb, mode):
    if mode == "cublas":
        torch.matmul(a.t() if ta else a, b.t() if tb else b, out=out)
    else:
        K.GEMM_KERNEL = mode
        K.gemm(a, b, trans_a=ta, trans_b=tb, out=out)


for buf in bufs:
    for mode in modes:
        for _ in range(3):
            run(*buf, mode)
torch.cuda.synchronize()
torch.cuda.profiler.start()
for buf in bufs:
    for mode in modes:
        run(*buf, mode)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
