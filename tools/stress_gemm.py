"""Race hunt for the tcgen05 GEMMs: repeat one shape many times (interleaved with other launches) and report where
mismatching elements sit (row / column ranges -> which CTA of a pair, which tile, which epilogue chunk)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.ops import kernels as K

torch.manual_seed(0)
M, N, Kk = 4096, 2048, 2048
a = torch.randn(M, Kk, device="cuda", dtype=torch.float16)
b = torch.randn(N, Kk, device="cuda", dtype=torch.float16)
ref = a.float() @ b.float().t()
small_a = torch.randn(1000, 200, device="cuda", dtype=torch.float16)
small_b = torch.randn(264, 200, device="cuda", dtype=torch.float16)
bad_runs = 0
for mode in ("pair", "single"):
    K.GEMM_KERNEL = mode
    for it in range(200):
        if it % 3 == 0:
            K.GEMM_KERNEL = "single"
            K.gemm(small_a, small_b)
            K.GEMM_KERNEL = mode
        out = K.gemm(a, b)
        err = (out.float() - ref).abs()
        bad = err > 0.3
        nb = int(bad.sum())
        if nb:
            bad_runs += 1
            idx = bad.nonzero()
            rows, cols = idx[:, 0], idx[:, 1]
            print(f"{mode} it={it}: {nb} bad, max err {err.max().item():.3f}, rows {rows.min().item()}..{rows.max().item()} "
                  f"cols {cols.min().item()}..{cols.max().item()}; row%256 set {sorted(set((rows % 256 // 32).tolist()))[:8]} "
                  f"col%256/64 set {sorted(set((cols % 256 // 64).tolist()))}; first {idx[:6].tolist()}", flush=True)
    print(mode, "done, bad runs so far", bad_runs, flush=True)
