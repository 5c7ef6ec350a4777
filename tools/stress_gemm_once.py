"""One fresh-process replay of the first GEMM tests (three small single-CTA problems in fp16/bf16, then the first
CTA-pair launch of the process) with a detailed mismatch report; run in a loop to hunt first-launch races."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.ops import kernels as K

bad_total = 0
for shape in [(128, 256, 64), (256, 512, 128), (1000, 264, 200), (4096, 2048, 2048), (384, 128, 320)]:
    for dt in (torch.float16, torch.bfloat16):
        M, N, Kk = shape
        torch.manual_seed(0)
        a = torch.randn(M, Kk, device="cuda", dtype=dt)
        b = torch.randn(N, Kk, device="cuda", dtype=dt)
        out = K.gemm(a, b)
        ref = a.float() @ b.float().t()
        err = (out.float() - ref).abs()
        tol = (2e-3 if dt == torch.float16 else 1.5e-2) * ref.abs().max()
        bad = (err > tol).nonzero()
        if bad.numel():
            bad_total += 1
            r, c = bad[:, 0], bad[:, 1]
            print(f"MISMATCH {shape} {dt}: {bad.shape[0]} elems, max err {err.max().item():.4f} (tol {tol.item():.4f}), "
                  f"rows {r.min().item()}..{r.max().item()} cols {c.min().item()}..{c.max().item()}, "
                  f"row%256//32 {sorted(set((r % 256 // 32).tolist()))}, col%256//64 {sorted(set((c % 256 // 64).tolist()))}, "
                  f"first {bad[:10].tolist()}, out {[out[i, j].item() for i, j in bad[:4].tolist()]} "
                  f"ref {[round(ref[i, j].item(), 3) for i, j in bad[:4].tolist()]}", flush=True)
print("ok" if bad_total == 0 else f"bad={bad_total}")
