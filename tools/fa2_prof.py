"""Tiny driver for ncu: a few forward + backward launches of the flash-attention kernels on the flagship shape."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import attention as A

B, S, H, D = (int(x) for x in (sys.argv[1:5] if len(sys.argv) >= 5 else (4, 2048, 16, 128)))
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.float16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
do = torch.randn(B, S, H, D, device="cuda", dtype=torch.float16)
for _ in range(3):
    out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
    A.flash_attn_bwd(do, q, k, v, out, lse, causal=True)
torch.cuda.synchronize()
print("ok")
