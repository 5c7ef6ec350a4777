"""Driver for an ncu capture of the flash-attention kernels (the default second-generation forward and backward) on the
flagship attention shape (batch 4, 16 heads, seq 2048, head dim 128, causal). Run plainly first, then under
`ncu --set full --clock-control none --import-source on --profile-from-start off`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import attention as A

B, S, H, D = 4, 2048, 16, 128
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.float16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
dout = torch.randn(B, S, H, D, device="cuda", dtype=torch.float16)


def once():
    A.set_flash_fwd_variant(2)           # flash_fwd2_kernel (default)
    out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
    A.set_flash_bwd_variant(2)           # flash_prep + flash_bwd2_kernel + flash_dq_finish (default)
    A.flash_attn_bwd(dout, q, k, v, out, lse, causal=True)


once(); once()
torch.cuda.synchronize()
torch.cuda.profiler.start()
once()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
