import sys, ctypes as C
sys.path.insert(0, "/root/repo")
import torch
from deeperspeed_b200.ops import attention as A
from deeperspeed_b200.ops import native as N
B, S, H, D = 4, 2048, 16, 128
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.float16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
dbg = torch.zeros(16, dtype=torch.int64, device="cuda")
lib = N.cuda_lib()
lib.dsb_flash_attn_debug_timers.argtypes = [C.c_void_p]
for _ in range(3):
    A.flash_attn_fwd(q, k, v, causal=True)
lib.dsb_flash_attn_debug_timers(dbg.data_ptr())
A.flash_attn_fwd(q, k, v, causal=True)
torch.cuda.synchronize()
d = dbg.tolist(); n = d[15]
print("blocks", n)
print("softmax warp cycles/block: wait_s %.0f  tmem_ld %.0f  math %.0f  wait_pv+rescale %.0f  store+fence+arrive %.0f" % tuple(x / n for x in d[0:5]))
print("mma thread cycles/block: wait_kv/s_free %.0f  issue_S %.0f  wait_p %.0f  issue_PV %.0f" % tuple(x / n for x in d[8:12]))
# ---- backward (CTA of key block 0: walks all 16 query blocks)
out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
do = torch.randn_like(out)
lib.dsb_flash_attn_debug_timers(None)
for _ in range(2):
    A.flash_attn_bwd(do, q, k, v, out, lse, causal=True)
dbg.zero_()
lib.dsb_flash_attn_debug_timers(dbg.data_ptr())
A.flash_attn_bwd(do, q, k, v, out, lse, causal=True)
torch.cuda.synchronize()
lib.dsb_flash_attn_debug_timers(None)
d = dbg.tolist(); n = d[15]
print("bwd query blocks", n)
print("bwd softmax warp cycles/block: wait_sdp %.0f  p/ds math+store %.0f  wait_mma2 %.0f  dq drain %.0f" % tuple(x / n for x in d[0:4]))
print("bwd mma thread cycles/block: wait_qdo %.0f  wait_s_free %.0f  issue S,dP %.0f  wait_pds %.0f  issue dV,dK,dQ %.0f" % tuple(x / n for x in d[8:13]))
