"""Phase timers (clock64) of the flash-attention v2 forward CTA (0, 0) under a few debug modes."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import attention as A
from deeperspeed_b200.ops import native as N
lib = N.cuda_lib()
lib.dsb_flash_attn_debug_mode.argtypes = [C.c_int]; lib.dsb_flash_attn_debug_mode.restype = None
lib.dsb_flash_attn_debug_timers.argtypes = [C.c_void_p]; lib.dsb_flash_attn_debug_timers.restype = None
B, S, H = 4, 2048, 16
qkv = torch.randn(B, S, 3, H, 128, device="cuda", dtype=torch.float16)
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
for _ in range(3):
    A.flash_attn_fwd(q, k, v, causal=True)
prof = torch.zeros(32, dtype=torch.int64, device="cuda")
for mode in (0,):
    prof.zero_()
    lib.dsb_flash_attn_debug_mode(mode)
    lib.dsb_flash_attn_debug_timers(prof.data_ptr())
    A.flash_attn_fwd(q, k, v, causal=True)
    torch.cuda.synchronize()
    lib.dsb_flash_attn_debug_timers(None)
    p = prof.tolist()
    n = max(p[29], 1)
    print("mode", mode, "blocks", p[28], p[29], "total cycles/block", p[30] // max(p[28], 1), p[31] // n)
    print("  softmax A per block: wait_s %d ld %d max %d exp %d st %d arrive %d" % tuple(x // n for x in p[0:6]))
    print("  softmax B per block: wait_s %d ld %d max %d exp %d st %d arrive %d" % tuple(x // n for x in p[8:14]))
    print("  mma A per block: wait_k %d issue_s %d wait_v %d wait_p %d issue_pv %d" % tuple(x // max(p[28], 1) for x in p[16:21]))
    print("  mma B per block: wait_k %d issue_s %d wait_v %d wait_p %d issue_pv %d" % tuple(x // n for x in p[22:27]))
lib.dsb_flash_attn_debug_mode(0)

out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
do = torch.randn_like(out)
for _ in range(2):
    A.flash_attn_bwd(do, q, k, v, out, lse, causal=True)
for mode in (0, 16, 32, 48, 64, 1, 1 + 64, 1 + 48 + 64, 4, 2):
    prof.zero_()
    lib.dsb_flash_attn_debug_mode(mode)
    lib.dsb_flash_attn_debug_timers(prof.data_ptr())
    A.flash_attn_bwd(do, q, k, v, out, lse, causal=True)
    torch.cuda.synchronize()
    lib.dsb_flash_attn_debug_timers(None)
    p = prof.tolist()
    n = max(p[31], 1)
    print("bwd mode", mode, "iterations", p[31], "cycles/iteration", p[30] // n)
    print("  mma A: wait_q %d wait_stfree %d issue_st %d wait_dpfree %d wait_do %d issue_dpt %d" % tuple(x // n for x in p[0:6]))
    print("  mma B: wait_pds %d issue_dv_dk %d wait_dqfree %d issue_dqt %d" % tuple(x // n for x in p[8:12]))
    print("  softmax: wait_s %d wait_dp %d ld %d math %d st+wait_dsfree %d smem+fence+arrive %d" % tuple(x // n for x in p[16:22]))
    print("  drain: wait_dq %d ld %d wait_prev_reduce %d sts+fence+issue %d" % tuple(x // n for x in p[24:28]))
lib.dsb_flash_attn_debug_mode(0)
