"""Times the flash-attention v2 kernels with parts of their work switched off (dsb_flash_attn_debug_mode) to find
which resource bounds them. Results under gpurun_out/fa2_modes.json."""
import ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import attention as A
from deeperspeed_b200.ops import native as N

lib = N.cuda_lib()
lib.dsb_flash_attn_debug_mode.argtypes = [C.c_int]
lib.dsb_flash_attn_debug_mode.restype = None
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=8, warm=2):
    ts = []
    for it in range(warm + iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if it >= warm:
            ts.append(e0.elapsed_time(e1))
    return round(sorted(ts)[len(ts) // 2], 4)


res = {}
for (B, S, H) in ((4, 2048, 16), (1, 8192, 16)):
    qkv = torch.randn(B, S, 3, H, 128, device="cuda", dtype=torch.float16)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
    do = torch.randn_like(out)
    o2 = torch.empty_like(out)
    for m in (0, 7, 7 + 16, 7 + 32, 7 + 48, 32, 48, 16):
        lib.dsb_flash_attn_debug_mode(m)
        res["fwd S=%d mode=%d" % (S, m)] = timed(lambda: A.flash_attn_fwd(q, k, v, causal=True, out=o2, return_lse=False))
    for m in (0, 11):
        lib.dsb_flash_attn_debug_mode(m)
        res["bwd S=%d mode=%d" % (S, m)] = timed(lambda: A.flash_attn_bwd(do, q, k, v, out, lse, causal=True))
    lib.dsb_flash_attn_debug_mode(0)
    # fixed costs around the backward kernel
    res["bwd S=%d zeros+empty only" % S] = timed(lambda: (torch.zeros(B, S, H, 128, dtype=torch.float32, device="cuda"),
                                                          torch.empty(B, S, H, 128, dtype=torch.float16, device="cuda")))
for k_, v_ in res.items():
    print(k_, v_, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/fa2_modes.json", "w"), indent=1)
