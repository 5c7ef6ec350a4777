"""Secondary: the tcgen05 flash-attention kernels vs the library SDPA (cuDNN) on the flagship attention shape
(batch 4, 16 heads, seq 2048, head dim 128, causal) and a long-sequence shape; CUDA events, L2 flushed between
iterations. FLOPs counted as 4 * B * H * S^2 * D (/2 when causal) for forward, 2.5 x that for backward."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from deeperspeed_b200.ops import attention as A

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=10, warm=3):
    ts = []
    for it in range(warm + iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if it >= warm:
            ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


res = []
for (B, S, H, D) in ((4, 2048, 16, 128), (1, 8192, 16, 128)):
    qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.float16)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    fl = 4.0 * B * H * S * S * D / 2
    A.set_flash_fwd_variant(128)
    t_128 = timed(lambda: A.flash_attn_fwd(q, k, v, causal=True, return_lse=True))
    A.set_flash_fwd_variant(64)
    t_64 = timed(lambda: A.flash_attn_fwd(q, k, v, causal=True, return_lse=True))
    A.set_flash_fwd_variant(2)            # the default: two query tiles per CTA, P kept in tensor memory
    t_ours = timed(lambda: A.flash_attn_fwd(q, k, v, causal=True, return_lse=True))
    qt, kt, vt = (t.transpose(1, 2) for t in (q, k, v))
    t_lib = timed(lambda: F.scaled_dot_product_attention(qt, kt, vt, is_causal=True))
    row = {"B": B, "S": S, "H": H, "D": D, "fwd_ours_ms": round(t_ours, 4), "fwd_gen1_64key_ms": round(t_64, 4), "fwd_gen1_128key_ms": round(t_128, 4), "fwd_sdpa_ms": round(t_lib, 4),
           "fwd_ours_tflops": round(fl / t_ours / 1e9, 1), "fwd_sdpa_tflops": round(fl / t_lib / 1e9, 1)}
    if hasattr(A, "flash_attn_bwd"):
        out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
        do = torch.randn_like(out)
        t_b = timed(lambda: A.flash_attn_bwd(do, q, k, v, out, lse, causal=True))     # incl. prep + dQ finish kernels
        A.set_flash_bwd_variant(1)
        t_b1 = timed(lambda: A.flash_attn_bwd(do, q, k, v, out, lse, causal=True))
        A.set_flash_bwd_variant(2)
        ql, kl, vl = (t.detach().clone().requires_grad_() for t in (qt, kt, vt))
        o = F.scaled_dot_product_attention(ql, kl, vl, is_causal=True)
        dot = do.transpose(1, 2)
        t_lb = timed(lambda: torch.autograd.grad(o, (ql, kl, vl), dot, retain_graph=True))
        row.update({"bwd_ours_ms": round(t_b, 4), "bwd_gen1_ms": round(t_b1, 4), "bwd_sdpa_ms": round(t_lb, 4),
                    "bwd_ours_tflops": round(2.5 * fl / t_b / 1e9, 1), "bwd_sdpa_tflops": round(2.5 * fl / t_lb / 1e9, 1)})
    res.append(row)
    print(row)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/flash_attn_bench.json", "w"), indent=1)
