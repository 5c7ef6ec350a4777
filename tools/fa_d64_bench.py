"""Head dim 64 on the flash kernels (dense, non-causal, as DeepSpeedTransformerLayer uses them) vs the library SDPA."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from deeperspeed_b200.ops.sparse_attention import fused

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
for (B, H, S, D) in ((16, 16, 512, 64), (64, 16, 128, 64), (2, 16, 8192, 64), (4, 16, 2048, 128)):
    qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.float16, requires_grad=True)
    q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))
    scale = D ** -0.5
    out = fused.BlockSparseFlashFn.apply(q, k, v, None, None, None, scale, None)
    g = torch.randn_like(out)
    t_f = timed(lambda: fused.BlockSparseFlashFn.apply(q, k, v, None, None, None, scale, None))
    t_b = timed(lambda: torch.autograd.grad(out, qkv, g, retain_graph=True))
    o2 = F.scaled_dot_product_attention(q, k, v)
    l_f = timed(lambda: F.scaled_dot_product_attention(q, k, v))
    l_b = timed(lambda: torch.autograd.grad(o2, qkv, g, retain_graph=True))
    row = {"B": B, "H": H, "S": S, "D": D, "ours_fwd_ms": round(t_f, 4), "ours_bwd_ms": round(t_b, 4),
           "sdpa_fwd_ms": round(l_f, 4), "sdpa_bwd_ms": round(l_b, 4)}
    res.append(row)
    print(row, flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/fa_d64_bench.json", "w"), indent=1)
