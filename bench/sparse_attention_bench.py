"""Secondary config: block-sparse (fixed layout) attention at long sequence vs dense-masked PyTorch attention on the same
GPU (the reference's Triton-0.4 kernels cannot be built any more). Forward + backward of one
attention call, CUDA events, median of 10."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops.sparse_attention import SparseSelfAttention, FixedSparsityConfig
from deeperspeed_b200.ops.sparse_attention.matmul import sparse_to_dense


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


res = []
for S in (2048, 4096, 8192):
    B, H, D, block = 2, 16, 64, 16
    cfg = FixedSparsityConfig(num_heads=H, block=block, num_local_blocks=8, num_global_blocks=1, attention="unidirectional")
    SparseSelfAttention.ops.clear()
    attn = SparseSelfAttention(cfg, max_seq_length=S).cuda()
    q, k, v = (torch.randn(B, H, S, D, device="cuda", dtype=torch.float16, requires_grad=True) for _ in range(3))
    layout = cfg.make_layout(S)
    density = layout.float().mean().item()

    def sparse():
        out = attn(q, k, v)
        out.sum().backward()

    mask = None

    def dense_masked():
        nonlocal_mask = dense_masked.mask
        sc = torch.matmul(q, k.transpose(-1, -2)) * D ** -0.5
        sc = sc.masked_fill(~nonlocal_mask, float("-inf"))
        out = torch.matmul(torch.softmax(sc, -1), v)
        out.sum().backward()

    dense_masked.mask = sparse_to_dense(torch.ones(1, int(layout.sum()), block, block), layout, block)[0].bool().cuda()[None]
    t_s = timeit(sparse)                       # one tcgen05 block-sparse flash kernel per direction
    SparseSelfAttention.use_fused = False
    t_u = timeit(sparse)                       # SDD + block-sparse softmax + DSD as separate operators
    SparseSelfAttention.use_fused = True
    try:
        t_d = timeit(dense_masked)
    except torch.cuda.OutOfMemoryError:
        t_d = None
    res.append({"seq": S, "density": round(density, 4), "sparse_ms": round(t_s, 3), "sparse_unfused_ms": round(t_u, 3),
                "dense_masked_ms": None if t_d is None else round(t_d, 3),
                "speedup": None if t_d is None else round(t_d / t_s, 2)})
    print(res[-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/sparse_attention_bench.json", "w"), indent=1)
