"""Secondary: weight-gradient GEMM + reduce-scatter in one kernel (tiles TMA-stored into the owner's staging over
NVLink while the tensor pipe keeps going, then a local sum) vs our GEMM followed by NCCL reduce_scatter, on the FF1 /
FF2 / QKV weight-gradient shapes of GPT-NeoX-1.3B (8192 tokens). Launch with torch.distributed.run.
Roofline: max(GEMM flops at the sustained cuBLAS rate, (W-1)/W of the result over NVLink at 770 GB/s)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import deeperspeed_b200 as deepspeed
from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.parallel.ag_gemm import GemmReduceScatter

deepspeed.init_distributed(verbose=False)
rank, world = dist.get_rank(), dist.get_world_size()
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
peaks = {}
try:
    peaks = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
except Exception:
    pass
tflops = float(peaks.get("bf16_tflops_sustained", 1400.0))
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tokens = 8192
shapes = {"ff1_wgrad": (8192, 2048), "ff2_wgrad": (2048, 8192), "qkv_wgrad": (6144, 2048)}
rs = GemmReduceScatter(8192 * 2048, dtype=torch.float16)


def timed(fn, iters=10, warm=3):
    ts = []
    for it in range(warm + iters):
        flush.zero_()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if it >= warm:
            ts.append(t.item())
    return sorted(ts)[len(ts) // 2]


res = {"world": world, "tokens": tokens, "shapes": {}}
for name, (out_f, in_f) in shapes.items():
    go = (torch.randn(tokens, out_f, device="cuda") * 0.05).half()
    x = (torch.randn(tokens, in_f, device="cuda") * 0.05).half()
    shard = torch.empty(out_f // world, in_f, dtype=torch.float16, device="cuda")

    def nccl():
        full = K.gemm(go, x, trans_a=True, trans_b=False)
        dist.reduce_scatter_tensor(shard, full, op=dist.ReduceOp.AVG)

    def fused():
        rs(go, x, trans_a=True, trans_b=False, out=shard)

    def gemm_only():
        K.gemm(go, x, trans_a=True, trans_b=False)

    t_n, t_f, t_g = timed(nccl), timed(fused), timed(gemm_only)
    flops_ms = 2.0 * tokens * out_f * in_f / (tflops * 1e12) * 1e3
    link_ms = (world - 1) / world * out_f * in_f * 2 / 770e9 * 1e3
    roof = max(flops_ms, link_ms)
    res["shapes"][name] = {"gemm_plus_nccl_rs_ms": round(t_n, 4), "fused_ms": round(t_f, 4), "gemm_alone_ms": round(t_g, 4),
                           "speedup": round(t_n / t_f, 3), "roofline_ms": round(roof, 4),
                           "fused_fraction_of_roofline": round(roof / t_f, 3)}
if rank == 0:
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/gemm_rs_bench_w{world}.json", "w"))
dist.barrier()
