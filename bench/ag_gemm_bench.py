"""Secondary: fused all-gather + GEMM (ZeRO-3 linear; weight tiles TMA-loaded straight from the owning rank's shard over
NVLink) vs the unfused path (NCCL all_gather_into_tensor of the weight, then our GEMM). Launch with
torch.distributed.run. Shapes: the GPT-NeoX-1.3B FF1 (8192 x 2048) and QKV (6144 x 2048) weights, 8192 tokens.
Roofline: the GEMM at the measured sustained tensor throughput (MEASURED_PEAKS.json) - the gathered bytes
((W-1)/W of the weight, a few MB) are two orders of magnitude below what NVLink moves in that time."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.parallel.ag_gemm import ShardedWeight, ag_linear, ag_linear_dgrad

deepspeed.init_distributed(verbose=False)
rank, world = dist.get_rank(), dist.get_world_size()
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
dt = torch.float16
tokens = 8192


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(t.item())
    return sorted(ts)[len(ts) // 2]


peak = 1381.1
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                             "MEASURED_PEAKS.json")))["bf16_tflops_sustained"])
except Exception:
    pass
res = {"world": world, "tokens": tokens, "peak_tflops": peak, "shapes": []}
for (Nn, Kk, tag) in [(8192, 2048, "ff1"), (6144, 2048, "qkv"), (2048, 8192, "ff2")]:
    if (Nn // world) % 256:
        continue
    torch.manual_seed(0)
    full = torch.randn(Nn, Kk, device="cuda", dtype=dt) * 0.02
    dist.broadcast(full, 0)
    sw = ShardedWeight(full, group=None)
    x = torch.randn(tokens, Kk, device="cuda", dtype=dt)
    gathered = torch.empty(Nn, Kk, device="cuda", dtype=dt)

    def unfused():
        dist.all_gather_into_tensor(gathered, sw.shard)
        return K.gemm(x, gathered)

    y_f = ag_linear(x, sw)
    y_u = unfused()
    err = ((y_f.float() - y_u.float()).abs().max() / y_u.float().abs().max()).item()
    y_d = ag_linear(x, sw, direct=True)
    err = max(err, ((y_d.float() - y_u.float()).abs().max() / y_u.float().abs().max()).item())
    t_f = timed(lambda: ag_linear(x, sw))
    t_d = timed(lambda: ag_linear(x, sw, direct=True), iters=8, warm=2)
    t_u = timed(unfused)
    t_local = timed(lambda: K.gemm(x, full))
    # input gradient of the same layer: the shards split the reduction dimension (dsb_gemm_ag_overlap_k)
    go = torch.randn(tokens, Nn, device="cuda", dtype=dt) * 0.1

    def unfused_dgrad():
        dist.all_gather_into_tensor(gathered, sw.shard)
        return K.gemm(go, gathered, trans_b=False)

    gd_f, gd_u = ag_linear_dgrad(go, sw), unfused_dgrad()
    err_d = ((gd_f.float() - gd_u.float()).abs().max() / gd_u.float().abs().max()).item()
    t_df = timed(lambda: ag_linear_dgrad(go, sw))
    t_du = timed(unfused_dgrad)
    fl = 2.0 * tokens * Nn * Kk
    res["shapes"].append({"tag": tag, "N": Nn, "K": Kk, "fused_ms": round(t_f, 4), "nccl_ag_plus_gemm_ms": round(t_u, 4),
                          "direct_remote_tma_ms": round(t_d, 4),
                          "local_gemm_ms": round(t_local, 4), "speedup": round(t_u / t_f, 2),
                          "fused_tflops": round(fl / t_f / 1e9, 1),
                          "fused_fraction_of_roofline": round(fl / t_f / 1e9 / peak, 3), "rel_err_vs_unfused": err,
                          "dgrad_fused_ms": round(t_df, 4), "dgrad_nccl_ag_plus_gemm_ms": round(t_du, 4),
                          "dgrad_speedup": round(t_du / t_df, 2),
                          "dgrad_fraction_of_roofline": round(fl / t_df / 1e9 / peak, 3),
                          "dgrad_rel_err_vs_unfused": err_d})
if rank == 0:
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/ag_gemm_bench_w{world}.json", "w"), indent=1)
dist.barrier()
dist.destroy_process_group()
