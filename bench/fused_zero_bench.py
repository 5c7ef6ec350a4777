"""Secondary: fused NVLink ZeRO step (RS + cast + norm, Adam + all-gather over peer memory) vs the NCCL path
(reduce_scatter + sumsq + adam + all_gather) on a GPT-NeoX-stage-sized arena. Launch with torch.distributed.run.
Roofline: bytes over NVLink per rank = 2 * (W-1)/W * 2N (fp16 in, fp16 out) at 770 GB/s measured peer bandwidth."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import deeperspeed_b200 as deepspeed
from deeperspeed_b200.ops.adam import FusedAdam
from deeperspeed_b200.runtime.zero.stage1 import FP16_DeepSpeedZeroOptimizer_Stage1

deepspeed.init_distributed(verbose=False)
rank, world = dist.get_rank(), dist.get_world_size()
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
N = int(os.environ.get("NUMEL", 700_000_000)) // (world * 8) * (world * 8)
res = {}
for name, fused in (("nccl", False), ("fused", True)):
    p = torch.nn.Parameter(torch.randn(N, device="cuda", dtype=torch.float16) * 0.02)
    opt = FP16_DeepSpeedZeroOptimizer_Stage1(FusedAdam([p], lr=1e-4), dynamic_loss_scale=False, static_loss_scale=1.0,
                                             dp_process_group=dist.group.WORLD, max_elements_per_comm=int(2e9),
                                             fused_nvlink=fused, clip_grad=1.0)
    ts = []
    for it in range(8):
        opt.arenas[0].grad.normal_(0, 1e-3)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); opt.step(); e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if it >= 3:
            ts.append(t.item())
    res[name] = round(sorted(ts)[len(ts) // 2], 3)
    del opt, p
    torch.cuda.empty_cache()
if rank == 0:
    nvlink_bytes = 2 * (world - 1) / world * 2 * N
    hbm_bytes = N / world * (2 * world + 4 + 4 + 12 + 12 + 2 * world)
    roof_ms = max(nvlink_bytes / 770e9, hbm_bytes / 6545e9) * 1e3
    out = {"world": world, "numel": N, "nccl_ms": res["nccl"], "fused_ms": res["fused"],
           "speedup": round(res["nccl"] / res["fused"], 2), "roofline_ms": round(roof_ms, 3),
           "fused_fraction_of_roofline": round(roof_ms / res["fused"], 3)}
    print(json.dumps(out))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open(f"gpurun_out/fused_zero_bench_w{world}.json", "w"))
dist.barrier()
