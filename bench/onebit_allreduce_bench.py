"""Secondary config: 1-bit compressed all-reduce, fused peer-memory kernel vs the torch.distributed/NCCL restatement
(N = 300 * 2^20 fp32 like reference tests/onebit/test_nccl_perf.py:32). Launch with torch.distributed.run."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import deeperspeed_b200 as deepspeed
from deeperspeed_b200.runtime.comm.nccl import NcclBackend

deepspeed.init_distributed(verbose=False)
rank, world = dist.get_rank(), dist.get_world_size()
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
n = 300 * 2 ** 20
n = n // (32 * world) * (32 * world)
res = {}
for name, fused in (("fused", True), ("nccl_restatement", False)):
    be = NcclBackend(use_fused=fused)
    a = torch.rand(n, device="cuda") - 0.5
    we, se = torch.zeros(n, device="cuda"), torch.zeros(n // world, device="cuda")
    ts = []
    for it in range(8):
        x = a.clone()
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); be.compressed_allreduce(x, we, se, rank); e1.record(); torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if it >= 3:
            ts.append(t.item())
    ms = sorted(ts)[len(ts) // 2]
    res[name] = {"ms": round(ms, 3), "algo_GBps": round(n * 4 / ms / 1e6, 1)}
    if fused:
        res[name]["phase_ms_last_call"] = be.phase_times_ms()
if rank == 0:
    res["speedup"] = round(res["nccl_restatement"]["ms"] / res["fused"]["ms"], 2)
    res["world"], res["numel"] = world, n
    # roofline: the kernel is HBM-bound (the wire carries n/8 bytes): phases A+B+E move 24 B/element, the server phases
    # C+D 16 B per chunk element; copy bandwidth from MEASURED_PEAKS.json (6545 GB/s)
    hbm_bytes = 24.0 * n + 16.0 * n / world
    res["hbm_roofline_ms"] = round(hbm_bytes / 6545e6, 3)
    res["fused_fraction_of_roofline"] = round(res["hbm_roofline_ms"] / res["fused"]["ms"], 3)
    print(json.dumps(res))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/onebit_bench_w{world}.json", "w"))
dist.barrier()
