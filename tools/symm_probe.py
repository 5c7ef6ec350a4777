"""Probe: torch symmetric memory (cuMem VMM + multicast object) on this box. Launch with torch.distributed.run."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank = int(os.environ["RANK"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
t = symm_mem.empty(1 << 22, dtype=torch.uint8, device="cuda")
hdl = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
info = {"rank": rank, "world": hdl.world_size, "buffer_ptrs": [hex(p) for p in hdl.buffer_ptrs],
        "multicast_ptr": hex(hdl.multicast_ptr), "signal_pad_ptrs": [hex(p) for p in hdl.signal_pad_ptrs],
        "has_multicast": bool(hdl.multicast_ptr)}
print(info, flush=True)
# functional check with torch's own multimem op (library) so that the address is known good before our kernels use it
x = t.view(torch.bfloat16)
x.fill_(rank + 1)
dist.barrier()
torch.cuda.synchronize()
try:
    out = torch.ops.symm_mem.multimem_all_reduce_(x[:4096], "sum", dist.group.WORLD.group_name)
    torch.cuda.synchronize()
    print(rank, "multimem_all_reduce_ ->", float(x[0]), "expect", sum(range(1, hdl.world_size + 1)), flush=True)
except Exception as e:
    print(rank, "multimem_all_reduce_ failed:", repr(e)[:300], flush=True)
dist.barrier()
dist.destroy_process_group()
