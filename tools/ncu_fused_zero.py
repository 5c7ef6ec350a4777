"""Driver for an ncu capture of the fused ZeRO step kernels on ONE GPU (R = 1, the GPU is its own peer; ncu must not
wrap multi-rank commands): reduce + norm, Adam + push, LAMB + push on 256M elements (64 tensors). Run plainly first,
then under `ncu --set full --clock-control none --import-source on --profile-from-start off`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.parallel.fused_zero_single import SingleRankFusedZero

n = 256 << 20
z = SingleRankFusedZero(n, tensor_sizes=[n // 64] * 64)
z.master.normal_(0, 0.02)
z.grad.normal_(0, 1e-3)


def once(step):
    z.rs_norm()
    z.adam_ag(1e-4, (0.9, 0.999), 1e-8, 0.01, True, step)
    z.rs_norm()
    z.lamb_ag(1e-4, (0.9, 0.999), 1e-8, 0.01, 10.0, 0.01, step)


once(1); once(2)
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
ev[0].record(); z.rs_norm(); ev[1].record(); z.adam_ag(1e-4, (0.9, 0.999), 1e-8, 0.01, True, 3); ev[2].record()
z.rs_norm(); ev[3].record(); z.lamb_ag(1e-4, (0.9, 0.999), 1e-8, 0.01, 10.0, 0.01, 3); ev[4].record()
torch.cuda.synchronize()
gb = n / 1e9
print("rs_norm %.3f ms (%.0f GB/s)  adam_ag %.3f ms (%.0f GB/s)  lamb_ag %.3f ms (%.0f GB/s)" % (
    ev[0].elapsed_time(ev[1]), gb * 6 / ev[0].elapsed_time(ev[1]) * 1e3,
    ev[1].elapsed_time(ev[2]), gb * 30 / ev[1].elapsed_time(ev[2]) * 1e3,
    ev[3].elapsed_time(ev[4]), gb * 42 / ev[3].elapsed_time(ev[4]) * 1e3))
torch.cuda.profiler.start()
once(4)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
