"""Driver for an ncu capture of the ONE-kernel ZeRO step (csrc/comm/fused_zero_step.cu) on one GPU (R = 1: the GPU is
its own peer; ncu must not wrap multi-rank commands): reduce + norm, in-kernel exchange, Adam, bulk-copy gather and
gradient clearing on 256M elements. Run plainly first, then under
`ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:fused_zero_step`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.parallel.fused_zero_single import SingleRankFusedZero

n = 256 << 20
z = SingleRankFusedZero(n)
z.master.normal_(0, 0.02)


def once(step):
    z.grad.normal_(0, 1e-3)
    return z.whole_step(1e-4, (0.9, 0.95), 1e-8, 0.01, True, step, loss_scale=1.0, clip=1.0)


once(1); once(2)
torch.cuda.synchronize()
z.grad.normal_(0, 1e-3)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
z.whole_step(1e-4, (0.9, 0.95), 1e-8, 0.01, True, 3, loss_scale=1.0, clip=1.0)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
# per element: grad read 2 + reduced write 2 + read 2 + p, m, v read 12 + write 12 + low-precision copy 2 + clearing 2
print("fused_zero_step %.3f ms on %d elements (%.0f GB/s of the 34 B / element it has to move)" % (ms, n, n * 34 / ms / 1e6))
torch.cuda.profiler.start()
once(4)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
