"""Secondary (host only): DeepSpeedCPUAdam (AVX-512/AVX2 + OpenMP, csrc/host/cpu_adam.cpp) vs torch.optim.AdamW on the
CPU, the comparison behind the reference's "5-7x faster than torch Adam" claim (deepspeed/ops/adam/cpu_adam.py:30)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.ops.adam import DeepSpeedCPUAdam
from deeperspeed_b200.ops.adam.cpu_adam import _lib

res = {"threads": torch.get_num_threads(), "simd_level": _lib().dsb_cpu_adam_simd_level(), "sizes": []}
for n in (1 << 20, 1 << 24, 1 << 26):
    p1 = torch.nn.Parameter(torch.randn(n))
    p2 = torch.nn.Parameter(p1.detach().clone())
    g = torch.randn(n)
    p1.grad, p2.grad = g.clone(), g.clone()
    ours = DeepSpeedCPUAdam([p1], lr=1e-3, weight_decay=0.01)
    ref = torch.optim.AdamW([p2], lr=1e-3, weight_decay=0.01)

    def timed(opt, reps):
        opt.step()
        t0 = time.perf_counter()
        for _ in range(reps):
            opt.step()
        return (time.perf_counter() - t0) / reps * 1e3

    reps = 20 if n <= (1 << 24) else 5
    t_ours, t_ref = timed(ours, reps), timed(ref, reps)
    # 4 fp32 streams read (p, g, m, v) + 3 written
    res["sizes"].append({"numel": n, "ours_ms": round(t_ours, 3), "torch_adamw_ms": round(t_ref, 3),
                         "speedup": round(t_ref / t_ours, 2), "ours_GBps": round(n * 28 / t_ours / 1e6, 1)})
    print(res["sizes"][-1], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/cpu_adam_bench.json", "w"), indent=1)
print(res["threads"], "threads, simd level", res["simd_level"])
