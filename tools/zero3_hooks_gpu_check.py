"""One-GPU sanity check of the ZeRO-3 gradient-hook path on CUDA: an in-place net and a net that returns a parameter
train under stage 3 (world 1, fp16) and follow the same model under stage 0 within fp16 tolerance."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29531")
os.environ.setdefault("RANK", "0")
os.environ.setdefault("WORLD_SIZE", "1")
os.environ.setdefault("LOCAL_RANK", "0")
import copy

import torch

import deeperspeed_b200 as deepspeed


class ReturnsBias(torch.nn.Linear):
    def forward(self, x):
        return super().forward(x), self.bias


class Net(torch.nn.Module):
    def __init__(self, d=256):
        super().__init__()
        self.a, self.b = torch.nn.Linear(d, d), torch.nn.Linear(d, d)
        self.c = ReturnsBias(d, d)
        self.act = torch.nn.ReLU(inplace=True)

    def forward(self, x):
        h = self.a(x)
        h += x
        h = self.act(h)
        h = torch.utils.checkpoint.checkpoint(self.b, h, use_reentrant=False)
        out, bias = self.c(h)
        return (out + bias).float().pow(2).mean()


def run(stage, model, xs, nvme=None):
    zero = {"stage": stage, "stage3_param_persistence_threshold": 0}
    conf = {"train_batch_size": 8, "steps_per_print": 1000, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
            "fp16": {"enabled": True, "loss_scale": 128}, "zero_optimization": zero}
    if nvme:        # ZeRO-Infinity: parameter shard arenas in files, two pinned buffers, optimizer state in host memory
        zero.update({"sub_group_size": 70000, "offload_optimizer": {"device": "cpu", "pin_memory": True},
                     "offload_param": {"device": "nvme", "nvme_path": nvme, "buffer_count": 2, "buffer_size": 70000}})
        conf["aio"] = {"block_size": 1048576, "queue_depth": 8, "thread_count": 1}
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=conf)
    losses = []
    for x in xs:
        loss = engine(x)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    if nvme:
        sw = engine.optimizer.param_swapper
        print("arena swapper", sw.stats, "arenas", len(engine.optimizer.sub_groups))
        assert sw.stats["evictions"] > 0 and sw.stats["writes"] > 0
    return losses


torch.manual_seed(0)
base = Net().half().cuda()
xs = [torch.randn(8, 256, device="cuda", dtype=torch.half) for _ in range(4)]
l0 = run(0, copy.deepcopy(base), xs)
l3 = run(3, copy.deepcopy(base), xs)
print("stage0", l0)
print("stage3", l3)
assert all(abs(a - b) <= 2e-2 * max(1.0, abs(a)) for a, b in zip(l0, l3)), (l0, l3)
print("ZERO3_HOOKS_GPU_OK")
import tempfile
with tempfile.TemporaryDirectory() as d:
    ln = run(3, copy.deepcopy(base), xs, nvme=d)
print("stage3+nvme", ln)
# CPU Adam vs the device Adam round differently in the last bits of an fp16 weight: close, not bitwise
assert all(abs(a - b) <= 2e-2 * max(1.0, abs(a)) for a, b in zip(l0, ln)), (l0, ln)
print("ZERO3_NVME_GPU_OK")
