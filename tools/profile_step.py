"""Kernel-time breakdown of one flagship train_batch (torch.profiler / CUPTI). Analysis only — never a bench number."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29544")
os.environ.setdefault("RANK", "0"); os.environ.setdefault("LOCAL_RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1")
import deeperspeed_b200 as deepspeed
from deeperspeed_b200.models.gpt_neox import GPTNeoXPipe, NeoXConfig
import bench

layers = int(sys.argv[1]) if len(sys.argv) > 1 else 24
deepspeed.init_distributed(verbose=False)
cfg = NeoXConfig.neox_1p3b(); cfg.num_layers = layers
model = GPTNeoXPipe(cfg, num_stages=1)
conf = bench.ds_config(1, 1, 1)
conf["gradient_accumulation_steps"] = 2; conf["train_batch_size"] = 8
engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=list(model.parameters()), config_params=conf)
it = bench.DeviceBatches(cfg.vocab_size, engine.device)
for _ in range(2):
    engine.train_batch(it)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU], record_shapes=True) as prof:
    engine.train_batch(it)
    torch.cuda.synchronize()
tab = prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=70)
os.makedirs("gpurun_out", exist_ok=True)
open("gpurun_out/profile_step.txt", "w").write(tab)
print(tab[-9000:])
# which framework (aten) ops still launch kernels, by input shape: candidates for fusion into our own kernels
ops = [e for e in prof.key_averages(group_by_input_shape=True) if e.key.startswith("aten::") and e.self_device_time_total > 0]
ops.sort(key=lambda e: -e.self_device_time_total)
lines = ["%-40s %10.1f us  x%-4d %s" % (e.key, e.self_device_time_total, e.count, str(e.input_shapes)[:110]) for e in ops[:30]]
open("gpurun_out/profile_step_aten_ops.txt", "w").write("\n".join(lines))
print("\n".join(lines))
