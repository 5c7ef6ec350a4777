"""BASELINE config #4: GPT-2 13B (40 layers, d 5120, 40 heads of 128) with ZeRO-3 + ZeRO-Offload (fp32 master and Adam
moments in pinned host memory, stepped by the AVX-512 CPU Adam, low-precision shards copied back on a side stream),
one node. The model is built under `zero.Init`, so no rank ever holds more than its 1 / N of the parameters.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/zero3_offload_bench.py

Env: LAYERS / HIDDEN / HEADS (defaults 40 / 5120 / 40 = 12.9 B parameters), SEQ (1024), MICRO (4), STEPS (2),
OFFLOAD=0 keeps the optimizer state on the GPUs (ZeRO-3 alone). Batches from pinned host memory, loss read back, CUDA
events, max over ranks. Writes gpurun_out/zero3_offload_bench_w<N>.json. Reference harness:
tests/model/Megatron_GPT2/run_perf_test.py; published figure: 13B on one V100-32GB at > 30 TFLOPS
(docs/_pages/features.md:116)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.models.gpt2 import GPT2Config, GPT2Model

LAYERS = int(os.environ.get("LAYERS", 40))
HIDDEN = int(os.environ.get("HIDDEN", 5120))
HEADS = int(os.environ.get("HEADS", 40))
SEQ = int(os.environ.get("SEQ", 1024))
MICRO = int(os.environ.get("MICRO", 4))
STEPS = int(os.environ.get("STEPS", 2))
OFFLOAD = os.environ.get("OFFLOAD", "1") != "0"


def main():
    deepspeed.init_distributed(verbose=False)
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    rank, world = dist.get_rank(), dist.get_world_size()
    cfg = GPT2Config(n_embd=HIDDEN, n_layer=LAYERS, n_head=HEADS, n_positions=SEQ)
    zero = {"stage": 3, "overlap_comm": True, "contiguous_gradients": True, "stage3_max_live_parameters": int(3e9),
            "stage3_prefetch_bucket_size": int(5e8), "stage3_param_persistence_threshold": int(1e5)}
    if OFFLOAD:
        zero["offload_optimizer"] = {"device": "cpu", "pin_memory": True}
    conf = {"train_batch_size": MICRO * world, "train_micro_batch_size_per_gpu": MICRO, "steps_per_print": 10 ** 9,
            "optimizer": {"type": "Adam", "params": {"lr": 1e-4, "betas": [0.9, 0.95], "weight_decay": 0.01}},
            "fp16": {"enabled": True, "loss_scale": 0, "initial_scale_power": 10}, "zero_optimization": zero,
            "gradient_clipping": 1.0}
    torch.manual_seed(1234)
    t0 = time.time()
    with deepspeed.zero.Init(param_dict=conf):
        model = GPT2Model(cfg).half()
    build_s = time.time() - t0
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=list(model.parameters()), config_params=conf)
    init_s = time.time() - t0
    g = torch.Generator().manual_seed(11 + rank)
    pool = []
    for _ in range(2):
        t = torch.randint(0, cfg.vocab_size, (MICRO, SEQ + 1), generator=g)
        pool.append((t[:, :-1].contiguous().pin_memory(), t[:, 1:].contiguous().pin_memory()))

    def step(i):
        x, y = pool[i % len(pool)]
        loss = engine(x.cuda(non_blocking=True), y.cuda(non_blocking=True))
        engine.backward(loss)
        engine.step()
        return float(loss)

    step(0)                                        # warm-up (allocations, first CPU Adam step)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.reset_peak_memory_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    wall0 = time.time()
    for i in range(STEPS):
        last = step(1 + i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / STEPS, (time.time() - wall0) * 1e3 / STEPS], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    peak = torch.tensor([torch.cuda.max_memory_allocated() / 2 ** 30], device="cuda")
    dist.all_reduce(peak, op=dist.ReduceOp.MAX)
    if rank == 0:
        params = cfg.num_params()
        step_s = ms[1].item() / 1e3                  # host wall clock: the CPU optimizer is part of the step
        tokens = MICRO * SEQ * world
        out = {"model": f"GPT-2 {params / 1e9:.1f}B ({LAYERS}L d{HIDDEN} h{HEADS}), seq {SEQ}", "params": params,
               "n_gpus": world, "micro_batch": MICRO, "steps": STEPS, "dtype": "fp16",
               "zero": "stage 3" + (" + optimizer offload (CPU Adam, pinned)" if OFFLOAD else ""),
               "ms_per_step_device": round(ms[0].item(), 1), "ms_per_step_wall": round(ms[1].item(), 1),
               "tokens_per_s": round(tokens / step_s), "model_tflops_per_gpu": round(6 * params * tokens / step_s / world / 1e12, 1),
               "peak_hbm_gib_max_rank": round(peak.item(), 2), "model_build_s": round(build_s, 1),
               "engine_init_s": round(init_s, 1), "last_loss": round(last, 4),
               "skipped_steps": int(engine.skipped_steps),
               "data": "synthetic tokens from pinned host memory, random-init weights"}
        print(json.dumps(out), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        tag = "" if OFFLOAD else "_nooffload"
        json.dump(out, open(f"gpurun_out/zero3_offload_bench_w{world}{tag}.json", "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
