"""BASELINE config #5 end to end: GPT-2 medium (24 layers, d 1024, 16 heads of 64), sequence 8192, block-sparse "fixed"
attention, trained through our engine (fp16, ZeRO-1, data parallel over all GPUs), against the SAME model and engine
with dense causal attention on the same box (the reference's Triton-C 0.4 kernels cannot be built any more): once as
materialised masked attention - the dense path the reference's sparse-attention speed-ups are quoted against - and once
as the library's fused causal kernel.

    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 bench/sparse_gpt2_bench.py

Every step: micro-batch of MICRO x 8192 tokens per GPU from pinned host memory, forward + backward + optimizer step,
loss read back. CUDA events on each rank, max over ranks. Writes gpurun_out/sparse_gpt2_bench_w<N>.json."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.models.gpt2 import GPT2Config, GPT2Model
from deeperspeed_b200.ops.sparse_attention import FixedSparsityConfig

SEQ = int(os.environ.get("SEQ", 8192))
MICRO = int(os.environ.get("MICRO", 2))
STEPS = int(os.environ.get("STEPS", 8))
WARMUP = int(os.environ.get("WARMUP", 3))
LAYERS = int(os.environ.get("LAYERS", 24))


def run(kind):
    torch.manual_seed(1234)
    sparsity = None
    if kind == "sparse":
        sparsity = FixedSparsityConfig(num_heads=16, block=16, num_local_blocks=8, num_global_blocks=1,
                                       attention="unidirectional")
    cfg = GPT2Config.medium(n_positions=SEQ, sparsity_config=sparsity,
                            attention_impl="masked" if kind == "dense_masked" else "flash",
                            checkpoint_blocks=(kind == "dense_masked"))      # 4.3 GB of scores per layer otherwise
    cfg.n_layer = LAYERS
    model = GPT2Model(cfg).half().cuda()
    world = dist.get_world_size()
    conf = {"train_batch_size": MICRO * world, "train_micro_batch_size_per_gpu": MICRO, "steps_per_print": 10 ** 9,
            "optimizer": {"type": "Adam", "params": {"lr": 1e-4, "betas": [0.9, 0.95], "weight_decay": 0.01}},
            "fp16": {"enabled": True, "loss_scale": 0, "initial_scale_power": 12},
            "zero_optimization": {"stage": 1}, "gradient_clipping": 1.0}
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=list(model.parameters()), config_params=conf)
    g = torch.Generator().manual_seed(7 + dist.get_rank())
    pool = []
    for _ in range(4):
        t = torch.randint(0, cfg.vocab_size, (MICRO, SEQ + 1), generator=g)
        pool.append((t[:, :-1].contiguous().pin_memory(), t[:, 1:].contiguous().pin_memory()))

    def step(i):
        x, y = pool[i % len(pool)]
        loss = engine(x.cuda(non_blocking=True), y.cuda(non_blocking=True))
        engine.backward(loss)
        engine.step()
        return float(loss)

    for i in range(WARMUP):
        step(i)
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(STEPS):
        last = step(i)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / STEPS], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    peak = torch.cuda.max_memory_allocated() / 2 ** 30
    density = float(sparsity.make_layout(SEQ).float().mean()) if sparsity is not None else 0.5
    del engine, model
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    return {"ms_per_step": round(ms.item(), 2), "tokens_per_s": round(MICRO * SEQ * world / (ms.item() / 1e3)),
            "last_loss": round(last, 4), "peak_gib": round(peak, 2), "layout_density": round(density, 4)}


def main():
    deepspeed.init_distributed(verbose=False)
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    world = dist.get_world_size()
    out = {"model": f"GPT-2 medium ({LAYERS}L d1024 h16), seq {SEQ}", "n_gpus": world, "micro_batch": MICRO,
           "steps": STEPS, "warmup": WARMUP, "dtype": "fp16", "parallelism": f"dp{world} zero1",
           "data": "synthetic tokens from pinned host memory, random-init weights"}
    for kind in ("sparse", "dense_flash", "dense_masked"):
        t0 = time.time()
        out[kind] = run(kind)
        out[kind]["wall_s"] = round(time.time() - t0, 1)
    out["speedup_sparse_over_dense_masked"] = round(out["dense_masked"]["ms_per_step"] / out["sparse"]["ms_per_step"], 3)
    out["speedup_sparse_over_dense_flash"] = round(out["dense_flash"]["ms_per_step"] / out["sparse"]["ms_per_step"], 3)
    out["arms"] = {"sparse": "ours: block-sparse flash kernels (tcgen05), fixed layout",
                   "dense_flash": "same model, fused causal attention from the library (cuDNN SDPA)",
                   "dense_masked": "same model, scores materialised + mask + softmax (the reference's dense path; "
                                   "blocks recomputed in backward, otherwise 24 x 4.3 GB of scores do not fit)"}
    if dist.get_rank() == 0:
        print(json.dumps(out), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump(out, open(f"gpurun_out/sparse_gpt2_bench_w{world}.json", "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
