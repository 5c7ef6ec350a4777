"""`bench.py --impl reference`: the UNMODIFIED reference (baseline/_ref) through its own public API
(deepspeed.initialize -> PipelineEngine.train_batch) on the same metric / config / timing protocol as our arm."""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")


def main(args, METRIC, SEQ, MICRO_BATCH, MB_PER_STAGE, topo_for, ds_config, ClockSampler, HostBatches, DeviceBatches):
    if not os.path.isdir(os.path.join(REF_ROOT, "deepspeed")):
        print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is not installed (pip install --target baseline/_ref /root/reference)"}))
        return
    os.environ.setdefault("TORCH_EXTENSIONS_DIR", os.path.join(REF_ROOT, "_torch_ext"))
    try:
        import ref_shims
        ref_shims.install(REF_ROOT)
        import torch
        import torch.distributed as dist
        if not torch.cuda.is_available():
            print(json.dumps({"impl": "reference", "unavailable": "reference hard-codes nccl + torch.cuda (engine.py:138,537-546); no GPU here"}))
            return
        import deepspeed
        assert os.path.abspath(deepspeed.__file__).startswith(os.path.abspath(REF_ROOT)), deepspeed.__file__
        from deepspeed.pipe import PipelineModule, LayerSpec
        from deepspeed.runtime.pipe.topology import PipeDataParallelTopology
        import ref_model
    except Exception as e:
        print(json.dumps({"impl": "reference", "unavailable": f"reference import failed: {e!r}"[:400]}))
        return

    n = args.gpus
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29512")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("WORLD_SIZE", str(n))
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    deepspeed.init_distributed(dist_backend="nccl")
    rank = dist.get_rank()
    pp, dp = topo_for(n)
    conf = ds_config(n, pp, dp)
    # the reference cannot JIT-build its CUDA ops for compute capability 10.0 (op_builder/builder.py:306): use its
    # own `torch_adam` switch; drop our extension key
    conf["optimizer"]["params"]["torch_adam"] = True
    conf["zero_optimization"].pop("fused_nvlink", None)
    layers = args.layers or 24
    vocab, hidden, heads = (1024, 256, 4) if args.tiny else (50432, 2048, 16)
    if args.tiny:
        layers = args.layers or 2
    torch.manual_seed(1234)
    model = PipelineModule(layers=ref_model.build_specs(LayerSpec, vocab, hidden, layers, heads),
                           topology=PipeDataParallelTopology(num_pp=pp, num_dp=dp), loss_fn=ref_model.ref_loss_fn,
                           partition_method="type:ParallelTransformerLayerPipe")
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=[p for p in model.parameters() if p.requires_grad],
                                           config_params=conf)
    device = engine.device
    tokens_per_step = conf["train_batch_size"] * SEQ
    dev_iter = DeviceBatches(vocab, device)
    host_iter = HostBatches(vocab)

    def sync():
        torch.cuda.synchronize()
        if dist.get_world_size() > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def mx(ms):
        t = torch.tensor([ms], dtype=torch.float64, device=device)
        if dist.get_world_size() > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        engine.train_batch(dev_iter)
    sync()
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", 0)))
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        engine.train_batch(dev_iter)
    e1.record()
    torch.cuda.synchronize()
    ms = mx(e0.elapsed_time(e1))
    sync()
    clocks = sampler.stop() if rank == 0 else None

    for _ in range(max(1, min(args.warmup, 2))):
        float(engine.train_batch(host_iter))
    sync()
    host_iter.bytes = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    lv = 0.0
    for _ in range(args.steps):
        lv = float(engine.train_batch(host_iter))
    e1.record()
    torch.cuda.synchronize()
    ms_e2e = mx(e0.elapsed_time(e1))
    h2d = torch.tensor([float(host_iter.bytes)], dtype=torch.float64, device=device)
    if dist.get_world_size() > 1:
        dist.all_reduce(h2d)
    sync()
    if rank == 0:
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": tokens_per_step / (ms / args.steps / 1e3), "unit": "tokens/s",
            "n_gpus": n, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp16",
            "data": "synthetic tokens, random-init weights",
            "config": {"model": "GPT-NeoX-1.3B (24L d2048 h16 seq2048 vocab50432)" if not args.tiny else "tiny",
                       "layers": layers, "global_batch": conf["train_batch_size"], "seq_len": SEQ,
                       "micro_batch": MICRO_BATCH, "grad_accum": conf["gradient_accumulation_steps"],
                       "parallelism": f"pp{pp}xdp{dp} zero1", "optimizer": "reference torch_adam path (AdamW)",
                       "l2": "working set per step >> 126 MB L2 (no flush needed)"},
            "clocks": clocks,
            "e2e": {"value": tokens_per_step / (ms_e2e / args.steps / 1e3), "unit": "tokens/s",
                    "h2d_bytes_per_step": float(h2d.item()) / args.steps, "d2h_bytes_per_step": 4.0,
                    "ms_per_step": ms_e2e / args.steps, "last_loss": lv},
            "gpu_launches": 0,
        }), flush=True)
    dist.barrier()
    dist.destroy_process_group()
