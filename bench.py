#!/usr/bin/env python
"""Headline benchmark: tokens/sec of GPT-NeoX-1.3B (24L, d=2048, seq 2048, fp16) with ZeRO-1 + 1F1B pipeline.

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched by torch.distributed.run, one rank / GPU)
  python bench.py --impl reference ...                      (unmodified reference from baseline/_ref)

Topologies: 1 GPU -> (pp 1, dp 1); 2 -> (2, 1); 4 -> (2, 2); 8 -> (2, 4). Weak scaling: micro-batch 4 x 2048 tokens,
8 micro-batches per pipeline stage per step. Synthetic tokens, random-init weights. `value` is timed on the device
with CUDA events over exactly K steps (max over ranks); `e2e` times the same public-API call with every step's
inputs copied from pinned host memory and the loss read back to the host."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "tokens/sec GPT-NeoX-1.3B ZeRO-1+PP"
SEQ = 2048
MICRO_BATCH = 4
MICRO_BATCHES_PER_STAGE = 8


def topo_for(n):
    return {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (2, 4)}.get(n, (2 if n % 2 == 0 else 1, max(n // 2, 1) if n % 2 == 0 else n))


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], None, set()
        for s in self.samples:
            f = [x.strip() for x in s.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def _attention_backend():
    try:
        from deeperspeed_b200.ops.attention import attention_backend
        return {"native": "tcgen05 flash attention (ours)", "sdpa": "library SDPA"}[attention_backend()]
    except Exception:
        return "unknown"


def _zero_backend(engine):
    opt = getattr(engine, "optimizer", None)
    fused = getattr(opt, "_fused", None) or getattr(opt, "fused", None)
    if fused is not None:
        how = "NVLS multimem.ld_reduce / multimem.st" if getattr(fused, "use_multicast", False) else "peer loads / stores"
        return f"fused_zero_step kernel (ours): reduce-scatter + norm + Adam + all-gather over the symmetric heap, {how}"
    return "NCCL reduce_scatter / all_gather + fused Adam kernel" if getattr(engine, "dp_world_size", 1) > 1 else \
        "single rank: fused Adam kernel, no collective"


def ds_config(n_gpus, pp, dp):
    gas = MICRO_BATCHES_PER_STAGE * pp
    return {
        "train_batch_size": MICRO_BATCH * gas * dp,
        "train_micro_batch_size_per_gpu": MICRO_BATCH,
        "gradient_accumulation_steps": gas,
        "steps_per_print": 1000000,
        "optimizer": {"type": "Adam", "params": {"lr": 1e-4, "betas": [0.9, 0.95], "eps": 1e-8, "weight_decay": 0.01}},
        "fp16": {"enabled": True, "type": "fp16", "loss_scale": 0, "initial_scale_power": 12, "loss_scale_window": 1000,
                 "hysteresis": 2, "min_loss_scale": 1},
        "zero_optimization": {"stage": 1, "reduce_scatter": True, "reduce_bucket_size": 500000000},
        "gradient_clipping": 1.0,
        "wall_clock_breakdown": False,
    }


class HostBatches:
    """Endless iterator of (tokens, labels) micro-batches living in pinned host memory (fresh buffer each time)."""

    def __init__(self, vocab, pool=16):
        import torch
        g = torch.Generator().manual_seed(1234 + int(os.environ.get("RANK", 0)))
        self.pool = []
        for _ in range(pool):
            t = torch.randint(0, vocab, (MICRO_BATCH, SEQ + 1), generator=g)
            x, y = t[:, :-1].contiguous(), t[:, 1:].contiguous()
            if torch.cuda.is_available():
                x, y = x.pin_memory(), y.pin_memory()
            self.pool.append((x, y))
        self.i = 0
        self.bytes = 0

    def __iter__(self):
        return self

    def __next__(self):
        x, y = self.pool[self.i % len(self.pool)]
        self.i += 1
        self.bytes += x.numel() * x.element_size() + y.numel() * y.element_size()
        return x, y


class DeviceBatches:
    def __init__(self, vocab, device, pool=4):
        import torch
        g = torch.Generator().manual_seed(99 + int(os.environ.get("RANK", 0)))
        self.pool = []
        for _ in range(pool):
            t = torch.randint(0, vocab, (MICRO_BATCH, SEQ + 1), generator=g)
            self.pool.append((t[:, :-1].contiguous().to(device), t[:, 1:].contiguous().to(device)))
        self.i = 0

    def __iter__(self):
        return self

    def __next__(self):
        self.i += 1
        return self.pool[self.i % len(self.pool)]


def _max_over_ranks(ms, device):
    import torch
    import torch.distributed as dist
    t = torch.tensor([ms], dtype=torch.float64, device=device)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def run_ours(args):
    import torch
    import torch.distributed as dist
    import deeperspeed_b200 as deepspeed
    from deeperspeed_b200.models.gpt_neox import GPTNeoXPipe, NeoXConfig
    from deeperspeed_b200.ops import kernels as K
    from deeperspeed_b200.runtime.pipe.topology import PipeDataParallelTopology

    n = args.gpus
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    os.environ.setdefault("RANK", "0")
    os.environ.setdefault("LOCAL_RANK", "0")
    os.environ.setdefault("WORLD_SIZE", str(n))
    assert int(os.environ["WORLD_SIZE"]) == n, "launch with torch.distributed.run --nproc-per-node N for --gpus N"
    deepspeed.init_distributed(verbose=False)
    rank = dist.get_rank()
    pp, dp = topo_for(n)
    cfg = NeoXConfig.tiny() if args.tiny else NeoXConfig.neox_1p3b()
    if args.layers:
        cfg.num_layers = args.layers
    torch.manual_seed(1234)
    if args.impl == "engine-ref-model":
        # OUR engine (pipeline executor, ZeRO-1 arenas, fused step) around the reference arm's eager PyTorch model
        # (baseline/ref_model.py): separates the engine's share of the speed-up from the model kernels'
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import ref_model
        from deeperspeed_b200.pipe import PipelineModule, LayerSpec
        vocab, hidden, heads = (1024, 256, 4) if args.tiny else (cfg.vocab_size, 2048, 16)
        model = PipelineModule(layers=ref_model.build_specs(LayerSpec, vocab, hidden, cfg.num_layers, heads),
                               topology=PipeDataParallelTopology(num_pp=pp, num_dp=dp), loss_fn=ref_model.ref_loss_fn,
                               partition_method="type:ParallelTransformerLayerPipe")
    else:
        model = GPTNeoXPipe(cfg, topology=PipeDataParallelTopology(num_pp=pp, num_dp=dp))
    conf = ds_config(n, pp, dp)
    if pp > 1 and args.edge_split > 1:
        # our engine's bubble trimming: the first and last micro-batch are fed as slices (same samples, same gradient)
        conf["pipeline"] = {"edge_split": args.edge_split}
    engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=[p for p in model.parameters() if p.requires_grad],
                                           config_params=conf)
    device = engine.device
    tokens_per_step = conf["train_batch_size"] * SEQ
    global SEQ_USED
    dev_iter = DeviceBatches(cfg.vocab_size, device)
    host_iter = HostBatches(cfg.vocab_size)

    def sync():
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        if dist.get_world_size() > 1:
            dist.barrier()
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    # one-off setup (outside every timed region): pick the GEMM kernel variant per layer shape, as the reference does
    # with its cuBLAS algorithm search when a transformer layer is built
    if torch.cuda.is_available() and not args.no_autotune:
        engine.autotune_gemms(MICRO_BATCH * SEQ, verbose=(rank == 0 and args.verbose))

    # ---- device-timed arm
    for _ in range(args.warmup):
        engine.train_batch(dev_iter)
    sync()
    sampler = ClockSampler(int(os.environ.get("LOCAL_RANK", 0)))
    if rank == 0:
        sampler.start()
    launches0 = K.launch_count()
    if torch.cuda.is_available():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.time()
    for _ in range(args.steps):
        engine.train_batch(dev_iter)
    if torch.cuda.is_available():
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.time() - t0) * 1e3
    launches = K.launch_count() - launches0
    sync()
    clocks = sampler.stop() if rank == 0 else None
    ms = _max_over_ranks(ms, device)
    ms_per_step = ms / args.steps

    # ---- end-to-end arm: pinned-host inputs every step + loss read back
    for _ in range(max(1, min(args.warmup, 2))):
        float(engine.train_batch(host_iter))
    sync()
    host_iter.bytes = 0
    loss_bytes = 0
    if torch.cuda.is_available():
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.time()
    for _ in range(args.steps):
        loss = engine.train_batch(host_iter)
        lv = float(loss)            # device -> host read of the step result
        loss_bytes += 4
    if torch.cuda.is_available():
        e1.record()
        torch.cuda.synchronize()
        ms_e2e = e0.elapsed_time(e1)
    else:
        ms_e2e = (time.time() - t0) * 1e3
    ms_e2e = _max_over_ranks(ms_e2e, device)
    h2d = torch.tensor([float(host_iter.bytes)], dtype=torch.float64, device=device)
    if dist.get_world_size() > 1:
        dist.all_reduce(h2d)
    sync()

    # ---- exposed communication: the same device-timed steps with the bulk transfers skipped (utils/comm_stub.py);
    # real - stubbed = transfer time (and the waiting it causes) the real step does not hide behind compute
    exposed = None
    if dist.get_world_size() > 1 and not args.no_exposed_comm:
        from deeperspeed_b200.utils import comm_stub
        comm_stub.set_enabled(True)
        for _ in range(max(1, min(args.warmup, 2))):
            engine.train_batch(dev_iter)
        sync()
        if torch.cuda.is_available():
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
        t0 = time.time()
        for _ in range(args.steps):
            engine.train_batch(dev_iter)
        if torch.cuda.is_available():
            e1.record()
            torch.cuda.synchronize()
            ms_stub = e0.elapsed_time(e1)
        else:
            ms_stub = (time.time() - t0) * 1e3
        comm_stub.set_enabled(False)
        sync()
        ms_stub = _max_over_ranks(ms_stub, device) / args.steps
        exposed = {"ms_per_step": ms_per_step - ms_stub, "stubbed_ms_per_step": ms_stub,
                   "how": "device-timed step minus the same step with pipeline p2p and ZeRO reduce-scatter / "
                          "all-gather skipped (max over ranks both)"}
    if rank == 0:
        value = tokens_per_step / (ms_per_step / 1e3)
        e2e_value = tokens_per_step / (ms_e2e / args.steps / 1e3)
        out = {
            "metric": METRIC, "value": value, "unit": "tokens/s", "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "fp16", "data": "synthetic tokens, random-init weights", "impl": args.impl,
            "config": {"model": "GPT-NeoX-1.3B (24L d2048 h16 seq2048 vocab50432)" if not args.tiny else "tiny",
                       "layers": cfg.num_layers, "global_batch": conf["train_batch_size"], "seq_len": SEQ,
                       "micro_batch": MICRO_BATCH, "grad_accum": conf["gradient_accumulation_steps"],
                       "parallelism": f"pp{pp}xdp{dp} zero1", "l2": "working set per step >> 126 MB L2 (no flush needed)",
                       "params": cfg.num_params(), "attention_backend": _attention_backend() if args.impl == "ours" else
                       "baseline/ref_model.py: eager PyTorch layers, library SDPA",
                       "pipeline_edge_split": getattr(engine, "edge_split", 1),
                       "zero_step_backend": _zero_backend(engine)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "tokens/s", "h2d_bytes_per_step": float(h2d.item()) / args.steps,
                    "d2h_bytes_per_step": loss_bytes / args.steps, "ms_per_step": ms_e2e / args.steps,
                    "last_loss": lv},
            "gpu_launches": launches,
            "zero_step_launches": dict(getattr(getattr(engine.optimizer, "fused", None), "launches", {}) or {}),
            "exposed_comm": exposed,
        }
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


def run_reference(args):
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    try:
        import ref_bench
    except Exception as e:  # pragma: no cover
        print(json.dumps({"impl": "reference", "unavailable": f"baseline harness import failed: {e!r}"[:300]}))
        return
    ref_bench.main(args, METRIC, SEQ, MICRO_BATCH, MICRO_BATCHES_PER_STAGE, topo_for, ds_config, ClockSampler,
                   HostBatches, DeviceBatches)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-exposed-comm", dest="no_exposed_comm", action="store_true",
                    help="skip the extra stubbed-communication pass that measures exposed comm per step (N > 1)")
    ap.add_argument("--no-autotune", dest="no_autotune", action="store_true", help="skip the GEMM variant autotune")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "engine-ref-model"],
                    help="ours; the unmodified reference; or our engine around the reference arm's eager model")
    ap.add_argument("--tiny", action="store_true", help="tiny model (CPU plumbing check only; not a valid bench)")
    ap.add_argument("--layers", type=int, default=0, help="debug override (invalidates the number)")
    ap.add_argument("--edge-split", dest="edge_split", type=int, default=1,
                    help="slices of the first / last micro-batch under pipeline parallelism (1 = plain 1F1B, the default: "
                         "on a power-capped B200 the trimmed bubble is paid back in clocks, profiles/README.md)")
    ap.add_argument("--local_rank", type=int, default=0)
    args = ap.parse_args()
    if args.tiny and os.environ.get("DSB_BENCH_TINY_SEQ"):
        # CPU plumbing runs only (tests/test_bench_contract_cpu.py): a shorter sequence for the tiny model
        global SEQ
        SEQ = int(os.environ["DSB_BENCH_TINY_SEQ"])
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
