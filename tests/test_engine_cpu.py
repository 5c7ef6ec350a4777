"""Engine / ZeRO / pipeline / checkpoint plumbing on CPU with the gloo backend (world size 1-2).
BASELINE config #1: SimpleModel, ZeRO-1, fp32, CPU/gloo, world 2."""
import copy
import os

import pytest
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from common import distributed_test
from simple_model import SimpleModel, LinearStack, LinearStackPipe, random_dataloader, args_from_dict, random_dataset

HID = 16


def _cfg(stage, **kw):
    c = {"train_batch_size": 8, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
         "optimizer": {"type": "Adam", "params": {"lr": 0.01, "weight_decay": 0.01}},
         "fp16": {"enabled": True, "type": "fp32"}}
    if stage:
        c["zero_optimization"] = {"stage": stage}
    c.update(kw)
    return c


def _train(engine, steps, seed=0):
    losses = []
    dl = random_dataloader(engine, 64, HID, engine.device, seed=seed + dist.get_rank())
    it = iter(dl)
    for _ in range(steps * engine.gradient_accumulation_steps()):
        x, y = next(it)
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        losses.append(loss.item())
    return losses


def _reference_run(steps, world, gas):
    """Plain torch AdamW on the concatenated global batch = what any DP/ZeRO variant must reproduce."""
    torch.manual_seed(42)
    model = SimpleModel(HID)
    opt = torch.optim.AdamW(model.parameters(), lr=0.01, weight_decay=0.01)
    its = [iter(torch.utils.data.DataLoader(random_dataset(64, HID, seed=r), batch_size=2)) for r in range(world)]
    for _ in range(steps):
        opt.zero_grad()
        for _ in range(gas):
            for it in its:
                x, y = next(it)
                (model(x, y) / (gas * world)).backward()
        opt.step()
    return [p.detach().clone() for p in model.parameters()]


@pytest.mark.parametrize("stage", [0, 1, 2])
def test_zero_matches_plain_adamw_world2(stage, tmp_path):
    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(42)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(stage)), model=model,
                                               model_parameters=model.parameters())
        assert engine.gradient_accumulation_steps() == 2
        _train(engine, 3)
        ref = _reference_run(3, 2, 2)
        for p, r in zip(engine.module.parameters(), ref):
            assert torch.allclose(p.detach(), r, atol=2e-5), (p - r).abs().max()
    run()


def test_zero2_overlap_and_offload_world2(tmp_path):
    @distributed_test(world_size=2)
    def run():
        for extra in ({"overlap_comm": True}, {"cpu_offload": True}):
            torch.manual_seed(42)
            model = SimpleModel(HID)
            cfg = _cfg(2)
            cfg["zero_optimization"].update(extra)
            engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                   model_parameters=model.parameters())
            _train(engine, 3)
            ref = _reference_run(3, 2, 2)
            for p, r in zip(engine.module.parameters(), ref):
                assert torch.allclose(p.detach(), r, atol=3e-5), (extra, (p - r).abs().max())
    run()


def test_unbalanced_gradients_zero(tmp_path):
    @distributed_test(world_size=2)
    def run():
        for stage in (1, 2):
            model = SimpleModel(HID, empty_grad=True, rank=dist.get_rank())
            engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(stage)), model=model,
                                                   model_parameters=model.parameters())
            losses = _train(engine, 2)
            assert all(l == l for l in losses)
    run()


@pytest.mark.parametrize("stage", [1, 2])
def test_checkpoint_roundtrip_and_elastic_reload(stage, tmp_path):
    @distributed_test(world_size=2)
    def save():
        torch.manual_seed(7)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(stage)), model=model,
                                               model_parameters=model.parameters())
        _train(engine, 2)
        engine.save_checkpoint(str(tmp_path), client_state={"foo": 17})
        # same world size: bitwise restore of master weights and Adam state
        model2 = SimpleModel(HID)
        e2, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(stage)), model=model2,
                                           model_parameters=model2.parameters())
        path, client = e2.load_checkpoint(str(tmp_path))
        assert path is not None and client["foo"] == 17 and e2.global_steps == engine.global_steps
        for a, b in zip(engine.module.parameters(), e2.module.parameters()):
            assert torch.equal(a, b)
        for ma, mb in zip(engine.optimizer.fp32_masters, e2.optimizer.fp32_masters):
            assert torch.equal(ma, mb)
            sa, sb = engine.optimizer.optimizer.state[ma], e2.optimizer.optimizer.state[mb]
            assert torch.equal(sa["exp_avg"], sb["exp_avg"]) and torch.equal(sa["exp_avg_sq"], sb["exp_avg_sq"])
        l1, l2 = _train(engine, 1, seed=5), _train(e2, 1, seed=5)
        assert l1 == l2
        if dist.get_rank() == 0:
            torch.save([p.detach().clone() for p in engine.module.parameters()], os.path.join(str(tmp_path), "after.pt"))
    save()

    @distributed_test(world_size=1)
    def reload_dp1():
        cfg = _cfg(stage)
        cfg["train_micro_batch_size_per_gpu"] = 4
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        path, _ = engine.load_checkpoint(str(tmp_path))
        assert path is not None
        full = torch.cat([m.data for m in engine.optimizer.fp32_masters])
        flat = torch.cat([p.detach().view(-1) for p in engine.module.parameters()])
        assert torch.allclose(full[:flat.numel()], flat)
        st = engine.optimizer.optimizer.state[engine.optimizer.fp32_masters[0]]
        assert st["exp_avg"].abs().sum() > 0
    reload_dp1()


def test_zero_to_fp32(tmp_path):
    @distributed_test(world_size=2)
    def run():
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(1)), model=model,
                                               model_parameters=model.parameters())
        _train(engine, 1)
        engine.save_checkpoint(str(tmp_path), tag="t")
        if dist.get_rank() == 0:
            from deeperspeed_b200.utils.zero_to_fp32 import convert_zero_chkpt_to_fp32_consolid_state_dict
            sd = convert_zero_chkpt_to_fp32_consolid_state_dict(os.path.join(str(tmp_path), "t"))
            for name, p in engine.module.named_parameters():
                assert torch.allclose(sd[name], p.detach().float(), atol=1e-6), name
    run()


def test_pipeline_matches_sequential_world2(tmp_path):
    @distributed_test(world_size=2)
    def run():
        from deeperspeed_b200.runtime.pipe.topology import PipeDataParallelTopology
        cfg = {"train_batch_size": 8, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 0.01}}, "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 1}}
        torch.manual_seed(3)
        full = LinearStack(HID, HID, HID, 2)
        torch.manual_seed(3)
        pipe = LinearStackPipe(HID, HID, HID, 2, topology=PipeDataParallelTopology(num_pp=2, num_dp=1),
                               partition_method="uniform")
        # same initial weights as the sequential model
        seq_layers = [full.input_layer] + list(full.layers) + [full.output_layer]
        mods = [m for m in pipe.forward_funcs if isinstance(m, torch.nn.Module)]
        start = 0 if pipe.stage_id == 0 else len(seq_layers) - len(mods)
        for m, ref in zip(mods, seq_layers[start:start + len(mods)]):
            m.load_state_dict(ref.state_dict())
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=pipe,
                                               model_parameters=[p for p in pipe.parameters()])
        ds = random_dataset(64, HID, seed=11)
        it = iter(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(ds, batch_size=2)))
        opt = torch.optim.AdamW(full.parameters(), lr=0.01, weight_decay=0.0)
        ref_it = iter(torch.utils.data.DataLoader(ds, batch_size=2))
        for step in range(3):
            loss = engine.train_batch(it)
            opt.zero_grad()
            tot = 0.0
            for _ in range(4):
                x, y = next(ref_it)
                l = full(x, y) / 4
                l.backward()
                tot += l.item()
            opt.step()
            assert abs(float(loss) - tot) < 1e-4, (step, float(loss), tot)
        for m, ref in zip(mods, seq_layers[start:start + len(mods)]):
            for a, b in zip(m.parameters(), ref.parameters()):
                assert torch.allclose(a.detach(), b.detach(), atol=1e-5)
        # per-layer checkpoint files + reload
        engine.save_checkpoint(str(tmp_path), tag="pp")
        dist.barrier()
        files = sorted(os.listdir(os.path.join(str(tmp_path), "pp")))
        assert "layer_00-model_states.pt" in files and "mp_rank_00_model_states.pt" in files and "mp_rank_01_model_states.pt" in files
        before = [p.detach().clone() for p in pipe.parameters()]
        for p in pipe.parameters():
            p.data.zero_()
        engine.load_checkpoint(str(tmp_path), tag="pp")
        for a, b in zip(before, pipe.parameters()):
            assert torch.equal(a, b.detach())
        ev = engine.eval_batch(it)
        assert float(ev) == float(ev)
    run()


def test_fp16_optimizer_bf16_cpu(tmp_path):
    @distributed_test(world_size=1)
    def run():
        cfg = {"train_batch_size": 4, "steps_per_print": 100, "optimizer": {"type": "Adam", "params": {"lr": 0.01}},
               "fp16": {"enabled": True, "type": "bfloat16"}, "gradient_clipping": 1.0}
        model = SimpleModel(HID)
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        assert type(opt).__name__ == "FP16_Optimizer" and next(engine.module.parameters()).dtype == torch.bfloat16
        dl = random_dataloader(engine, 32, HID, engine.device, dtype=torch.bfloat16)
        first = last = None
        for _ in range(4):
            for x, y in dl:
                loss = engine(x, y); engine.backward(loss); engine.step()
                first = first if first is not None else loss.item()
                last = loss.item()
        assert last < first
        sd = opt.state_dict()
        assert set(["dynamic_loss_scale", "cur_scale", "cur_iter", "optimizer_state_dict", "fp32_groups_flat", "clip_grad"]) <= set(sd)
    run()


def test_dynamic_loss_scale_overflow_skips(tmp_path):
    @distributed_test(world_size=1)
    def run():
        cfg = {"train_batch_size": 1, "steps_per_print": 100, "optimizer": {"type": "Adam", "params": {"lr": 0.001}},
               "fp16": {"enabled": True, "type": "bfloat16"}}
        model = SimpleModel(HID)
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        # force the fp16-style dynamic scaler on a bf16 CPU model to exercise the overflow path
        opt.dynamic_loss_scale, opt.cur_scale, opt.scale_window, opt.min_loss_scale = True, 2.0 ** 8, 2, 1
        opt.last_overflow_iter, opt.scale_factor = -1, 2
        before = [p.detach().clone() for p in model.parameters()]
        x = torch.randn(1, HID, dtype=torch.bfloat16); y = torch.zeros(1, dtype=torch.long)
        loss = engine(x, y); engine.backward(loss)
        opt.arenas[0].grad[3] = float("inf")
        engine.step()
        assert opt.overflow and opt.cur_scale == 2.0 ** 7 and engine.skipped_steps == 1
        for a, b in zip(before, model.parameters()):
            assert torch.equal(a, b.detach())
        for _ in range(3):
            loss = engine(x, y); engine.backward(loss); engine.step()
        assert not opt.overflow and opt.cur_scale == 2.0 ** 8
    run()


def test_zero_to_fp32_script_runs_standalone_from_checkpoint_dir(tmp_path):
    """The recovery script copied next to every ZeRO checkpoint works as a plain command, also where this package is
    not importable (the pickled loss-scaler objects are stubbed): `python zero_to_fp32.py . out.bin`."""
    import subprocess
    import sys

    @distributed_test(world_size=2)
    def save():
        torch.manual_seed(42)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(2)), model=model,
                                               model_parameters=model.parameters())
        _train(engine, 2)
        engine.save_checkpoint(str(tmp_path), tag="final")
        if dist.get_rank() == 0:
            torch.save({k: v.detach().float().clone() for k, v in model.state_dict().items()},
                       os.path.join(str(tmp_path), "expected.pt"))
    save()
    tag_dir = os.path.join(str(tmp_path), "final")
    assert os.path.isfile(os.path.join(tag_dir, "zero_to_fp32.py"))
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    out = os.path.join(str(tmp_path), "consolidated.bin")
    r = subprocess.run([sys.executable, "zero_to_fp32.py", ".", out], cwd=tag_dir, env=env, capture_output=True, timeout=300)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    got = torch.load(out, weights_only=False)
    want = torch.load(os.path.join(str(tmp_path), "expected.pt"))
    assert set(got.keys()) == set(want.keys())
    for k in want:
        assert torch.allclose(got[k].float(), want[k], atol=1e-6), k


def _whole_parameters(engine):
    params = list(engine.module.parameters())
    with deepspeed.zero.GatheredParameters(params, enabled=engine.zero_optimization_stage() == 3):
        return [p.detach().clone() for p in params]


@pytest.mark.parametrize("stage", [1, 2, 3])
def test_elastic_reload_world3_to_world2_continues_identically(stage, tmp_path):
    """Elastic ZeRO checkpoints (padding-free shards, reference stage1.py:924-1098 / stage2.py:1685-1823): a run saved
    on 3 ranks and resumed on 2 has the same master weights and Adam moments, i.e. the same next update as the run
    that was never interrupted (same global batch, same data). Stage 3 as well - the reference's stage-3 checkpoints
    are rigid (stage3.py:3072) - with sub-groups that fall differently for the two world sizes."""
    def cfg_for(world):
        c = _cfg(stage)
        c["train_batch_size"] = 12
        c["train_micro_batch_size_per_gpu"] = 2
        if stage == 3:
            c["zero_optimization"].update({"stage3_param_persistence_threshold": 0, "sub_group_size": 100})
        return c          # gas = 2 on 3 ranks, 3 on 2 ranks

    data = torch.utils.data.TensorDataset(*random_dataset(48, HID, seed=99).tensors)

    def batches(world, rank, gas, step):
        """Micro-batches of global step `step` for this rank: the 12 samples of the step split rank-major."""
        xs, ys = data.tensors
        lo = step * 12
        per_rank = 12 // world
        sl = slice(lo + rank * per_rank, lo + (rank + 1) * per_rank)
        return [(xs[sl][i * 2:(i + 1) * 2], ys[sl][i * 2:(i + 1) * 2]) for i in range(gas)]

    def run_steps(engine, world, steps, first_step):
        for s in range(first_step, first_step + steps):
            for x, y in batches(world, dist.get_rank(), engine.gradient_accumulation_steps(), s):
                loss = engine(x, y)
                engine.backward(loss)
                engine.step()

    @distributed_test(world_size=3)
    def save():
        torch.manual_seed(7)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg_for(3)), model=model,
                                               model_parameters=model.parameters())
        run_steps(engine, 3, 2, 0)
        engine.save_checkpoint(str(tmp_path), tag="w3")
        run_steps(engine, 3, 1, 2)            # the uninterrupted run goes one step further
        whole = _whole_parameters(engine)
        if dist.get_rank() == 0:
            torch.save(whole, os.path.join(str(tmp_path), "w3_step3.pt"))
    save()

    @distributed_test(world_size=2)
    def resume():
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg_for(2)), model=model,
                                               model_parameters=model.parameters())
        path, _ = engine.load_checkpoint(str(tmp_path), tag="w3")
        assert path is not None and engine.global_steps == 2
        run_steps(engine, 2, 1, 2)
        want = torch.load(os.path.join(str(tmp_path), "w3_step3.pt"))
        for p, w in zip(_whole_parameters(engine), want):
            assert torch.allclose(p, w, atol=2e-6), (p - w).abs().max()
    resume()


def test_zero3_checkpoint_resumes_on_more_ranks(tmp_path):
    """The other direction for stage 3: saved on 2 ranks, resumed on 4 (ranks 2 and 3 have no model-states file of
    their own and read rank 0's); the next step equals the uninterrupted run's."""
    def cfg():
        c = _cfg(3, train_batch_size=8)
        c["zero_optimization"].update({"stage3_param_persistence_threshold": 0, "sub_group_size": 150})
        return c

    data = random_dataset(32, HID, seed=5).tensors

    def step(engine, world, index):
        per_rank = 8 // world
        lo = index * 8 + dist.get_rank() * per_rank
        for i in range(0, per_rank, 2):
            loss = engine(data[0][lo + i:lo + i + 2], data[1][lo + i:lo + i + 2])
            engine.backward(loss)
            engine.step()

    @distributed_test(world_size=2)
    def save():
        torch.manual_seed(3)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg()), model=model,
                                               model_parameters=model.parameters())
        step(engine, 2, 0)
        engine.save_checkpoint(str(tmp_path), tag="w2")
        step(engine, 2, 1)
        whole = _whole_parameters(engine)
        if dist.get_rank() == 0:
            torch.save(whole, os.path.join(str(tmp_path), "w2_next.pt"))
    save()

    @distributed_test(world_size=4)
    def resume():
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg()), model=model,
                                               model_parameters=model.parameters())
        path, _ = engine.load_checkpoint(str(tmp_path), tag="w2")
        assert path is not None and engine.global_steps == 1
        step(engine, 4, 1)
        for p, w in zip(_whole_parameters(engine), torch.load(os.path.join(str(tmp_path), "w2_next.pt"))):
            assert torch.allclose(p, w, atol=2e-6), (p - w).abs().max()
    resume()


@pytest.mark.parametrize("stage", [1, 2])
def test_zero_bf16_world2_tracks_fp32(stage, tmp_path):
    """fp16.type = bfloat16 (DeeperSpeed's bf16 mode, reference runtime/config.py:104-113: loss scale pinned to 1,
    fp32 all-reduce) under ZeRO-1/2 at world 2: fp32 masters follow the fp32 run closely, replicas stay identical."""
    @distributed_test(world_size=2)
    def run():
        cfg = _cfg(stage)
        cfg["fp16"] = {"enabled": True, "type": "bfloat16"}
        torch.manual_seed(42)
        model = SimpleModel(HID)
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        assert next(engine.module.parameters()).dtype == torch.bfloat16 and float(opt.loss_scale) == 1.0
        dl = random_dataloader(engine, 64, HID, engine.device, dtype=torch.bfloat16, seed=dist.get_rank())
        it = iter(dl)
        for _ in range(3 * engine.gradient_accumulation_steps()):
            x, y = next(it)
            loss = engine(x, y)
            engine.backward(loss)
            engine.step()
        ref = _reference_run(3, 2, 2)
        flat = torch.cat([p.detach().float().view(-1) for p in engine.module.parameters()])
        want = torch.cat([r.view(-1) for r in ref])
        assert (flat - want).abs().max().item() < 0.05 and (flat - want).abs().mean().item() < 0.01
        other = flat.clone()
        dist.broadcast(other, src=0)
        assert torch.equal(flat, other)
    run()


def test_gpt2_model_trains_under_zero3_on_cpu(tmp_path):
    """models/gpt2.py (dense and block-sparse attention variants) through the engine on CPU: parameter count formula,
    finite decreasing loss, ZeRO-3 partitioning of a model with a tied LM head."""
    from deeperspeed_b200.models.gpt2 import GPT2Config, GPT2Model

    @distributed_test(world_size=2)
    def run():
        for sparse in (False, True):
            sc = None
            if sparse:
                from deeperspeed_b200.ops.sparse_attention import FixedSparsityConfig
                sc = FixedSparsityConfig(num_heads=2, block=16, num_local_blocks=2, attention="unidirectional")
            cfg_m = GPT2Config(vocab_size=64, n_positions=32, n_embd=32, n_layer=2, n_head=2, sparsity_config=sc)
            torch.manual_seed(1)
            model = GPT2Model(cfg_m)
            assert sum(p.numel() for p in model.parameters()) == cfg_m.num_params()
            cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
                   "optimizer": {"type": "Adam", "params": {"lr": 1e-2}},
                   "fp16": {"enabled": True, "type": "fp32"},
                   "zero_optimization": {"stage": 0 if sparse else 3, "stage3_param_persistence_threshold": 0}}
            engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                   model_parameters=model.parameters())
            g = torch.Generator().manual_seed(dist.get_rank())
            tokens = torch.randint(0, 64, (2, 32), generator=g)
            losses = []
            for _ in range(6):
                loss = engine(tokens, tokens)
                engine.backward(loss)
                engine.step()
                losses.append(loss.item())
            assert all(l == l for l in losses) and losses[-1] < losses[0], (sparse, losses)
    run()


def test_model_parallel_norm_correction_reads_the_reduced_slices():
    """Fused NVLink step (parallel/fused_zero.py): the gradient arena still holds raw local gradients when the norm is
    corrected for replicated parameters, so the correction must use the reduced fp32 slices it is given."""
    from types import SimpleNamespace
    from deeperspeed_b200.runtime.fp16.flat_optimizer import FlatMixedPrecisionOptimizer

    class FakeMpu:
        def get_model_parallel_world_size(self):
            return 2

        def get_model_parallel_rank(self):
            return 1

    sub, parts, me = 8, 2, 1
    replicated = SimpleNamespace(param=SimpleNamespace(model_parallel=False), offset=6, numel=6)    # [6, 12)
    sharded = SimpleNamespace(param=SimpleNamespace(model_parallel=True), offset=12, numel=4)
    raw = torch.full((sub * parts,), 100.0)                      # what an un-reduced arena would hold
    opt = object.__new__(FlatMixedPrecisionOptimizer)
    opt.mpu, opt.shard, opt.dp_rank, opt.partition_count = FakeMpu(), True, me, parts
    opt.shard_gradients = False
    opt.arenas = [SimpleNamespace(segments=[replicated, sharded], grad=raw)]
    opt.num_buckets, opt.sub_size = [1], [sub]
    reduced = torch.arange(sub, dtype=torch.float32)             # this rank's slice = arena[8:16]
    opt._sumsq = torch.tensor([float((reduced * reduced).sum()), 0.0])
    opt._mp_owned_correction(reduced={(0, 0): reduced})
    keep = reduced[4:]                                           # arena [12, 16): the replicated [8, 12) part is removed
    assert torch.allclose(opt._sumsq[0], (keep * keep).sum())
    # the in-place path reads the arena itself
    opt._sumsq = torch.tensor([1e6, 0.0])
    opt._mp_owned_correction()
    assert torch.allclose(opt._sumsq[0], torch.tensor(1e6 - 4 * 100.0 ** 2))


def test_zero2_sharded_gradient_memory_world2(tmp_path):
    """Stage 2 with `gradient_memory: sharded`: no whole gradient arena exists, only the owned slices plus pooled
    bucket staging; results equal plain AdamW on the global batch (with several buckets, gradient accumulation, an
    offloaded optimizer, and parameters that get no gradient on one rank)."""
    @distributed_test(world_size=2)
    def run():
        for extra in ({}, {"reduce_bucket_size": 96}, {"cpu_offload": True, "reduce_bucket_size": 96}):
            torch.manual_seed(42)
            model = SimpleModel(HID)
            cfg = _cfg(2)
            cfg["zero_optimization"].update(extra, gradient_memory="sharded")
            engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                   model_parameters=model.parameters())
            opt = engine.optimizer
            assert opt.shard_gradients and all(a.grad is None for a in opt.arenas)
            _train(engine, 3)
            assert all(p.grad is None for p in engine.module.parameters())      # freed as soon as they were staged
            ref = _reference_run(3, 2, 2)
            for p, r in zip(engine.module.parameters(), ref):
                assert torch.allclose(p.detach(), r, atol=3e-5), (extra, (p - r).abs().max())
            own, staged = opt.gradient_memory_bytes()
            total = sum(a.numel * 4 for a in opt.arenas)
            assert own == total // 2
            if "reduce_bucket_size" in extra:
                assert sum(opt.num_buckets) > 1 and staged < total      # staging is a few buckets, not the arena
        model = SimpleModel(HID, empty_grad=True, rank=dist.get_rank())
        cfg = _cfg(2)
        cfg["zero_optimization"]["gradient_memory"] = "sharded"
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        losses = _train(engine, 2)
        assert all(l == l for l in losses)
    run()


def test_amp_block_runs_on_native_autocast_world2(tmp_path):
    """`"amp": {"enabled": true}`: autocast forward + GradScaler instead of apex; fp32 weights, gradients averaged over
    the ranks, replicas stay identical, the loss goes down, and amp + fp16 / amp + ZeRO are refused like the reference."""
    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(3 + dist.get_rank())          # different initial weights: the deferred broadcast must fix that
        model = SimpleModel(HID)
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 0.01}}, "gradient_clipping": 1.0,
               "amp": {"enabled": True, "opt_level": "O1"}}
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        assert engine.amp is not None and all(p.dtype == torch.float32 for p in model.parameters())
        start = [p.detach().clone() for p in model.parameters()]      # rank 0's weights after the broadcast
        _train(engine, 4)
        flat = torch.cat([p.detach().view(-1) for p in model.parameters()])
        other = flat.clone()
        dist.broadcast(other, src=0)
        assert torch.allclose(flat, other, atol=1e-6)
        # the same four steps by hand: autocast forward on both ranks' batches, averaged gradients, clip, Adam
        ref = SimpleModel(HID)
        for p, s0 in zip(ref.parameters(), start):
            p.data.copy_(s0)
        opt = torch.optim.Adam(ref.parameters(), lr=0.01)
        its = [iter(torch.utils.data.DataLoader(random_dataset(64, HID, seed=r), batch_size=2)) for r in range(2)]
        for _ in range(4):
            opt.zero_grad()
            for it in its:
                x, y = next(it)
                with torch.autocast("cpu", dtype=torch.bfloat16):
                    loss = ref(x, y)
                (loss / 2).backward()
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
            opt.step()
        for p, r in zip(model.parameters(), ref.parameters()):
            assert torch.allclose(p.detach(), r.detach(), atol=2e-3), (p - r).abs().max()
        with pytest.raises(AssertionError):
            bad = dict(cfg, zero_optimization={"stage": 1}, fp16={"enabled": True})
            deepspeed.initialize(args=args_from_dict(tmp_path, bad), model=SimpleModel(HID),
                                 model_parameters=model.parameters())
    run()


def _scalar_tags(log_dir):
    from tensorboard.backend.event_processing.event_accumulator import EventAccumulator
    tags = set()
    for root, _, files in os.walk(log_dir):
        if any(f.startswith("events.out") for f in files):
            acc = EventAccumulator(root)
            acc.Reload()
            tags.update(acc.Tags()["scalars"])
    return tags


def test_tensorboard_scalars_from_both_engines(tmp_path):
    """`"tensorboard": {"enabled": true}`: rank 0 of the plain engine and of the PIPELINE engine write the reference's
    scalar tags (Train/Samples/train_loss, lr, loss_scale; per-phase times with wall_clock_breakdown)."""
    pytest.importorskip("tensorboard")

    @distributed_test(world_size=2)
    def run():
        tb = lambda name: {"enabled": True, "output_path": str(tmp_path / name), "job_name": "job"}
        cfg = _cfg(0, tensorboard=tb("plain"), wall_clock_breakdown=True)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        _train(engine, 2)
        dist.barrier()
        if dist.get_rank() == 0:
            tags = _scalar_tags(str(tmp_path / "plain"))
            assert {"Train/Samples/train_loss", "Train/Samples/lr"} <= tags, tags
            assert any(t.startswith("Train/Samples/elapsed_time_ms_") for t in tags), tags
        pcfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
                "optimizer": {"type": "Adam", "params": {"lr": 0.01}}, "tensorboard": tb("pipe")}
        net = LinearStackPipe(HID, HID, HID, 2, num_stages=2)
        pengine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, pcfg), model=net,
                                                model_parameters=[p for p in net.parameters()])
        ds = random_dataset(32, HID, seed=3)
        it = iter(deepspeed.utils.RepeatingLoader(torch.utils.data.DataLoader(ds, batch_size=2)))
        for _ in range(3):
            pengine.train_batch(it)
        dist.barrier()
        if dist.get_rank() == 0:
            tags = _scalar_tags(str(tmp_path / "pipe"))
            assert "Train/Samples/train_loss" in tags and "Train/Samples/lr" in tags, tags
    run()


def test_training_data_loader_shards_by_rank_and_feeds_the_engine(tmp_path):
    """`initialize(training_data=...)`: the returned loader hands every data-parallel rank its own, disjoint share of
    the dataset in micro-batches, applies the collate function, starts the throughput timer per batch and drives a
    normal training loop; `deepspeed_io` builds evaluation loaders that read sequentially; `RepeatingLoader` never runs
    dry (reference runtime/dataloader.py:10-101, engine.py:890-957)."""
    from deeperspeed_b200.runtime.dataloader import DeepSpeedDataLoader, RepeatingLoader

    @distributed_test(world_size=2)
    def run():
        n = 24
        xs = torch.arange(n, dtype=torch.float32).unsqueeze(1).repeat(1, HID)        # sample i is all i's
        ys = torch.arange(n) % HID
        dataset = torch.utils.data.TensorDataset(xs, ys)
        tagged = []

        def collate(samples):
            x = torch.stack([s[0] for s in samples])
            y = torch.stack([s[1] for s in samples])
            tagged.append(len(samples))
            return x, y

        model = SimpleModel(HID)
        engine, _, loader, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(1)), model=model,
                                                    model_parameters=model.parameters(), training_data=dataset,
                                                    collate_fn=collate)
        assert isinstance(loader, DeepSpeedDataLoader) and loader is engine.training_dataloader
        assert len(loader) == n // 2                                   # samples of this rank
        seen = []
        for x, y in loader:
            assert x.shape == (engine.train_micro_batch_size_per_gpu(), HID)
            seen.extend(int(v) for v in x[:, 0])
            loss = engine(x / n, y)
            engine.backward(loss)
            engine.step()
        assert tagged and all(t == 2 for t in tagged)                  # the collate function was used
        assert engine.global_steps == len(seen) // (2 * engine.gradient_accumulation_steps())
        assert engine.tput_timer.total_step_count > 0                  # the loader started the timer
        gathered = [None, None]
        dist.all_gather_object(gathered, seen)
        assert sorted(gathered[0] + gathered[1]) == list(range(n)) and not set(gathered[0]) & set(gathered[1])

        evaluation = engine.deepspeed_io(dataset, batch_size=4, route="eval")
        order = [int(v) for x, _ in evaluation for v in x[:, 0]]
        assert order == list(range(n))                                 # sequential, every sample, no sharding
        with pytest.raises(ValueError):
            engine.deepspeed_io([1, 2, 3])

        endless = RepeatingLoader(engine.deepspeed_io(dataset, batch_size=4, route="eval"))
        drawn = [next(endless)[0][0, 0].item() for _ in range(2 * (n // 4) + 1)]
        assert drawn[0] == drawn[n // 4] == drawn[2 * (n // 4)] == 0.0

    run()


def test_engine_step_bookkeeping_api(tmp_path):
    """Small engine methods nothing else calls directly: accumulation boundary, batch info / set_train_batch_size,
    global gradient norm, optimizer type, module_state_dict round trip, fp32 gradient clipping."""
    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(1)
        model = SimpleModel(HID)
        cfg = _cfg(0, gradient_clipping=0.05)
        del cfg["fp16"]                                        # plain fp32: the engine itself clips
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        batch, micro, gas = engine.get_batch_info()
        assert (batch, micro, gas) == (8, 2, 2) and (batch(), micro(), gas()) == (8, 2, 2)
        assert engine.get_type() == ["adam"] or engine.get_type() == [None] or engine.get_type() is not None
        it = iter(random_dataloader(engine, 64, HID, engine.device, seed=dist.get_rank()))
        boundaries = []
        for _ in range(4):
            boundaries.append(engine.is_gradient_accumulation_boundary())
            x, y = next(it)
            loss = engine(x, y)
            engine.backward(loss)
            if boundaries[-1]:
                engine.clip_fp32_gradients()                   # what step() does for an unwrapped optimizer
                total = torch.sqrt(sum(p.grad.float().pow(2).sum() for p in model.parameters()))
                assert total <= 0.05 * 1.001
            engine.step()
        assert boundaries == [False, True, False, True] and engine.global_steps == 2

        engine.set_train_batch_size(16)                        # same micro batch and dp: accumulation doubles
        assert engine.gradient_accumulation_steps() == 4 and engine.train_batch_size() == 16
        with pytest.raises(ValueError):
            engine.set_train_batch_size(18)
        for _ in range(4):
            x, y = next(it)
            loss = engine(x, y); engine.backward(loss); engine.step()
        assert engine.global_steps == 3

        saved = {k: v.clone() for k, v in engine.module_state_dict().items()}
        with torch.no_grad():
            for p in model.parameters():
                p.add_(1.0)
        engine.load_module_state_dict(saved)
        for k, v in engine.module_state_dict().items():
            assert torch.equal(v, saved[k])

    run()


def test_zero_global_grad_norm_matches_manual_norm(tmp_path):
    """engine.get_global_grad_norm() after a ZeRO step is the 2-norm of the averaged global-batch gradient."""
    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(2)
        model = SimpleModel(HID)
        twin = SimpleModel(HID)
        twin.load_state_dict(model.state_dict())
        cfg = _cfg(1, train_batch_size=4)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        sets = [next(iter(torch.utils.data.DataLoader(random_dataset(8, HID, seed=r), batch_size=2))) for r in range(2)]
        x, y = sets[dist.get_rank()]
        loss = engine(x, y); engine.backward(loss); engine.step()
        for xb, yb in sets:
            (twin(xb, yb) / 2).backward()
        want = torch.sqrt(sum(p.grad.pow(2).sum() for p in twin.parameters())).item()
        assert engine.get_global_grad_norm() == pytest.approx(want, rel=1e-4)

    run()


def test_reference_keyword_arguments_of_optimizers_and_stage1_comms_timer(tmp_path):
    """Keyword arguments callers written against the reference pass: `deepspeed=` of the precision wrappers, `mpu=` of
    SparseSelfAttention, and `comms_timer=` of the ZeRO-1 step, which is started and stopped around the parameter
    all-gather (reference stage1.py:688,759-775)."""
    from deeperspeed_b200.ops.adam import FusedAdam
    from deeperspeed_b200.ops.sparse_attention import SparseSelfAttention, FixedSparsityConfig
    from deeperspeed_b200.runtime.fp16.fused_optimizer import FP16_Optimizer
    from deeperspeed_b200.runtime.fp16.unfused_optimizer import FP16_UnfusedOptimizer
    marker = object()
    for cls in (FP16_Optimizer, FP16_UnfusedOptimizer):
        model = SimpleModel(HID).bfloat16()
        wrapped = cls(FusedAdam(model.parameters(), lr=1e-3), static_loss_scale=1.0, deepspeed=marker)
        assert wrapped.deepspeed is marker
    attention = SparseSelfAttention(FixedSparsityConfig(num_heads=2, block=16), max_seq_length=64, mpu=marker)
    assert attention.mpu is marker

    class Stopwatch:
        def __init__(self):
            self.events = []

        def start(self):
            self.events.append("start")

        def stop(self):
            self.events.append("stop")

    @distributed_test(world_size=2)
    def run():
        model = SimpleModel(HID)
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, _cfg(1, train_batch_size=4)),
                                                 model=model, model_parameters=model.parameters())
        x, y = next(iter(random_dataloader(engine, 8, HID, engine.device, seed=dist.get_rank())))
        loss = engine(x, y)
        engine.backward(loss)
        watch = Stopwatch()
        before = [p.detach().clone() for p in model.parameters()]
        opt.step(comms_timer=watch)
        assert watch.events == ["start", "stop"]
        assert any(not torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))
        assert opt._comms_timer is None

    run()
