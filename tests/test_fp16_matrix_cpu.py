"""Optimizer / precision / ZeRO combinations the reference exercises in tests/unit/test_fp16.py that the other CPU
files do not reach: client-built optimizers under every ZeRO stage, ZeRO-2 with `reduce_scatter: false`, the
`torch_adam` switch for Adam and AdamW, a model handed to ZeRO-3 without `zero.Init` (parameters are scattered when the
engine is built), LAMB with gradient clipping in fp32, and the fp32 empty-gradient case. Everything runs on CPU (gloo);
"fp16" blocks use `"type": "fp32"` / bf16 because CPU half matmuls are not the point here."""
import pytest
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from common import distributed_test
from simple_model import SimpleModel, random_dataloader, args_from_dict

HID = 10


def _run_steps(engine, steps=4, dtype=torch.float32):
    losses = []
    loader = random_dataloader(engine, 16 * engine.train_micro_batch_size_per_gpu(), HID, engine.device, dtype=dtype,
                               seed=dist.get_rank())
    for i, (x, y) in enumerate(loader):
        if i == steps:
            break
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        losses.append(float(loss))
    assert all(l == l for l in losses), losses        # no NaN
    return losses


@pytest.mark.parametrize("zero_stage", [1, 2, 3])
@pytest.mark.parametrize("make", ["torch_adam", "fused_adam", "cpu_adam"])
def test_zero_accepts_client_built_adam(tmp_path, zero_stage, make):
    """A client may hand in torch.optim.Adam, FusedAdam or DeepSpeedCPUAdam and still get the ZeRO wrapper
    (reference test_zero_supported_client_optimizer)."""
    conf = {"train_batch_size": 2, "steps_per_print": 100, "fp16": {"enabled": True, "type": "fp32"},
            "zero_optimization": {"stage": zero_stage}}

    @distributed_test(world_size=1)
    def run():
        from deeperspeed_b200.ops.adam import FusedAdam, DeepSpeedCPUAdam
        torch.manual_seed(0)
        model = SimpleModel(HID)
        build = {"torch_adam": lambda p: torch.optim.Adam(p, lr=1e-2), "fused_adam": lambda p: FusedAdam(p, lr=1e-2),
                 "cpu_adam": lambda p: DeepSpeedCPUAdam(p, lr=1e-2)}[make]
        client = build(model.parameters())
        before = [p.detach().clone() for p in model.parameters()]
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model, optimizer=client)
        assert engine.zero_optimization_stage() == zero_stage
        assert type(opt).__name__.lower().find("zero") >= 0 or "Stage" in type(opt).__name__ or hasattr(opt, "partition_count")
        _run_steps(engine, 3)
        if zero_stage < 3:
            assert any(not torch.equal(a, b.detach().float()) for a, b in zip(before, model.parameters()))

    run()


def test_zero2_without_reduce_scatter_matches_reduce_scatter(tmp_path):
    """`reduce_scatter: false` (all-reduce the bucket, keep the owned part) trains to the same weights as the
    reduce-scatter path (reference test_zero2_reduce_scatter_off only checks that it runs)."""
    def conf(rs):
        return {"train_batch_size": 4, "steps_per_print": 100, "gradient_clipping": 1.0,
                "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "fp16": {"enabled": True, "type": "fp32"},
                "zero_optimization": {"stage": 2, "contiguous_gradients": True, "allgather_bucket_size": 2000000000,
                                      "reduce_bucket_size": 200000000, "overlap_comm": False, "reduce_scatter": rs}}

    @distributed_test(world_size=2)
    def run():
        finals = []
        for rs in (False, True):
            torch.manual_seed(5)
            model = SimpleModel(HID)
            engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, conf(rs)), model=model,
                                                   model_parameters=model.parameters())
            _run_steps(engine, 5)
            finals.append([p.detach().float().clone() for p in model.parameters()])
        for a, b in zip(*finals):
            assert torch.allclose(a, b, atol=1e-6), (a - b).abs().max()

    run()


@pytest.mark.parametrize("adam_type, w_mode", [("Adam", None), ("Adam", False), ("AdamW", None)])
@pytest.mark.parametrize("torch_impl", [True, False])
def test_adam_types_and_torch_adam_switch(tmp_path, adam_type, w_mode, torch_impl):
    """`optimizer.params.torch_adam` picks torch.optim over the fused optimizer; "Adam" means decoupled weight decay
    unless `adam_w_mode: false`; every combination trains (reference test_fp16_adam_types, engine.py:710-735)."""
    params = {"lr": 1e-2, "torch_adam": torch_impl}
    if w_mode is not None:
        params["adam_w_mode"] = w_mode
    conf = {"train_batch_size": 1, "steps_per_print": 100,
            "fp16": {"enabled": True, "type": "fp32", "initial_scale_power": 10},
            "optimizer": {"type": adam_type, "params": params}}
    decoupled = adam_type == "AdamW" or w_mode is not False

    @distributed_test(world_size=1)
    def run():
        torch.manual_seed(0)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model,
                                               model_parameters=model.parameters())
        basic = engine.basic_optimizer
        if torch_impl:
            assert type(basic) is (torch.optim.AdamW if decoupled else torch.optim.Adam), type(basic)
        else:
            assert type(basic).__name__ == "FusedAdam", type(basic)
            assert bool(basic.adam_w_mode) == decoupled
        losses = _run_steps(engine, 8)
        assert losses[-1] < losses[0] + 0.5

    run()


def test_zero3_scatters_a_model_built_outside_init(tmp_path):
    """A plain module handed to a stage-3 engine is partitioned when the engine is built; parameters are gathered
    for each sub-module and released again (reference test_zero3_lazyscatter)."""
    conf = {"train_batch_size": 2, "steps_per_print": 100,
            "fp16": {"enabled": True, "type": "fp32", "initial_scale_power": 10},
            "optimizer": {"type": "AdamW", "params": {"lr": 1e-2}},
            "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}   # nothing stays gathered

    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        model = SimpleModel(HID)
        full = {n: p.detach().clone() for n, p in model.named_parameters()}
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model,
                                               model_parameters=model.parameters())
        for n, p in model.named_parameters():
            assert hasattr(p, "ds_id") and p.ds_numel == full[n].numel()
            assert p.numel() in (0, 1), (n, p.shape)                 # released: only the shard persists
            with deepspeed.zero.GatheredParameters(p):
                assert torch.equal(p.detach().float(), full[n])
        _run_steps(engine, 4)
        with deepspeed.zero.GatheredParameters(list(model.parameters())):
            assert any(not torch.equal(p.detach().float(), full[n]) for n, p in model.named_parameters())

    run()


def test_lamb_fp32_with_gradient_clipping(tmp_path):
    """LAMB in fp32 with gradient clipping through the plain engine path (reference test_lamb_fp32_grad_clip)."""
    conf = {"train_batch_size": 2, "steps_per_print": 100, "gradient_clipping": 1.0,
            "optimizer": {"type": "Lamb", "params": {"lr": 1e-3}}}

    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        model = SimpleModel(HID)
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model,
                                                 model_parameters=model.parameters())
        assert type(opt).__name__ == "FusedLamb"
        before = [p.detach().clone() for p in model.parameters()]
        _run_steps(engine, 4)
        assert any(not torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))
        for a, b in zip(*[[p.detach().clone() for p in model.parameters()]] * 2):
            assert torch.isfinite(a).all() and torch.equal(a, b)

    run()


@pytest.mark.parametrize("opt_type", ["Adam", "Lamb"])
def test_fp32_empty_grad_on_some_ranks(tmp_path, opt_type):
    """A layer that only rank 0 uses leaves `grad is None` elsewhere; the data-parallel reduction still lines up
    (reference test_adam_fp32_empty_grad / test_lamb_fp16_empty_grad)."""
    conf = {"train_batch_size": 2, "steps_per_print": 100,
            "optimizer": {"type": opt_type, "params": {"lr": 1e-3}}}

    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        model = SimpleModel(HID, empty_grad=True, rank=dist.get_rank())
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model,
                                               model_parameters=model.parameters())
        _run_steps(engine, 4)
        # replicas stay identical although rank 1 never produced a gradient for linear2
        for p in model.parameters():
            mine = p.detach().float().clone()
            other = mine.clone()
            dist.broadcast(other, src=0)
            assert torch.allclose(mine, other, atol=1e-7)

    run()
