"""ZeRO-3 on CPU/gloo: zero.Init scatter, GatheredParameters, external params, TiledLinear, stage-3 training parity."""
import os
import pytest
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.runtime.zero.partition_parameters import ZeroParamStatus
from common import distributed_test
from simple_model import SimpleModel, random_dataset, args_from_dict

HID = 16


def test_zero_init_scatter_gather():
    @distributed_test(world_size=2)
    def run():
        with deepspeed.zero.Init():
            lin = torch.nn.Linear(10, 7)
        assert lin.weight.ds_status == ZeroParamStatus.NOT_AVAILABLE and lin.weight.numel() == 1
        assert lin.weight.ds_tensor.numel() == 35 and lin.bias.ds_tensor.numel() == 4
        with deepspeed.zero.GatheredParameters(lin.weight, modifier_rank=0):
            assert lin.weight.shape == (7, 10)
            if dist.get_rank() == 0:
                lin.weight.data.fill_(3.0)
        assert lin.weight.ds_status == ZeroParamStatus.NOT_AVAILABLE
        with deepspeed.zero.GatheredParameters(lin.weight):
            assert torch.all(lin.weight == 3.0)
    run()


def test_tiled_linear_matches_linear():
    torch.manual_seed(0)
    base = torch.nn.Linear(24, 20)
    tiled = deepspeed.zero.TiledLinear(24, 20, in_splits=3, out_splits=2, init_linear=base)
    x = torch.randn(5, 24, requires_grad=True)
    x2 = x.detach().clone().requires_grad_()
    y, y2 = base(x), tiled(x2)
    assert torch.allclose(y, y2, atol=1e-5)
    y.sum().backward(); y2.sum().backward()
    assert torch.allclose(x.grad, x2.grad, atol=1e-5)
    with pytest.raises(RuntimeError):
        deepspeed.zero.TiledLinear(8, 8, in_splits=9)


@pytest.mark.parametrize("offload", [False, True, "params"])
def test_stage3_matches_adamw(offload, tmp_path):
    @distributed_test(world_size=2)
    def run():
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 0.01, "weight_decay": 0.01}},
               "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 20}}
        if offload:
            cfg["zero_optimization"]["offload_optimizer"] = {"device": "cpu"}
        if offload == "params":   # parameter shards on the host as well (ZeRO-3 Offload)
            cfg["zero_optimization"]["offload_param"] = {"device": "cpu", "pin_memory": False}
        torch.manual_seed(42)
        model = SimpleModel(HID)
        ref = SimpleModel(HID)
        ref.load_state_dict(model.state_dict())
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        assert type(opt).__name__ == "FP16_DeepSpeedZeroOptimizer_Stage3"
        ropt = torch.optim.AdamW(ref.parameters(), lr=0.01, weight_decay=0.01)
        its = [iter(torch.utils.data.DataLoader(random_dataset(64, HID, seed=r), batch_size=2)) for r in range(2)]
        for _ in range(3):
            x, y = next(its[dist.get_rank()])
            loss = engine(x, y); engine.backward(loss); engine.step()
            ropt.zero_grad()
            its2 = its  # the reference consumes both ranks' batches
            for r in range(2):
                xb, yb = (x, y) if r == dist.get_rank() else next(its[r])
                (ref(xb, yb) / 2).backward()
            ropt.step()
        with deepspeed.zero.GatheredParameters(list(model.parameters())):
            for p, r in zip(model.parameters(), ref.parameters()):
                assert torch.allclose(p.detach(), r.detach(), atol=3e-5), (p - r).abs().max()
        # checkpoint round trip
        engine.save_checkpoint(str(tmp_path), tag="z3")
        path, _ = engine.load_checkpoint(str(tmp_path), tag="z3")
        assert path is not None
    run()


@pytest.mark.parametrize("sub_group", [150, 40])      # one sub-group per rank / several
@pytest.mark.parametrize("pipeline", [False, True])
def test_stage3_nvme_offload_matches_adamw(pipeline, sub_group, tmp_path):
    """ZeRO-Infinity on files: optimizer state (and fp16 parameter shards) live under nvme_path and stream through the
    async I/O engine sub-group by sub-group (reference swap_tensor/*, stage3.py:1569-1700); result equals AdamW.
    (The single-sub-group pipelined case once read a state file whose asynchronous write was still in flight.)"""
    @distributed_test(world_size=2)
    def run():
        nvme = os.path.join(str(tmp_path), "nvme")
        os.makedirs(nvme, exist_ok=True)
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 0.01, "weight_decay": 0.01}},
               "fp16": {"enabled": True, "type": "fp32"},
               "aio": {"block_size": 65536, "queue_depth": 4, "thread_count": 1, "single_submit": False,
                       "overlap_events": True},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 20, "sub_group_size": sub_group,
                                     "offload_optimizer": {"device": "nvme", "nvme_path": nvme, "buffer_count": 4,
                                                           "pipeline_read": pipeline, "pipeline_write": pipeline},
                                     "offload_param": {"device": "nvme", "nvme_path": nvme, "buffer_count": 5,
                                                       "buffer_size": 4096}}}
        torch.manual_seed(42)
        model = SimpleModel(HID)
        ref = SimpleModel(HID)
        ref.load_state_dict(model.state_dict())
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        ropt = torch.optim.AdamW(ref.parameters(), lr=0.01, weight_decay=0.01)
        its = [iter(torch.utils.data.DataLoader(random_dataset(64, HID, seed=r), batch_size=2)) for r in range(2)]
        for _ in range(3):
            x, y = next(its[dist.get_rank()])
            loss = engine(x, y); engine.backward(loss); engine.step()
            ropt.zero_grad()
            for r in range(2):
                xb, yb = (x, y) if r == dist.get_rank() else next(its[r])
                (ref(xb, yb) / 2).backward()
            ropt.step()
        with deepspeed.zero.GatheredParameters(list(model.parameters())):
            for p, r in zip(model.parameters(), ref.parameters()):
                assert torch.allclose(p.detach(), r.detach(), atol=3e-5), (p - r).abs().max()
        # something really went through the files
        swapped = [os.path.join(dp, f) for dp, _, fs in os.walk(nvme) for f in fs]
        assert len(swapped) > 0 and sum(os.path.getsize(f) for f in swapped) > 0, swapped
    run()


class _ReturnsBias(torch.nn.Linear):
    """The skip-bias-add idiom: the caller adds the bias itself."""

    def forward(self, x):
        return super().forward(x), self.bias


class _Result:
    def __init__(self, out, bias):
        self.out, self.bias = out, bias


class _ReturnsBiasInObject(_ReturnsBias):
    def forward(self, x):
        return _Result(*super().forward(x))


class _ReturnsBiasInDict(_ReturnsBias):
    def forward(self, x):
        out, bias = super().forward(x)
        return {"out": out, "bias": bias}


class _Middle(torch.nn.Linear):
    """Passes the inner module's bias further up, so the parameter must end up external two levels above its owner."""

    def __init__(self, dim, as_object):
        super().__init__(dim, dim)
        kinds = {False: _ReturnsBias, True: _ReturnsBiasInObject, "dict": _ReturnsBiasInDict}
        self.inner = kinds[as_object](dim, dim)
        self.as_object = as_object

    def forward(self, x):
        got = self.inner(super().forward(x))
        if self.as_object == "dict":
            return got["out"], got["bias"]
        return (got.out, got.bias) if self.as_object else got


class _Outer(torch.nn.Module):
    def __init__(self, dim, as_object):
        super().__init__()
        self.first = torch.nn.Linear(dim, dim)
        self.middle = _Middle(dim, as_object)

    def forward(self, x):
        out, bias = self.middle(self.first(x))
        assert bias.numel() == self.first.in_features          # still whole although its owner has finished
        return (out + bias).pow(2).sum()


class _Root(torch.nn.Module):
    def __init__(self, dim=HID, as_object=False):
        super().__init__()
        self.outer = _Outer(dim, as_object)

    def forward(self, x):
        return self.outer(x)


@pytest.mark.parametrize("as_object", [False, True, "dict"])
def test_stage3_parameter_returned_by_a_module_stays_gathered_for_the_caller(as_object):
    """Reference tests/unit/test_zero_context.py test_ext_param_return(obj): a parameter handed back in a module's
    output is registered as external to the module that receives it (at the right level of the call stack, once),
    survives its owner's release, trains like the unpartitioned model, and everything is released again afterwards.
    Also under no_grad, where nothing but this bookkeeping keeps the tensor alive."""
    import copy

    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        net = _Root(as_object=as_object)
        plain = copy.deepcopy(net)
        plain_opt = torch.optim.Adam(plain.parameters(), lr=1e-2)
        cfg = {"train_batch_size": 2, "train_micro_batch_size_per_gpu": 1, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "adam_w_mode": False}},
               "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
        engine, _, _, _ = deepspeed.initialize(model=net, model_parameters=net.parameters(), config_params=cfg)
        g = torch.Generator().manual_seed(1)
        xs = [torch.randn(2, HID, generator=g) for _ in range(4)]      # row r belongs to rank r
        with torch.no_grad():
            assert torch.allclose(engine(xs[0][dist.get_rank()]), plain(xs[0][dist.get_rank()]), atol=1e-5)
        for x in xs:
            loss = engine(x[dist.get_rank()])
            engine.backward(loss)
            engine.step()
            plain_opt.zero_grad()
            (0.5 * (plain(x[0]) + plain(x[1]))).backward()
            plain_opt.step()
            held = {name: len(getattr(m, "_external_params", {})) for name, m in net.named_modules()}
            assert held["outer"] == 1 and sum(held.values()) == 1, held
        for (name, p), q in zip(net.named_parameters(), plain.parameters()):
            assert p.ds_status == ZeroParamStatus.NOT_AVAILABLE and p.ds_active_sub_modules == 0, name
            with deepspeed.zero.GatheredParameters(p):
                assert torch.allclose(p.detach(), q.detach(), atol=1e-5), (name, (p - q).abs().max())

    run()


# ------------------------------------------------------------------------------------------------- TiledLinear grid
def _tile_slices(tiled, row, col):
    return (slice(tiled.out_parts[row], tiled.out_parts[row + 1]), slice(tiled.in_parts[col], tiled.in_parts[col + 1]))


@pytest.mark.parametrize("splits", [1, 2, 5, 32])
def test_tiled_linear_tiles_are_the_blocks_of_the_full_weight(splits):
    """reference tests/unit/test_zero_tiled.py test_tiled_init / test_tiled_baddim."""
    import copy
    from deeperspeed_b200.runtime.zero.tiling import TiledLinear
    base = torch.nn.Linear(32, 40)
    tiled = TiledLinear(32, 40, bias=True, init_linear=copy.deepcopy(base), out_splits=splits, in_splits=splits)
    for row in range(splits):
        for col in range(splits):
            tile = tiled.linears[row][col]
            assert isinstance(tile, torch.nn.Linear)
            assert torch.equal(tile.weight.data, base.weight.data[_tile_slices(tiled, row, col)])
            if col == splits - 1:             # the bias of an output slice lives in its last tile only
                assert torch.equal(tile.bias.data, base.bias.data[_tile_slices(tiled, row, col)[0]])
            else:
                assert tile.bias is None
    for bad in (0, 33):
        with pytest.raises(RuntimeError):
            TiledLinear(32, 32, out_splits=bad, in_splits=bad)


class _LinearReturningBias(torch.nn.Linear):
    def forward(self, x):
        return super().forward(x), self.bias


@pytest.mark.parametrize("bias", [False, True])
@pytest.mark.parametrize("splits", [1, 2, 3])
@pytest.mark.parametrize("in_f, out_f", [(32, 32), (23, 29), (29, 23)])
@pytest.mark.parametrize("return_bias", [False, True])
def test_tiled_linear_forward_backward_grid(bias, splits, in_f, out_f, return_bias):
    """reference test_tiled_forward / test_tiled_backward / test_tiled_returnbias_backward: outputs and every tile's
    gradients equal the matching block of the full layer's, for uneven feature counts too."""
    import copy
    from deeperspeed_b200.runtime.zero.tiling import TiledLinear, TiledLinearReturnBias
    torch.manual_seed(in_f * out_f + splits)
    if return_bias:
        base = _LinearReturningBias(in_f, out_f, bias=bias)
        tiled = TiledLinearReturnBias(in_f, out_f, bias=bias, linear_cls=_LinearReturningBias,
                                      init_linear=copy.deepcopy(base), out_splits=splits, in_splits=splits)
    else:
        base = torch.nn.Linear(in_f, out_f, bias=bias)
        tiled = TiledLinear(in_f, out_f, bias=bias, init_linear=copy.deepcopy(base), out_splits=splits,
                            in_splits=splits)
    x = torch.rand(3, in_f)
    want, got = base(x.clone()), tiled(x.clone())
    if return_bias:
        (want, want_b), (got, got_b) = want, got
        assert (want_b is None) == (got_b is None) == (not bias)
        if bias:
            assert torch.allclose(want_b, got_b, atol=1e-6)
            want, got = want + want_b, got + got_b
    assert torch.allclose(want, got, atol=1e-5)
    want.sum().backward()
    got.sum().backward()
    for row in range(splits):
        for col in range(splits):
            tile = tiled.linears[row][col]
            rows, cols = _tile_slices(tiled, row, col)
            assert torch.allclose(base.weight.grad[rows, cols], tile.weight.grad, atol=1e-5)
            if tile.bias is not None:
                assert torch.allclose(base.bias.grad[rows], tile.bias.grad, atol=1e-5)


class _InPlaceNet(torch.nn.Module):
    """In-place updates of sub-module outputs, an in-place activation module, a block that is recomputed in backward
    and a module that is called twice."""

    def __init__(self, dim=HID, recompute=True):
        super().__init__()
        self.a, self.b, self.c = (torch.nn.Linear(dim, dim) for _ in range(3))
        self.act = torch.nn.ReLU(inplace=True)
        self.recompute = recompute

    def _block(self, h):
        return self.c(torch.tanh(self.b(h)))

    def forward(self, x):
        h = self.a(x)
        h += x                                  # in place on a sub-module's output
        h = self.act(h)                         # in-place module
        h = self.b(h)
        h.mul_(0.5)
        if self.recompute:
            h = torch.utils.checkpoint.checkpoint(self._block, h, use_reentrant=False)
        else:
            h = self._block(h)
        return h.pow(2).mean()


@pytest.mark.parametrize("recompute", [False, True])
def test_stage3_tolerates_in_place_updates_of_submodule_outputs(recompute):
    """The backward side of stage 3 hangs on gradient hooks, not on wrapped outputs: `h += x`, ReLU(inplace=True), a
    module used twice and a recomputed block all train exactly like the unpartitioned model (torch's module backward
    hooks would raise 'is a view and is being modified inplace'; the reference's wrappers tolerate it too)."""
    import copy

    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        net = _InPlaceNet(recompute=recompute)
        plain = copy.deepcopy(net)
        plain_opt = torch.optim.Adam(plain.parameters(), lr=1e-2)
        cfg = {"train_batch_size": 2, "train_micro_batch_size_per_gpu": 1, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "adam_w_mode": False}},
               "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
        engine, _, _, _ = deepspeed.initialize(model=net, model_parameters=net.parameters(), config_params=cfg)
        g = torch.Generator().manual_seed(2)
        for _ in range(4):
            x = torch.randn(2, 3, HID, generator=g)
            loss = engine(x[dist.get_rank()])
            engine.backward(loss)
            engine.step()
            plain_opt.zero_grad()
            (0.5 * (plain(x[0]) + plain(x[1]))).backward()
            plain_opt.step()
        for (name, p), q in zip(net.named_parameters(), plain.parameters()):
            assert p.ds_status == ZeroParamStatus.NOT_AVAILABLE, name
            with deepspeed.zero.GatheredParameters(p):
                assert torch.allclose(p.detach(), q.detach(), atol=1e-5), (name, (p - q).abs().max())

    run()


def test_stage3_returned_parameter_is_still_the_parameter():
    """What the caller gets back is the ds parameter itself (status AVAILABLE while it is in use), as the reference's
    test_ext_param_return asserts inside its forward."""
    seen = []

    class Outer(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.inner = _ReturnsBias(HID, HID)

        def forward(self, x):
            out, bias = self.inner(x)
            seen.append((bias is self.inner.bias, getattr(bias, "ds_status", None), bias.numel()))
            return (out + bias).sum()

    @distributed_test(world_size=1)
    def run():
        net = torch.nn.Sequential(Outer())
        cfg = {"train_batch_size": 1, "steps_per_print": 100, "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
               "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
        engine, _, _, _ = deepspeed.initialize(model=net, model_parameters=net.parameters(), config_params=cfg)
        for _ in range(2):
            loss = engine(torch.rand(HID))
            engine.backward(loss)
            engine.step()
        assert seen == [(True, ZeroParamStatus.AVAILABLE, HID)] * 2, seen

    run()


@pytest.mark.parametrize("family, stage", [("gpt2", 3), ("bert", 3), ("gpt2", 2), ("gpt2", 1)])
def test_stage3_trains_a_transformers_model_like_plain_adam(family, stage):
    """A model this package did not write: Hugging Face GPT-2 (tied input / output embeddings, Conv1D projections,
    ModelOutput results) and BERT (tuple-returning blocks, LayerNorm everywhere) built from a config, wrapped by a
    stage-3 engine on two ranks, follow plain Adam on the global batch step for step."""
    transformers = pytest.importorskip("transformers")
    import copy

    def build():
        torch.manual_seed(0)
        if family == "gpt2":
            cfg = transformers.GPT2Config(n_layer=2, n_head=2, n_embd=32, vocab_size=128, n_positions=32,
                                          resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0)
            return transformers.GPT2LMHeadModel(cfg)
        cfg = transformers.BertConfig(num_hidden_layers=2, num_attention_heads=2, hidden_size=32, intermediate_size=64,
                                      vocab_size=128, max_position_embeddings=32, hidden_dropout_prob=0.0,
                                      attention_probs_dropout_prob=0.0)
        return transformers.BertForMaskedLM(cfg)

    class LossOnly(torch.nn.Module):
        def __init__(self, inner):
            super().__init__()
            self.inner = inner

        def forward(self, ids):
            return self.inner(input_ids=ids, labels=ids).loss

    @distributed_test(world_size=2)
    def run():
        model = build()
        plain = copy.deepcopy(model)
        plain_opt = torch.optim.Adam(plain.parameters(), lr=1e-3)
        conf = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
                "optimizer": {"type": "Adam", "params": {"lr": 1e-3, "adam_w_mode": False}},
                "fp16": {"enabled": True, "type": "fp32"},
                "zero_optimization": {"stage": stage}}
        if stage == 3:
            conf["zero_optimization"]["stage3_param_persistence_threshold"] = 0
        engine, _, _, _ = deepspeed.initialize(model=LossOnly(model), model_parameters=model.parameters(),
                                               config_params=conf)
        g = torch.Generator().manual_seed(1)
        for _ in range(3):
            ids = torch.randint(0, 128, (4, 16), generator=g)
            mine = ids[2 * dist.get_rank():2 * dist.get_rank() + 2]
            loss = engine(mine)
            engine.backward(loss)
            engine.step()
            plain_opt.zero_grad()
            want = 0.5 * (plain(input_ids=ids[:2], labels=ids[:2]).loss + plain(input_ids=ids[2:], labels=ids[2:]).loss)
            want.backward()
            plain_opt.step()
            mean = loss.detach().clone()
            dist.all_reduce(mean)
            assert float(mean) / 2 == pytest.approx(float(want), rel=1e-5)
        for (name, p), q in zip(model.named_parameters(), plain.parameters()):
            with deepspeed.zero.GatheredParameters(p, enabled=stage == 3):
                assert torch.allclose(p.detach(), q.detach(), atol=2e-5), (name, (p - q).abs().max())

    run()


def test_stage3_parameter_shards_live_in_files_and_rotate_through_the_buffer_pool(tmp_path):
    """offload_param nvme: every sub-group's low-precision shard arena is a file under
    <nvme>/zero_stage_3/*params/rank<R>/, at most `buffer_count` of them are in host memory at once (the others have
    empty `ds_tensor`s), arenas are read back for gathers and rewritten by the optimizer step, a modification made
    under GatheredParameters reaches the file, a checkpoint round trip restores them - and training equals AdamW."""
    import copy
    from simple_model import LinearStack

    @distributed_test(world_size=2)
    def run():
        nvme = os.path.join(str(tmp_path), "nvme")
        os.makedirs(nvme, exist_ok=True)
        dim, layers = 24, 5
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 0.01, "weight_decay": 0.01}},
               "fp16": {"enabled": True, "type": "fp32"},
               "aio": {"block_size": 65536, "queue_depth": 4, "thread_count": 1},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0, "sub_group_size": 300,
                                     "offload_optimizer": {"device": "cpu"},
                                     "offload_param": {"device": "nvme", "nvme_path": nvme, "buffer_count": 2,
                                                       "buffer_size": 1024}}}
        torch.manual_seed(7)
        model = LinearStack(dim, dim, dim, num_layers=layers)
        ref = LinearStack(dim, dim, dim, num_layers=layers)
        ref.load_state_dict(model.state_dict())
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        swapper = opt.param_swapper
        arenas = len(opt.sub_groups)
        assert swapper is not None and arenas >= 5 and all(a is None for a in opt.fp16_partitioned_groups_flat)
        folder = swapper.folder
        assert sorted(os.listdir(folder)) == sorted(f"arena_{i}.swp" for i in range(arenas))
        assert f"rank{dist.get_rank()}" in folder and "zero_stage_3" in folder

        def resident_params():
            return sum(p.ds_tensor.numel() > 0 for p in model.parameters())

        ropt = torch.optim.AdamW(ref.parameters(), lr=0.01, weight_decay=0.01)
        g = torch.Generator().manual_seed(3)
        for _ in range(3):
            x = torch.randn(4, dim, generator=g)
            y = torch.randint(0, dim, (4,), generator=g)
            mine = slice(2 * dist.get_rank(), 2 * dist.get_rank() + 2)
            loss = engine(x[mine], y[mine]); engine.backward(loss); engine.step()
            assert len(swapper.resident_arenas()) <= 2 and all(h == 0 for h in swapper.holders.values())
            assert resident_params() < len(list(model.parameters()))
            ropt.zero_grad()
            (0.5 * (ref(x[:2], y[:2]) + ref(x[2:], y[2:]))).backward()
            ropt.step()
        assert swapper.stats["evictions"] > 0 and swapper.stats["reads"] > 0 and swapper.stats["writes"] >= 4 * arenas
        assert swapper.stats["prefetch_hits"] > 0          # reads started ahead along the recorded module trace

        def check_against(reference):
            for (name, p), r in zip(model.named_parameters(), reference.parameters()):
                with deepspeed.zero.GatheredParameters(p):
                    assert torch.allclose(p.detach(), r.detach(), atol=3e-5), (name, (p - r).abs().max())

        check_against(ref)
        # checkpoint round trip into a fresh engine that swaps to its own folder
        engine.save_checkpoint(os.path.join(str(tmp_path), "ckpt"), tag="t")
        dist.barrier()
        cfg2 = copy.deepcopy(cfg)
        cfg2["zero_optimization"]["offload_param"]["nvme_path"] = os.path.join(str(tmp_path), "nvme2")
        torch.manual_seed(99)
        model2 = LinearStack(dim, dim, dim, num_layers=layers)
        engine2, opt2, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg2), model=model2,
                                                   model_parameters=model2.parameters())
        path, _ = engine2.load_checkpoint(os.path.join(str(tmp_path), "ckpt"), tag="t")
        assert path is not None
        for p2, r in zip(model2.parameters(), ref.parameters()):
            with deepspeed.zero.GatheredParameters(p2):
                assert torch.allclose(p2.detach(), r.detach(), atol=3e-5)
        assert opt2.param_swapper.folder != swapper.folder and opt2.param_swapper.stats["writes"] >= 2 * arenas
        # a modification through GatheredParameters lands in the file (the arena is evicted and read again below)
        first = next(model.parameters())
        with deepspeed.zero.GatheredParameters(first, modifier_rank=0):
            if dist.get_rank() == 0:
                with torch.no_grad():
                    first.fill_(0.125)
        for p in list(model.parameters())[1:]:
            with deepspeed.zero.GatheredParameters(p):
                pass
        with deepspeed.zero.GatheredParameters(first):
            assert torch.all(first == 0.125)

    run()


@pytest.mark.parametrize("offload", [None, "cpu", "nvme"])
def test_stage3_weight_edit_after_initialize_survives_the_next_step(offload, tmp_path):
    """Weights written through GatheredParameters AFTER the engine exists (loading pretrained weights, re-initialising
    a head) are taken over by the fp32 master - also when it lives in host memory or in a swap file -, so the next
    optimizer step starts from them instead of undoing the edit."""
    @distributed_test(world_size=2)
    def run():
        zero = {"stage": 3, "stage3_param_persistence_threshold": 0, "sub_group_size": 200}
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-3, "adam_w_mode": False}},
               "fp16": {"enabled": True, "type": "fp32"}, "zero_optimization": zero}
        if offload == "cpu":
            zero["offload_optimizer"] = {"device": "cpu"}
        elif offload == "nvme":
            cfg["aio"] = {"block_size": 65536, "queue_depth": 4, "thread_count": 1}
            zero["offload_optimizer"] = {"device": "nvme", "nvme_path": str(tmp_path), "buffer_count": 4}
        torch.manual_seed(0)
        model = SimpleModel(HID)
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        torch.manual_seed(5)
        new_weight = torch.randn(HID, HID)              # same on both ranks
        with deepspeed.zero.GatheredParameters(model.linear.weight, modifier_rank=0):
            if dist.get_rank() == 0:
                with torch.no_grad():
                    model.linear.weight.copy_(new_weight)
        x, y = next(iter(torch.utils.data.DataLoader(random_dataset(8, HID, seed=dist.get_rank()), batch_size=2)))
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
        with deepspeed.zero.GatheredParameters(model.linear.weight):
            drift = (model.linear.weight.detach() - new_weight).abs().max()
        assert 0 < drift <= 1.01e-3, drift           # one Adam step of size lr away from the EDITED weights
        # the explicit refresh (reference API) is a no-op now: masters already equal the shards
        before = [m.data.clone() if m.data.numel() else None for m in engine.optimizer.fp32_partitioned_groups_flat]
        engine.optimizer.refresh_fp32_params()
        for m, b in zip(engine.optimizer.fp32_partitioned_groups_flat, before):
            assert b is None or torch.equal(m.data, b)

    run()


def test_execution_trace_replay_reuse_distance_and_rerecording():
    """ExecutionTrace alone: records one iteration, answers look-ahead and reuse-distance queries on replay, and
    re-records from the point where a later iteration leaves the recorded order."""
    from deeperspeed_b200.runtime.zero.stage3 import ExecutionTrace
    sizes = {0: 0, 1: 10, 2: 20, 3: 30}
    trace = ExecutionTrace()
    first = [0, 1, 2, 3, 3, 2, 1, 0]                   # forward 0..3, backward 3..0
    trace.begin_pass()
    for mid in first:
        assert trace.upcoming(4) == [] and trace.reuse_distance(mid) == float("inf")      # nothing known yet
        trace.advance(mid)
    trace.end_iteration(sizes.__getitem__)
    assert trace.ready and trace.order == first
    trace.begin_pass()
    trace.advance(0)
    assert trace.upcoming(3) == [1, 2, 3]
    trace.advance(1); trace.advance(2); trace.advance(3)
    assert trace.upcoming(8) == [3, 2, 1, 0]           # the look-ahead runs into the backward pass
    assert trace.reuse_distance(3) == 0                # runs again immediately
    assert trace.reuse_distance(2) == 30 + 30          # both fetches of module 3 lie in between
    assert trace.reuse_distance(1) == 20 + 30 + 30 + 20
    for mid in (3, 2, 1, 0):
        trace.advance(mid)
    assert trace.reuse_distance(0) == float("inf")     # last use of the iteration
    trace.end_iteration(sizes.__getitem__)
    # an iteration that skips module 2 on the way back: truncated at the deviation and recorded again
    trace.begin_pass()
    for mid in (0, 1, 2, 3, 3):
        trace.advance(mid)
    assert trace.ready
    trace.advance(1)
    assert not trace.ready and trace.order == [0, 1, 2, 3, 3, 1]
    trace.advance(0)
    trace.end_iteration(sizes.__getitem__)
    assert trace.ready and trace.order == [0, 1, 2, 3, 3, 1, 0]


def test_stage3_keeps_the_forward_tail_for_backward_and_prefetches_in_backward(tmp_path):
    """From the second iteration on (the first records the trace): the last layer is still gathered when the forward
    pass ends, because its backward comes first (`stage3_max_reuse_distance`); with a reuse distance of 0 it is
    released like every other layer. Backward issues asynchronous gathers ahead of the module it is in. Either way
    the result equals plain Adam."""
    import copy
    from simple_model import LinearStack

    @distributed_test(world_size=2)
    def run():
        dim = 16
        torch.manual_seed(0)
        base = LinearStack(dim, dim, dim, num_layers=3)

        def train(reuse):
            model = copy.deepcopy(base)
            cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
                   "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "adam_w_mode": False}},
                   "fp16": {"enabled": True, "type": "fp32"},
                   "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0,
                                         "stage3_max_reuse_distance": reuse, "stage3_prefetch_bucket_size": 10 ** 6}}
            engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                     model_parameters=model.parameters())
            pc = opt.param_coordinator
            from deeperspeed_b200.runtime.zero.partition_parameters import Init
            gathers = []                                  # (iteration, "forward" | "backward", asynchronous?)
            where = {"iteration": 0, "phase": "forward"}
            plain_gather = Init._all_gather

            def counting(init, param_list, async_op=False, hierarchy=None):
                needed = [p for p in param_list if p.ds_status == ZeroParamStatus.NOT_AVAILABLE]
                if needed:
                    gathers.append((where["iteration"], where["phase"], bool(async_op)))
                return plain_gather(init, param_list, async_op=async_op, hierarchy=hierarchy)

            Init._all_gather = counting
            g = torch.Generator().manual_seed(4)
            gathered_after_forward = []
            for _ in range(3):
                x = torch.randn(4, dim, generator=g)
                y = torch.randint(0, dim, (4,), generator=g)
                mine = slice(2 * dist.get_rank(), 2 * dist.get_rank() + 2)
                where["phase"] = "forward"
                loss = engine(x[mine], y[mine])
                gathered_after_forward.append(sorted(n for n, p in model.named_parameters()
                                                     if p.ds_status == ZeroParamStatus.AVAILABLE))
                where["phase"] = "backward"
                engine.backward(loss)
                where["phase"] = "step"
                engine.step()
                where["iteration"] += 1
            Init._all_gather = plain_gather
            return model, gathered_after_forward, gathers, pc.trace

        kept, after_fwd, _, trace = train(reuse=10 ** 9)
        half = len(trace.order) // 2                                    # every module that ran: forward, then backward
        assert trace.ready and half >= 6 and sorted(trace.order[:half]) == sorted(trace.order[half:])
        assert after_fwd[0] == []                                       # first iteration: nothing known, all released
        assert "output_layer.weight" in after_fwd[1] and "output_layer.weight" in after_fwd[2]
        strict, after_fwd0, gathers, _ = train(reuse=0)
        assert after_fwd0 == [[], [], []]
        # first iteration: every gather is synchronous, in forward and again in backward; once the trace is known the
        # look-ahead (which runs across the turn into backward) leaves only the very first gather synchronous
        first = [g for g in gathers if g[0] == 0]
        assert {g[1] for g in first} == {"forward", "backward"} and not any(g[2] for g in first)
        later = [g for g in gathers if g[0] == 2]
        assert sum(not g[2] for g in later) <= 1 and sum(g[2] for g in later) >= 2
        assert not [g for g in later if g[1] == "backward" and not g[2]]
        plain = copy.deepcopy(base)
        popt = torch.optim.Adam(plain.parameters(), lr=1e-2)
        g = torch.Generator().manual_seed(4)
        for _ in range(3):
            x = torch.randn(4, dim, generator=g)
            y = torch.randint(0, dim, (4,), generator=g)
            popt.zero_grad()
            (0.5 * (plain(x[:2], y[:2]) + plain(x[2:], y[2:]))).backward()
            popt.step()
        for model in (kept, strict):
            for (name, p), q in zip(model.named_parameters(), plain.parameters()):
                with deepspeed.zero.GatheredParameters(p):
                    assert torch.allclose(p.detach(), q.detach(), atol=1e-5), name

    run()


def test_stage3_gradient_accumulation_matches_adam_on_the_global_batch(tmp_path):
    """Two micro-batches per step on each of two ranks (every micro-batch is one pass of the execution trace, the
    gradient shards accumulate across them) equal one Adam step on the four micro-batches together."""
    import copy
    from simple_model import LinearStack

    @distributed_test(world_size=2)
    def run():
        dim = 16
        torch.manual_seed(0)
        model = LinearStack(dim, dim, dim, num_layers=2)
        plain = copy.deepcopy(model)
        popt = torch.optim.Adam(plain.parameters(), lr=1e-2)
        cfg = {"train_batch_size": 8, "train_micro_batch_size_per_gpu": 2, "gradient_accumulation_steps": 2,
               "steps_per_print": 100, "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "adam_w_mode": False}},
               "fp16": {"enabled": True, "type": "fp32"}, "gradient_clipping": 1.0,
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
        engine, _, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                               model_parameters=model.parameters())
        g = torch.Generator().manual_seed(9)
        for _ in range(3):
            x = torch.randn(8, dim, generator=g)
            y = torch.randint(0, dim, (8,), generator=g)
            for micro in range(2):                       # rank r takes micro-batches 2r and 2r + 1
                lo = 4 * dist.get_rank() + 2 * micro
                loss = engine(x[lo:lo + 2], y[lo:lo + 2])
                engine.backward(loss)
                engine.step()
            popt.zero_grad()
            sum(plain(x[i:i + 2], y[i:i + 2]) for i in range(0, 8, 2)).div(4).backward()
            torch.nn.utils.clip_grad_norm_(plain.parameters(), 1.0)
            popt.step()
        assert engine.global_steps == 3
        for (name, p), q in zip(model.named_parameters(), plain.parameters()):
            with deepspeed.zero.GatheredParameters(p):
                assert torch.allclose(p.detach(), q.detach(), atol=2e-5), (name, (p - q).abs().max())

    run()


def test_stage3_evaluation_pass_between_training_steps_releases_everything(tmp_path):
    """A no-grad forward pass in the middle of training follows the forward part of the recorded trace, leaves no
    parameter gathered (the tail kept for a backward that is not coming is released) and does not disturb training."""
    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        model = SimpleModel(HID)
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-2}}, "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
        engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                 model_parameters=model.parameters())
        data = list(torch.utils.data.DataLoader(random_dataset(16, HID, seed=dist.get_rank()), batch_size=2))
        losses = []
        for i, (x, y) in enumerate(data[:4]):
            loss = engine(x, y); engine.backward(loss); engine.step()
            losses.append(float(loss))
            engine.eval()
            with torch.no_grad():
                first = float(engine(*data[5]))
                assert all(p.ds_status == ZeroParamStatus.NOT_AVAILABLE for p in model.parameters())
                assert float(engine(*data[5])) == first            # twice in a row, same answer
            engine.train()
            assert opt.param_coordinator.trace.ready and opt.param_coordinator.trace.position == -1
        assert all(l == l for l in losses)

    run()


def test_arena_swapper_alone(tmp_path):
    """ArenaSwapper without an engine: contents survive eviction, the least recently used free arena goes first, held
    arenas are never evicted (a clear error when every buffer is held), oversized arenas are refused, a prefetch is
    picked up by the next acquire, and an asynchronous write is finished before its file is read back."""
    from deeperspeed_b200.runtime.swap_tensor.arena_swapper import ArenaSwapper
    sw = ArenaSwapper(str(tmp_path), torch.float32, buffer_count=2, buffer_size=1000,
                      aio_config={"block_size": 65536, "queue_depth": 4, "thread_count": 1})
    seen = {}
    for idx in range(3):
        sw.register(idx, 700 + idx, lambda view, idx=idx: seen.__setitem__(idx, view))
        buf = sw.acquire(idx, load=False)
        buf.fill_(float(idx + 1))
        sw.write_back(idx, async_op=(idx == 2))          # the last one is still on its way out
        sw.release(idx)
    assert sw.resident_arenas() == [1, 2] and seen[0] is None and seen[2].numel() == 702
    with pytest.raises(ValueError):
        sw.register(9, 5000, lambda view: None)
    back = sw.acquire(0)                                 # evicts 1 (least recently used), reads the file
    assert torch.all(back == 1.0) and back.numel() == 700 and sw.resident_arenas() == [2, 0]
    held_too = sw.acquire(2)
    assert torch.all(held_too == 3.0)
    with pytest.raises(RuntimeError):
        sw.acquire(1)                                    # both buffers are held
    sw.release(2)
    assert sw.prefetch(1) and not sw.prefetch(1)         # takes over arena 2's buffer; a second call has nothing to do
    one = sw.acquire(1)
    assert torch.all(one == 2.0) and sw.stats["prefetch_hits"] == 1
    sw.release(1)
    sw.release(0)
    again = sw.acquire(2)                                # written asynchronously above, evicted since
    assert torch.all(again == 3.0)
    sw.release(2)
    sw.flush()
    assert all(h == 0 for h in sw.holders.values())
    assert sorted(os.listdir(sw.folder)) == ["arena_0.swp", "arena_1.swp", "arena_2.swp"]


class _Branchy(torch.nn.Module):
    """Takes a different path through its layers from one step to the next."""

    def __init__(self, dim=HID):
        super().__init__()
        self.first, self.detour, self.last = (torch.nn.Linear(dim, dim) for _ in range(3))
        self.calls = 0

    def forward(self, x):
        self.calls += 1
        h = torch.tanh(self.first(x))
        if self.calls % 2 == 0:
            h = torch.tanh(self.detour(h)) + h
        return self.last(h).pow(2).mean()


def test_stage3_follows_control_flow_that_changes_between_steps():
    """The recorded execution trace is a hint, not a contract: a model whose path changes from step to step (one layer
    is skipped every other step, so it has no gradient then) trains exactly like plain Adam; the trace is cut at the
    first deviation and recorded again."""
    import copy

    @distributed_test(world_size=2)
    def run():
        torch.manual_seed(0)
        net = _Branchy()
        plain = copy.deepcopy(net)
        popt = torch.optim.Adam(plain.parameters(), lr=1e-2)
        cfg = {"train_batch_size": 4, "train_micro_batch_size_per_gpu": 2, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-2, "adam_w_mode": False}},
               "fp16": {"enabled": True, "type": "fp32"},
               "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 0}}
        engine, opt, _, _ = deepspeed.initialize(model=net, model_parameters=net.parameters(), config_params=cfg)
        g = torch.Generator().manual_seed(6)
        lengths = []
        for _ in range(5):
            x = torch.randn(4, HID, generator=g)
            loss = engine(x[2 * dist.get_rank():2 * dist.get_rank() + 2])
            engine.backward(loss)
            engine.step()
            lengths.append(len(opt.param_coordinator.trace.order))
            popt.zero_grad()
            plain.calls += 1
            took_detour = plain.calls % 2 == 0
            plain.calls -= 1                                  # both halves of the global batch take the same path
            a = plain(x[:2])
            plain.calls -= 1
            b = plain(x[2:])
            (0.5 * (a + b)).backward()
            if not took_detour:                               # the engine's Adam sees a zero gradient there, not None
                for p in plain.detour.parameters():
                    p.grad = torch.zeros_like(p)
            popt.step()
        assert len(set(lengths)) == 2, lengths                # the trace was re-recorded as the path changed
        for (name, p), q in zip(net.named_parameters(), plain.parameters()):
            assert p.ds_status == ZeroParamStatus.NOT_AVAILABLE
            with deepspeed.zero.GatheredParameters(p):
                assert torch.allclose(p.detach(), q.detach(), atol=1e-5), (name, (p - q).abs().max())

    run()
