"""Remaining scenarios of the reference's unit suite, restated against this package: launcher resource filters
(tests/unit/test_run.py), elasticity error paths incl. the engine-level ones (test_elastic.py), LR scheduler wiring
through the engine (test_lr_schedulers.py) and the pipeline schedules' stage queries and per-stage instruction streams
(test_pipe_schedule.py). CPU only."""
import copy
import json
import os

import pytest
import torch

import deeperspeed_b200 as deepspeed
from common import distributed_test
from simple_model import SimpleModel, random_dataloader, args_from_dict

# ------------------------------------------------------------------------------------------------- launcher filters
ONE = {"worker-0": [0, 1, 2, 3]}
TWO = {"worker-0": [0, 1, 2, 3], "worker-1": [0, 1, 2, 3]}


@pytest.mark.parametrize("hosts, kwargs, want", [
    (ONE, {}, ONE),
    (ONE, {"exclude_str": "worker-0:1"}, {"worker-0": [0, 2, 3]}),
    (ONE, {"exclude_str": "worker-0:1,2"}, {"worker-0": [0, 3]}),
    (ONE, {"include_str": "worker-0:1"}, {"worker-0": [1]}),
    (ONE, {"include_str": "worker-0:1,1"}, {"worker-0": [1]}),                       # repeated slot
    (ONE, {"include_str": "worker-0:1@worker-0:0,1"}, {"worker-0": [0, 1]}),         # repeated host
    (ONE, {"include_str": "worker-0"}, ONE),                                         # bare host = all its slots
    (ONE, {"exclude_str": "worker-0"}, {}),
    (ONE, {"exclude_str": "worker-0:0,1,2,3"}, {}),
    (TWO, {}, TWO),
    (TWO, {"include_str": "worker-1:0,3"}, {"worker-1": [0, 3]}),
    (TWO, {"exclude_str": "worker-1"}, {"worker-0": [0, 1, 2, 3]}),
    (TWO, {"exclude_str": "worker-0:0,1@worker-1:3"}, {"worker-0": [2, 3], "worker-1": [0, 1, 2]}),
])
def test_resource_filter_results(hosts, kwargs, want):
    from deeperspeed_b200.launcher import runner
    assert dict(runner.parse_resource_filter(copy.deepcopy(hosts), **kwargs)) == want


@pytest.mark.parametrize("kwargs", [
    {"include_str": "A", "exclude_str": "B"},              # mutually exclusive
    {"include_str": "jeff"}, {"exclude_str": "jeff"},      # unknown host
    {"include_str": "worker-1:4"}, {"exclude_str": "worker-1:4"},      # unknown slot
    {"exclude_str": "worker-1@worker-0:1@5"},              # malformed
])
def test_resource_filter_errors(kwargs):
    from deeperspeed_b200.launcher import runner
    with pytest.raises(ValueError):
        runner.parse_resource_filter(copy.deepcopy(TWO), **kwargs)


@pytest.mark.parametrize("argv", ["--num_nodes 1 -i localhost foo.py", "--num_nodes 1 --num_gpus 1 -i localhost foo.py",
                                  "--num_gpus 1 -i localhost foo.py", "--num_nodes 1 -e localhost foo.py",
                                  "--num_nodes 1 --num_gpus 1 -e localhost foo.py", "--num_gpus 1 -e localhost foo.py"])
def test_runner_rejects_counts_together_with_filters(argv):
    from deeperspeed_b200.launcher import runner
    with pytest.raises(ValueError):
        runner.main(args=argv.split())


# ------------------------------------------------------------------------------------------------- elasticity
ELASTIC = {"enabled": True, "max_train_batch_size": 10000, "micro_batch_sizes": [8, 12, 16, 17], "min_gpus": 32,
           "max_gpus": 1500, "min_time": 20, "version": 0.1}


def _elastic(drop=(), **changes):
    block = {k: v for k, v in ELASTIC.items() if k not in drop}
    block.update(changes)
    return {"elasticity": block}


@pytest.mark.parametrize("config", [
    _elastic(enabled=False), _elastic(version="0.2"), _elastic(drop=("max_train_batch_size",)),
    _elastic(drop=("micro_batch_sizes",)), {"elasticity": {"enabled": True}},
    _elastic(micro_batch_sizes=[1, 4, -1, 2, -10]), _elastic(min_gpus=-1), _elastic(max_gpus=-1),
    _elastic(micro_batch_sizes=5), _elastic(micro_batch_sizes=["a", None, 0.5]), _elastic(micro_batch_sizes=[2, 0.5, 4]),
])
def test_elastic_config_rejections(config):
    from deeperspeed_b200.elasticity import compute_elastic_config
    from deeperspeed_b200.elasticity.config import ElasticityError
    with pytest.raises(ElasticityError):
        compute_elastic_config(ds_config=config, target_deepspeed_version=deepspeed.__version__)


def test_elastic_micro_batch_choice_for_a_world_size():
    from deeperspeed_b200.elasticity import compute_elastic_config
    from deeperspeed_b200.elasticity.config import ElasticityError, ElasticityIncompatibleWorldSize
    version = deepspeed.__version__
    assert compute_elastic_config(_elastic(), version, world_size=64)[2] == 17
    with pytest.raises(ElasticityIncompatibleWorldSize):
        compute_elastic_config(_elastic(), version, world_size=128)
    small = _elastic(max_train_batch_size=32, micro_batch_sizes=[1, 2, 3, 7], min_gpus=1)
    assert compute_elastic_config(small, version, world_size=7)[2] == 3       # largest micro batch that still fits
    with pytest.raises(ElasticityError):
        compute_elastic_config(_elastic(), "0.2")                             # package older than elasticity itself


def _engine_elastic_conf(**more):
    block = {"enabled": True, "max_train_batch_size": 4, "micro_batch_sizes": [1, 2, 3, 4], "min_gpus": 1,
             "max_gpus": 4, "min_time": 20, "version": 0.1}
    block.update(more)
    return {"train_batch_size": 2, "steps_per_print": 100, "gradient_clipping": 1.0,
            "optimizer": {"type": "Lamb", "params": {"lr": 1.5e-4}}, "elasticity": block}


@pytest.mark.parametrize("world", [1, 2])
def test_engine_with_elasticity_and_explicit_batch_sizes(world, tmp_path, monkeypatch):
    """Batch parameters next to an enabled elasticity block are an error unless `ignore_non_elastic_batch_info`; then
    elasticity decides the batch sizes. A scheduler-side elastic config (environment) that differs from the runtime
    one is an error too (reference test_non_elastic_batch_params*, test_elastic_config_changed)."""
    from deeperspeed_b200.elasticity.config import ElasticityError
    monkeypatch.delenv("DEEPSPEED_ELASTICITY_CONFIG", raising=False)

    @distributed_test(world_size=world)
    def run():
        def build(conf):
            model = SimpleModel(10)
            return deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model,
                                        model_parameters=model.parameters())[0]

        with pytest.raises(ElasticityError):
            build(_engine_elastic_conf())
        engine = build(_engine_elastic_conf(ignore_non_elastic_batch_info=True))
        # elasticity's choice, not the explicit 2: the least common multiple of the micro batch sizes is a candidate
        # even above max_train_batch_size, and it is the one every GPU count 1..4 divides (same as the reference)
        assert engine.train_batch_size() == 12
        assert engine.train_batch_size() == (engine.train_micro_batch_size_per_gpu() * world
                                             * engine.gradient_accumulation_steps())
        changed = _engine_elastic_conf(ignore_non_elastic_batch_info=True, max_train_batch_size=27)
        os.environ["DEEPSPEED_ELASTICITY_CONFIG"] = json.dumps(changed)
        try:
            with pytest.raises(ElasticityError):
                build(_engine_elastic_conf(ignore_non_elastic_batch_info=True))
        finally:
            del os.environ["DEEPSPEED_ELASTICITY_CONFIG"]

    run()


# ------------------------------------------------------------------------------------------------- LR schedulers
SCHEDULERS = {
    "LRRangeTest": {"lr_range_test_min_lr": 1e-3, "lr_range_test_step_size": 5, "lr_range_test_step_rate": 2.0},
    "OneCycle": {"cycle_min_lr": 1e-3, "cycle_max_lr": 1e-2, "cycle_first_step_size": 5},
    "WarmupLR": {"warmup_min_lr": 1e-4, "warmup_max_lr": 1e-2, "warmup_num_steps": 5},
    "WarmupDecayLR": {"warmup_min_lr": 1e-4, "warmup_max_lr": 1e-2, "warmup_num_steps": 5, "total_num_steps": 20},
}


@pytest.mark.parametrize("name", sorted(SCHEDULERS))
def test_engine_scheduler_wiring(name, tmp_path):
    """reference test_get_lr_before_train / test_scheduler_optimizer_parity: the scheduler the engine builds drives the
    engine's own optimizer, answers get_lr() before the first step, and the optimizer's lr follows it step by step."""
    conf = {"train_batch_size": 2, "steps_per_print": 100, "optimizer": {"type": "Adam", "params": {"lr": 1.5e-4}},
            "scheduler": {"type": name, "params": SCHEDULERS[name]}, "gradient_clipping": 1.0}

    @distributed_test(world_size=1)
    def run():
        model = SimpleModel(10)
        engine, optimizer, _, scheduler = deepspeed.initialize(args=args_from_dict(tmp_path, conf), model=model,
                                                               model_parameters=model.parameters())
        assert scheduler is engine.lr_scheduler and type(scheduler).__name__ == name
        assert scheduler.optimizer is optimizer
        before = scheduler.get_lr()
        assert len(before) == len(optimizer.param_groups) and all(lr >= 0 for lr in before)   # warm-up: [0.0] + warning
        seen = []
        for i, (x, y) in enumerate(random_dataloader(engine, 24, 10, engine.device)):
            loss = engine(x, y)
            engine.backward(loss)
            engine.step()
            assert engine.get_lr() == [g["lr"] for g in optimizer.param_groups] == list(scheduler.get_last_lr())
            seen.append(engine.get_lr()[0])
        if name == "LRRangeTest":
            assert seen == sorted(seen) and seen[-1] > seen[0]            # keeps growing
        elif name == "WarmupLR":
            assert seen[:5] == sorted(seen[:5]) and seen[5:] == [pytest.approx(1e-2)] * len(seen[5:])
        elif name == "WarmupDecayLR":
            peak = seen.index(max(seen))
            assert 3 <= peak <= 5 and seen[peak:] == sorted(seen[peak:], reverse=True) and seen[-1] < 0.8 * seen[peak]
        else:
            peak = seen.index(max(seen))
            assert max(seen) == pytest.approx(1e-2, rel=0.05) and 3 <= peak <= 5 and seen[-1] < seen[peak]

    run()


# ------------------------------------------------------------------------------------------------- pipeline schedules
def _names(cmds):
    return [type(c).__name__ for c in cmds]


def test_schedule_stage_and_buffer_queries():
    """reference test_pipe_stagequery + the buffer-count rule of TrainSchedule (schedule.py:243-249)."""
    from deeperspeed_b200.runtime.pipe import schedule
    for stages in (1, 2, 4):
        for stage in range(stages):
            s = schedule.TrainSchedule(micro_batches=8, stages=stages, stage_id=stage)
            assert s.stage == stage and s.num_stages == stages and s.num_micro_batches == 8
            assert s.is_first_stage == (stage == 0) and s.is_last_stage == (stage == stages - 1)
            assert s.prev_stage == stage - 1 and s.next_stage == stage + 1
            assert s._valid_stage(stage) and not s._valid_stage(-1) and not s._valid_stage(stages)
            assert s._valid_micro_batch(0) and s._valid_micro_batch(7)
            assert not s._valid_micro_batch(-1) and not s._valid_micro_batch(8)
            assert s.num_pipe_buffers() == max(2, min(stages - stage + 1, 8))
    assert schedule.InferenceSchedule(micro_batches=4, stages=3, stage_id=1).num_pipe_buffers() == 2
    assert schedule.DataParallelSchedule(micro_batches=4, stages=1, stage_id=0).num_pipe_buffers() == 1


@pytest.mark.parametrize("micro_batches", [1, 3, 8])
@pytest.mark.parametrize("stages", [1, 2, 3, 4])
def test_train_schedule_streams_are_complete_and_consistent(micro_batches, stages):
    """Every stage loads / forwards / backwards each micro batch exactly once in order, never runs a backward before
    that micro batch's forward, sends exactly what its neighbour receives in the same step, reduces and steps once at
    the very end, and never uses a buffer index outside num_pipe_buffers()
    (the properties behind reference test_pipe_schedule_firststage / laststage / singlestage)."""
    from deeperspeed_b200.runtime.pipe import schedule as S
    streams = [list(S.TrainSchedule(micro_batches, stages, stage)) for stage in range(stages)]
    assert len({len(s) for s in streams}) == 1                                  # lock step
    assert len(streams[0]) == 2 * (micro_batches + stages - 1)
    for stage, steps in enumerate(streams):
        sched = S.TrainSchedule(micro_batches, stages, stage)
        flat = [c for step in steps for c in step]
        count = lambda kind: sum(isinstance(c, kind) for c in flat)
        assert count(S.ForwardPass) == count(S.BackwardPass) == micro_batches
        assert count(S.LoadMicroBatch) == (micro_batches if stage in (0, stages - 1) else 0)
        assert count(S.SendActivation) == count(S.RecvGrad) == (micro_batches if stage < stages - 1 else 0)
        assert count(S.RecvActivation) == count(S.SendGrad) == (micro_batches if stage > 0 else 0)
        assert _names(steps[-1])[-3:] == ["ReduceTiedGrads", "ReduceGrads", "OptimizerStep"]
        assert count(S.OptimizerStep) == count(S.ReduceGrads) == count(S.ReduceTiedGrads) == 1
        forwards_done = backwards_done = 0
        for c in flat:
            if hasattr(c, "buffer_id"):
                assert 0 <= c.buffer_id < sched.num_pipe_buffers()
            if isinstance(c, S.ForwardPass):
                forwards_done += 1
            if isinstance(c, S.BackwardPass):
                backwards_done += 1
                assert backwards_done <= forwards_done
        # 1F1B: never more forwards in flight than there are stages after (and including) this one
        in_flight = peak = 0
        for c in flat:
            in_flight += isinstance(c, S.ForwardPass) - isinstance(c, S.BackwardPass)
            peak = max(peak, in_flight)
        assert peak <= min(stages - stage, micro_batches)
    for step in range(len(streams[0])):
        for stage in range(stages - 1):
            here, there = streams[stage][step], streams[stage + 1][step]
            assert (sum(isinstance(c, S.SendActivation) for c in here)
                    == sum(isinstance(c, S.RecvActivation) for c in there))
            assert (sum(isinstance(c, S.RecvGrad) for c in here) == sum(isinstance(c, S.SendGrad) for c in there))


@pytest.mark.parametrize("stages, stage", [(1, 0), (3, 0), (3, 1), (3, 2)])
def test_inference_schedule_per_stage_streams(stages, stage):
    """reference test_pipe_inference_schedule_{single,first,mid,last}stage: which instructions a stage issues, how
    many steps, and the two alternating buffers."""
    from deeperspeed_b200.runtime.pipe import schedule as S
    micro_batches = 4
    steps = list(S.InferenceSchedule(micro_batches, stages, stage))
    assert len(steps) == micro_batches + stages - 1
    flat = [c for step in steps for c in step]
    count = lambda kind: sum(isinstance(c, kind) for c in flat)
    assert count(S.ForwardPass) == micro_batches
    assert count(S.LoadMicroBatch) == (micro_batches if stage in (0, stages - 1) else 0)
    assert count(S.RecvActivation) == (micro_batches if stage > 0 else 0)
    assert count(S.SendActivation) == (micro_batches if stage < stages - 1 else 0)
    assert count(S.BackwardPass) == count(S.OptimizerStep) == 0
    assert {c.buffer_id for c in flat if hasattr(c, "buffer_id")} <= {0, 1}
    busy = [i for i, step in enumerate(steps) if any(isinstance(c, S.ForwardPass) for c in step)]
    assert busy == list(range(stage, stage + micro_batches))           # stage s starts s steps late
