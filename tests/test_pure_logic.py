"""Pure-logic goldens: topology, partitioning, schedules, elasticity, LR schedules, loss scaler, config
(oracles harvested from the reference's unit tests)."""
import json
import math

import pytest
import torch

import deeperspeed_b200 as deepspeed
from deeperspeed_b200.runtime.pipe.topology import ProcessTopology, PipeDataParallelTopology, PipeModelDataParallelTopology, _prime_factors
from deeperspeed_b200.runtime.pipe import schedule
from deeperspeed_b200.runtime.utils import partition_uniform, partition_balanced, prefix_sum_inc, call_to_str
from deeperspeed_b200.runtime.fp16.loss_scaler import DynamicLossScaler, LossScaler
from deeperspeed_b200.runtime import lr_schedules
from deeperspeed_b200.runtime.config import DeepSpeedConfig, dict_raise_error_on_duplicate_keys
from deeperspeed_b200.elasticity import compute_elastic_config, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize


def test_topology_rank_math():
    topo = ProcessTopology(axes=["pipe", "data", "model"], dims=[2, 2, 2])
    for p in range(2):
        for d in range(2):
            for m in range(2):
                assert topo.get_rank(pipe=p, data=d, model=m) == 4 * p + 2 * d + m
    assert topo.filter_match(pipe=0, data=1) == [2, 3]
    topo = ProcessTopology(axes=["a", "b", "c"], dims=[2, 2, 2])
    assert topo.get_axis_list("b", 1) == [2, 3, 6, 7]
    assert topo.get_axis_comm_lists("c") == [[0, 1], [2, 3], [4, 5], [6, 7]]
    assert topo.get_axis_comm_lists("jeff") == []


def test_topology_rank_repr_and_grids():
    topo = PipeModelDataParallelTopology(num_pp=2, num_mp=2, num_dp=2)
    assert topo.get_rank_repr(rank=0) == "model_00"
    assert topo.get_rank_repr(rank=1) == "model_01"
    assert topo.get_rank_repr(rank=0, omit_axes=[]) == "pipe_00-data_00-model_00"
    assert topo.get_rank_repr(rank=3, omit_axes=[], inner_sep="+", outer_sep="|") == "pipe+00|data+01|model+01"
    topo = PipeDataParallelTopology(num_pp=2, num_dp=4)
    assert topo.get_rank_repr(rank=5) == ""
    assert topo.get_rank(pipe=1, data=2) == 6
    assert topo.get_axis_comm_lists("pipe") == [[0, 4], [1, 5], [2, 6], [3, 7]]
    assert topo.get_axis_comm_lists("data") == [[0, 1, 2, 3], [4, 5, 6, 7]]
    assert _prime_factors(12) == [2, 2, 3]


def test_partitioning_goldens():
    assert prefix_sum_inc([3, 4, 5]) == [3, 7, 12]
    assert partition_balanced([0, 1, 2, 3, 3, 3], 4) == [0, 3, 4, 5, 6]
    assert partition_balanced([0., 1.1, 1.9, 3., 3., 3.], 4) == [0, 3, 4, 5, 6]
    assert partition_balanced([0., 1.1, 30, 3.], 3) == [0, 2, 3, 4]
    parts = partition_uniform(8, 4)
    assert [parts[i + 1] - parts[i] for i in range(4)] == [2, 2, 2, 2]
    weights = [52559872] + [12596224] * 24 + [0, 52559872]
    parts = partition_balanced(weights, 8)
    assert parts[0] == 0 and parts[-1] == len(weights) and all(parts[i] <= parts[i + 1] for i in range(8))
    assert call_to_str("foo", 1, "a", k=2) == "foo(1, 'a', k=2)"


def test_train_schedule_golden_2stage_4mb():
    def names(stage):
        out = []
        for cmds in schedule.TrainSchedule(micro_batches=4, stages=2, stage_id=stage):
            out.append([f"{type(c).__name__}[{c.kwargs.get('buffer_id', '')}]" for c in cmds])
        return out
    s0, s1 = names(0), names(1)
    assert s0[0] == ["LoadMicroBatch[0]", "ForwardPass[0]"]
    assert s0[3] == ["SendActivation[1]", "RecvGrad[0]", "BackwardPass[0]"]
    assert s0[9] == ["RecvGrad[0]", "BackwardPass[0]", "ReduceTiedGrads[]", "ReduceGrads[]", "OptimizerStep[]"]
    assert s1[0] == []
    assert s1[1] == ["RecvActivation[0]", "LoadMicroBatch[0]", "ForwardPass[0]"]
    assert s1[5] == ["RecvActivation[0]", "SendGrad[1]", "LoadMicroBatch[0]", "ForwardPass[0]"]
    assert s1[9] == ["SendGrad[1]", "ReduceTiedGrads[]", "ReduceGrads[]", "OptimizerStep[]"]
    assert schedule.TrainSchedule(4, 2, 0).num_pipe_buffers() == 3
    assert schedule.TrainSchedule(4, 2, 1).num_pipe_buffers() == 2


def test_inference_schedule_single_stage():
    sched = schedule.InferenceSchedule(micro_batches=4, stages=1, stage_id=0)
    full = list(iter(sched))
    assert len(full) == 4
    for cmds in full:
        assert len(cmds) == 2
        assert isinstance(cmds[0], schedule.LoadMicroBatch) and isinstance(cmds[1], schedule.ForwardPass)
        assert cmds[0].buffer_id == cmds[1].buffer_id
    assert sched.num_pipe_buffers() == 2 and sched.is_first_stage and sched.is_last_stage


@pytest.mark.parametrize("mb", [1, 3, 8, 10])
def test_inference_schedule_pairs_send_recv(mb):
    stages = 3
    per_stage = [list(iter(schedule.InferenceSchedule(mb, stages, s))) for s in range(stages)]
    for step in range(mb + stages - 1):
        for s in range(stages - 1):
            sends = [c for c in per_stage[s][step] if isinstance(c, schedule.SendActivation)]
            recvs = [c for c in per_stage[s + 1][step] if isinstance(c, schedule.RecvActivation)]
            assert len(sends) == len(recvs)


def test_elasticity_goldens():
    base = {"elasticity": {"enabled": True, "max_train_batch_size": 10000, "micro_batch_sizes": [8, 12, 16, 17],
                           "min_gpus": 32, "max_gpus": 1500, "min_time": 20, "version": 0.1}}
    final_batch, valid = compute_elastic_config(base, "0.3.15")
    assert final_batch == 9792 and len(valid) == 23
    for g in valid:
        assert final_batch % g == 0 and any((final_batch // g) % mb == 0 for mb in [8, 12, 16, 17])
    with pytest.raises(ElasticityError):
        compute_elastic_config(base, "0.2")
    bad = json.loads(json.dumps(base)); bad["elasticity"]["version"] = 0.2
    with pytest.raises(ElasticityError):
        compute_elastic_config(bad, "0.3.15")
    bad = json.loads(json.dumps(base)); del bad["elasticity"]["max_train_batch_size"]
    with pytest.raises(ElasticityConfigError):
        compute_elastic_config(bad, "0.3.15")
    with pytest.raises(ElasticityIncompatibleWorldSize):
        compute_elastic_config(base, "0.3.15", world_size=7)
    fb, valid, mb = compute_elastic_config(base, "0.3.15", world_size=valid[0])
    assert (fb // valid[0]) % mb == 0


def test_dynamic_loss_scaler_semantics():
    s = DynamicLossScaler(init_scale=2 ** 8, scale_window=2, min_scale=1, delayed_shift=2)
    s.update_scale(True)       # hysteresis absorbs the first overflow
    assert s.cur_scale == 2 ** 8
    s.update_scale(True)
    assert s.cur_scale == 2 ** 7
    for _ in range(2):
        s.update_scale(False)
    assert s.cur_scale == 2 ** 8
    st = LossScaler(138.)
    st.update_scale(True)
    assert st.loss_scale == 138.
    m = DynamicLossScaler(init_scale=2, min_scale=1, delayed_shift=1)
    m.update_scale(True)
    with pytest.raises(Exception):
        m.update_scale(True)


def test_lr_schedules():
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([p], lr=0.1)
    w = lr_schedules.WarmupLR(opt, warmup_min_lr=0.0, warmup_max_lr=0.01, warmup_num_steps=10)
    seen = []
    for _ in range(15):
        w.step()
        seen.append(w.get_lr()[0])
    assert abs(seen[0] - 0.0) < 1e-12 and abs(seen[4] - 0.01 * math.log(5) / math.log(10)) < 1e-12
    assert all(abs(v - 0.01) < 1e-12 for v in seen[9:])
    d = lr_schedules.WarmupDecayLR(opt, total_num_steps=20, warmup_max_lr=0.01, warmup_num_steps=10)
    vals = []
    for _ in range(20):
        d.step()
        vals.append(d.get_lr()[0])
    assert abs(vals[14] - 0.01 * (20 - 14) / 10) < 1e-12 and vals[-1] <= vals[10]
    r = lr_schedules.LRRangeTest(opt, lr_range_test_min_lr=1e-3, lr_range_test_step_size=5, lr_range_test_step_rate=2.0,
                                 lr_range_test_staircase=True)
    r.step(); assert abs(r.get_lr()[0] - 1e-3) < 1e-12
    for _ in range(5):
        r.step()
    assert abs(r.get_lr()[0] - 1e-3 * 3) < 1e-12
    o = lr_schedules.OneCycle(opt, cycle_min_lr=0.01, cycle_max_lr=0.1, cycle_first_step_size=10, cycle_min_mom=0.8, cycle_max_mom=0.9)
    lrs, moms = [], []
    for _ in range(20):
        o.step()
        lrs.append(o.get_lr()[0]); moms.append(opt.param_groups[0]["betas"][0])
    assert max(lrs) <= 0.1 + 1e-9 and lrs.index(max(lrs)) in (8, 9, 10) and moms[lrs.index(max(lrs))] == min(moms)
    assert w.state_dict() == {"last_batch_iteration": 14}


def test_config_batch_triple_and_duplicates(tmp_path):
    assert DeepSpeedConfig(None, param_dict={"train_batch_size": 32, "train_micro_batch_size_per_gpu": 16,
                                             "gradient_accumulation_steps": 2}).train_batch_size == 32
    c = DeepSpeedConfig(None, param_dict={"train_batch_size": 32, "gradient_accumulation_steps": 4})
    assert c.train_micro_batch_size_per_gpu == 8
    with pytest.raises(AssertionError):
        DeepSpeedConfig(None, param_dict={"train_batch_size": 33, "train_micro_batch_size_per_gpu": 17,
                                          "gradient_accumulation_steps": 2})
    with pytest.raises(AssertionError):
        DeepSpeedConfig(None, param_dict={"train_batch_size": 4, "zero_optimization": {"stage": 1}})
    f = tmp_path / "dup.json"
    f.write_text('{"train_batch_size": 2, "train_batch_size": 4}')
    with pytest.raises(ValueError):
        DeepSpeedConfig(str(f))
    c = DeepSpeedConfig(None, param_dict={"train_batch_size": 2, "fp16": {"enabled": True, "type": "bfloat16"},
                                          "zero_optimization": True})
    assert c.precision == torch.bfloat16 and c.loss_scale == 1.0 and c.fp32_allreduce and c.zero_optimization_stage == 1
    c = DeepSpeedConfig(None, param_dict={"train_batch_size": 2, "fp16": {"enabled": True, "hysteresis": 4},
                                          "sparse_attention": {"mode": "bigbird", "block": 32}})
    assert c.dynamic_loss_scale_args["delayed_shift"] == 4 and c.sparse_attention["num_random_blocks"] == 1


def test_fused_zero_lamb_owned_segments_cover_the_owned_range():
    """Segment tables handed to the fused ZeRO LAMB kernel: every rank's owned range is tiled exactly, 8-aligned,
    and the union over ranks gives every tensor its full (padded) extent once."""
    import torch
    from deeperspeed_b200.parallel.fused_zero import FusedZeroStep
    from deeperspeed_b200.runtime.flat import FlatArena
    R, count = 3, 2
    params = [torch.nn.Parameter(torch.zeros(n)) for n in (5, 64, 1001, 8, 333)]
    arena = FlatArena(params, shard_count=R * count, shard_align=8)
    sub = arena.numel // (R * count)
    seen = {}
    for me in range(R):
        starts, tids = FusedZeroStep.owned_segments(arena, sub, count, R, me)
        assert starts[0] == 0 and starts[-1] == count * sub and len(starts) == len(tids) + 1
        assert all(s % 8 == 0 for s in starts) and all(a < b for a, b in zip(starts, starts[1:]))
        for (a, b), t in zip(zip(starts, starts[1:]), tids):
            assert a // sub == (b - 1) // sub, "a segment never crosses a bucket boundary"
            seen[t] = seen.get(t, 0) + (b - a)
    pad_tid = len(params)
    for ti, sg in enumerate(arena.segments):
        nxt = arena.segments[ti + 1].offset if ti + 1 < len(arena.segments) else arena.numel_used
        assert seen[ti] == nxt - sg.offset
    assert seen.get(pad_tid, 0) == arena.numel - arena.numel_used
