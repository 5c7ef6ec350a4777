"""Smaller components: launcher parsing, CSR tensors, RepeatingLoader, PLD, flops profiler, activation checkpointing,
optimizer selection matrix, argument parsing, deepspeed alias, batch-size scheduler, checkpoint tag validation."""
import argparse
import os
import pickle

import pytest
import torch

import deeperspeed_b200 as deepspeed
from common import distributed_test
from simple_model import SimpleModel, args_from_dict, random_dataloader


def test_runner_filters_and_commands():
    from deeperspeed_b200.launcher import runner as R
    from deeperspeed_b200.launcher.multinode_runner import PDSHRunner, OpenMPIRunner, SlurmRunner
    hosts = {"worker-0": [0, 1, 2, 3], "worker-1": [0, 1, 2, 3]}
    assert R.parse_resource_filter(hosts, include_str="worker-1:0,2") == {"worker-1": [0, 2]}
    assert R.parse_resource_filter(hosts, exclude_str="worker-0") == {"worker-1": [0, 1, 2, 3]}
    with pytest.raises(ValueError):
        R.parse_resource_filter(hosts, include_str="worker-1", exclude_str="worker-0")
    with pytest.raises(ValueError):
        R.parse_resource_filter(hosts, include_str="worker-9")
    with pytest.raises(ValueError):
        R.parse_resource_filter(hosts, include_str="worker-1:7")
    args = R.parse_args(["--num_nodes", "1", "--num_gpus", "2", "--master_addr", "10.0.0.1", "train.py", "--foo", "bar"])
    assert args.user_script == "train.py" and args.user_args == ["--foo", "bar"]
    wi = R.encode_world_info(hosts)
    cmd = PDSHRunner(args, wi).get_cmd({}, hosts)
    assert cmd[0] == "pdsh" and "--node_rank=%n" in cmd and any("launcher.launch" in c for c in cmd)
    args.hostfile = "/job/hostfile"
    cmd = OpenMPIRunner(R.parse_args(["train.py"]), wi, {"a": 4, "b": 4}).get_cmd({}, hosts)
    assert cmd[:3] == ["mpirun", "-n", "8"]
    sargs = R.parse_args(["--launcher", "slurm", "--comment", "neox", "train.py"])
    cmd = SlurmRunner(sargs, wi, {"a": 4, "b": 4}).get_cmd({}, hosts)
    assert cmd[:3] == ["srun", "-n", "8"] and "--comment" in cmd and cmd[cmd.index("--comment") + 1] == "neox"


def test_hostfile(tmp_path):
    from deeperspeed_b200.launcher.runner import fetch_hostfile
    f = tmp_path / "hostfile"
    f.write_text("worker-0 slots=4\n# c\nworker-1 slots=2\n")
    assert fetch_hostfile(str(f)) == {"worker-0": 4, "worker-1": 2}
    f.write_text("worker-0 slots=4\nworker-0 slots=2\n")
    with pytest.raises(ValueError):
        fetch_hostfile(str(f))
    assert fetch_hostfile(str(tmp_path / "nope")) is None


def test_gpu_topology_parse():
    from deeperspeed_b200.launcher.gpu_topology import detect_nvlink_pairs
    text = ("\tGPU0\tGPU1\tGPU2\tGPU3\tCPU Affinity\n"
            "GPU0\t X \tPIX\tNV2\tSYS\t0-15\nGPU1\tPIX\t X \tSYS\tNV2\t0-15\n"
            "GPU2\tNV2\tSYS\t X \tPIX\t0-15\nGPU3\tSYS\tNV2\tPIX\t X \t0-15\n")
    assert detect_nvlink_pairs(text) == [0, 2, 1, 3]


def test_gpu_topology_reference_names():
    """The reference's launcher/gpu_topology.py function names on top of the same parser."""
    from deeperspeed_b200.launcher import gpu_topology as g
    txt = ("        GPU0    GPU1    GPU2    GPU3    CPU Affinity\n"
           "GPU0     X      SYS     NV2     SYS     0-15\n"
           "GPU1    SYS      X      SYS     NV1     0-15\n"
           "GPU2    NV2     SYS      X      SYS     16-31\n"
           "GPU3    SYS     NV1     SYS      X      16-31\n\nLegend:\n  X = Self\n")
    topo = g.topology_from_string(txt)
    assert topo[0] == ["X", "SYS", "NV2", "SYS"] and len(topo) == 4
    assert g.contains_nvlinks(topo) and g.is_nvlink("NV18") and not g.is_nvlink("SYS")
    assert sorted(g.get_nvlink_pairs(topo)) == [(0, 2), (1, 3)]
    assert g.get_cuda_visible_device_mapping(g.get_nvlink_pairs(topo)) == "0,2,1,3"
    assert g.get_cuda_visible_device_mapping(g.get_nvlink_pairs(topo), local_gpu_ids=[1, 3]) == "1,3"
    assert not g.contains_nvlinks([["X", "SYS"], ["SYS", "X"]])


def test_csr_and_repeating_loader_and_pld():
    from deeperspeed_b200.runtime.csr_tensor import CSRTensor
    from deeperspeed_b200.runtime.progressive_layer_drop import ProgressiveLayerDrop
    import math
    dense = torch.zeros(10, 4)
    dense[3] = 1.0
    dense[7] = 2.0
    c = CSRTensor(dense)
    assert c.indices.tolist() == [3, 7] and torch.equal(c.to_dense(), dense)
    c.add(CSRTensor(dense))
    assert torch.equal(c.to_dense(), 2 * dense)
    loader = deepspeed.utils.RepeatingLoader([1, 2, 3])
    assert [next(loader) for _ in range(7)] == [1, 2, 3, 1, 2, 3, 1]
    pld = ProgressiveLayerDrop(theta=0.5, gamma=0.01)
    pld.update_state(100)
    assert abs(pld.get_theta() - (0.5 * math.exp(-1.0) + 0.5)) < 1e-9
    assert pld.get_state() == {"progressive_layer_drop": True, "pld_theta": pld.get_theta()}


def test_flops_profiler_lenet():
    import torch.nn as nn
    import torch.nn.functional as F
    from deeperspeed_b200.profiling.flops_profiler import get_model_profile

    class LeNet5(nn.Module):
        def __init__(self):
            super().__init__()
            self.f = nn.Sequential(nn.Conv2d(1, 6, 5), nn.Tanh(), nn.AvgPool2d(2), nn.Conv2d(6, 16, 5), nn.Tanh(),
                                   nn.AvgPool2d(2), nn.Conv2d(16, 120, 5), nn.Tanh())
            self.c = nn.Sequential(nn.Linear(120, 84), nn.Tanh(), nn.Linear(84, 10))

        def forward(self, x):
            return F.softmax(self.c(torch.flatten(self.f(x), 1)), dim=1)

    macs, params = get_model_profile(LeNet5(), (1024, 1, 32, 32), print_profile=False)
    assert macs == "439.56 MMACs" and params == "61.71 k"


def test_activation_checkpointing_matches_plain():
    ckpt = deepspeed.checkpointing
    torch.manual_seed(0)
    lin = torch.nn.Sequential(torch.nn.Linear(16, 16), torch.nn.Dropout(0.3), torch.nn.Linear(16, 16))
    x = torch.randn(4, 16, requires_grad=True)
    mask = torch.ones(4, 16, dtype=torch.bool)

    def fn(a, m, flag):
        out = lin(a)
        return out * m, (1 if flag else 0)

    torch.manual_seed(5)
    y_ref, _ = fn(x, mask, True)
    y_ref.sum().backward()
    gref = [p.grad.clone() for p in lin.parameters()]
    xg = x.grad.clone()
    for p in lin.parameters():
        p.grad = None
    x.grad = None
    torch.manual_seed(5)
    out = ckpt.checkpoint(fn, x, mask, True)
    assert isinstance(out, tuple) and out[1] == 1
    out[0].sum().backward()
    assert torch.allclose(out[0], y_ref) and torch.allclose(x.grad, xg)
    for p, g in zip(lin.parameters(), gref):
        assert torch.allclose(p.grad, g)
    ckpt.configure(None, partition_activations=False, profile=False)
    assert ckpt.is_configured()
    ckpt.get_cuda_rng_tracker().reset()
    ckpt.model_parallel_cuda_manual_seed(123)
    with ckpt.get_cuda_rng_tracker().fork():
        a = torch.rand(3)
    with ckpt.get_cuda_rng_tracker().fork():
        b = torch.rand(3)
    assert not torch.equal(a, b)


def test_optimizer_selection_matrix(tmp_path):
    @distributed_test(world_size=1)
    def run():
        from deeperspeed_b200.ops.adam import FusedAdam, DeepSpeedCPUAdam
        cases = [("Adam", {}, False, FusedAdam, 1), ("AdamW", {}, False, FusedAdam, 1),
                 ("Adam", {"adam_w_mode": False}, False, FusedAdam, 0),
                 ("Adam", {"torch_adam": True}, False, torch.optim.AdamW, None),
                 ("Adam", {"torch_adam": True, "adam_w_mode": False}, False, torch.optim.Adam, None),
                 ("Adam", {}, True, DeepSpeedCPUAdam, True), ("Lamb", {}, False, None, None)]
        for name, extra, offload, cls, wmode in cases:
            cfg = {"train_batch_size": 2, "steps_per_print": 100,
                   "optimizer": {"type": name, "params": dict({"lr": 0.001}, **extra)},
                   "fp16": {"enabled": True, "type": "bfloat16"}}
            if offload:
                cfg["zero_optimization"] = {"stage": 2, "cpu_offload": True}
            model = SimpleModel(8)
            engine, opt, _, _ = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                     model_parameters=model.parameters())
            inner = opt.optimizer
            if cls is None:
                assert type(opt).__name__ == "FP16_UnfusedOptimizer" and type(inner).__name__ == "FusedLamb"
            else:
                assert type(inner) is cls, (name, extra, type(inner))
                if wmode is not None:
                    assert inner.adam_w_mode == wmode
    run()


def test_add_config_arguments_and_alias():
    parser = deepspeed.add_config_arguments(argparse.ArgumentParser())
    args = parser.parse_args(["--deepspeed", "--deepspeed_config", "foo.json"])
    assert args.deepspeed and args.deepspeed_config == "foo.json" and args.deepscale_config is None
    deepspeed.install_deepspeed_alias()
    import deepspeed as ds_alias
    from deepspeed.runtime.fp16.loss_scaler import DynamicLossScaler
    from deepspeed.runtime.pipe.schedule import TrainSchedule
    from deepspeed.pipe import PipelineModule, LayerSpec
    assert ds_alias.initialize is deepspeed.initialize
    s = pickle.loads(pickle.dumps(DynamicLossScaler(init_scale=8)))
    assert s.cur_scale == 8
    import deepspeed.pt.loss_scaler as legacy
    assert legacy.DynamicLossScaler is DynamicLossScaler


def test_bs_scheduler_and_gns():
    from deeperspeed_b200.runtime.bs_schedules import BatchSizeScheduler
    s = BatchSizeScheduler(final_batch_size=64, start_batch_size=8, num_intervals=7, steps_per_interval=10,
                           min_batch_size_multiplier=8)
    seen = [s.step() for _ in range(75)]
    assert seen[0] == 8 and seen[-1] == 64 and sorted(seen) == seen
    from deeperspeed_b200.runtime.utils import GradientNoiseScale
    m = torch.nn.Linear(4, 4)
    gns = GradientNoiseScale(m, batch_size_small=2, n_batches=2)
    for _ in range(6):
        m.zero_grad()
        m(torch.randn(2, 4)).sum().backward()
        gns.update()
    assert gns.n_updates == 6


def test_engine_lr_scheduler_and_tag_validation(tmp_path):
    @distributed_test(world_size=2)
    def run():
        import torch.distributed as dist
        cfg = {"train_batch_size": 4, "steps_per_print": 100, "optimizer": {"type": "Adam", "params": {"lr": 0.01}},
               "scheduler": {"type": "WarmupLR", "params": {"warmup_min_lr": 0, "warmup_max_lr": 0.01, "warmup_num_steps": 4}},
               "checkpoint": {"tag_validation": "FAIL"}}
        model = SimpleModel(8)
        engine, _, _, sched = deepspeed.initialize(args=args_from_dict(tmp_path, cfg), model=model,
                                                   model_parameters=model.parameters())
        assert type(sched).__name__ == "WarmupLR"
        for x, y in random_dataloader(engine, 16, 8, engine.device):
            loss = engine(x, y); engine.backward(loss); engine.step()
        assert engine.get_lr()[0] == pytest.approx(0.01)
        with pytest.raises(AssertionError):
            engine.save_checkpoint(str(tmp_path), tag=f"rank{dist.get_rank()}")
        engine.save_checkpoint(str(tmp_path), tag="same")
        dist.barrier()   # rank 0 writes `latest` after the collective part of save_checkpoint
        if dist.get_rank() == 0:
            assert open(os.path.join(str(tmp_path), "latest")).read() == "same"
        assert engine.load_checkpoint(str(tmp_path / "missing")) == (None, None)
    run()


def test_timer_chrome_trace_export(tmp_path):
    """Timer intervals can be dumped as a Chrome / Perfetto trace (device timestamps on CUDA, host time here)."""
    import json
    import time
    from deeperspeed_b200.utils.timer import SynchronizedWallClockTimer
    timers = SynchronizedWallClockTimer()
    timers.enable_trace()
    for _ in range(3):
        timers("forward").start()
        time.sleep(0.002)
        timers("forward").stop()
        timers("step").start()
        timers("step").stop()
    assert timers("forward").elapsed(reset=True) >= 0.006
    path = tmp_path / "trace.json"
    n = timers.export_chrome_trace(str(path), rank=1)
    ev = json.load(open(path))["traceEvents"]
    spans = [e for e in ev if e["ph"] == "X"]
    assert n == len(ev) and len(spans) == 6
    assert {e["name"] for e in spans} == {"forward", "step"} and all(e["pid"] == 1 for e in spans)
    fw = sorted(e["ts"] for e in spans if e["name"] == "forward")
    assert fw[0] < fw[1] < fw[2] and all(e["dur"] >= 0 for e in spans)


def test_launch_spawns_ranks_and_propagates_failure(tmp_path):
    """launcher/launch.py end to end on this host: two local ranks rendezvous over gloo with the environment the
    launcher sets; when one rank fails, its sibling is killed and the exit code is propagated
    (reference launcher/launch.py:139-175)."""
    import base64
    import json
    import subprocess
    import sys
    import time
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ok_script = tmp_path / "ok.py"
    ok_script.write_text(
        "import os, torch, torch.distributed as dist\n"
        "dist.init_process_group('gloo')\n"
        "t = torch.tensor([float(os.environ['RANK']) + 1.0])\n"
        "dist.all_reduce(t)\n"
        "assert int(os.environ['WORLD_SIZE']) == 2 and os.environ['LOCAL_RANK'] == os.environ['RANK']\n"
        "open(os.path.join(os.path.dirname(__file__), 'out%s.txt' % os.environ['RANK']), 'w').write(str(t.item()))\n")
    bad_script = tmp_path / "bad.py"
    bad_script.write_text(
        "import os, sys, time\n"
        "if os.environ['RANK'] == '1':\n"
        "    sys.exit(7)\n"
        "time.sleep(120)\n")
    world = base64.urlsafe_b64encode(json.dumps({"localhost": [0, 1]}).encode()).decode()
    env = dict(os.environ, PYTHONPATH=repo + os.pathsep + os.environ.get("PYTHONPATH", ""))
    base = [sys.executable, "-u", "-m", "deeperspeed_b200.launcher.launch", f"--world_info={world}",
            "--master_addr=127.0.0.1"]
    r = subprocess.run(base + ["--master_port=29677", str(ok_script)], env=env, cwd=repo, capture_output=True, timeout=120)
    assert r.returncode == 0, r.stderr.decode()[-2000:]
    assert (tmp_path / "out0.txt").read_text() == "3.0" and (tmp_path / "out1.txt").read_text() == "3.0"
    t0 = time.time()
    r = subprocess.run(base + ["--master_port=29678", str(bad_script)], env=env, cwd=repo, capture_output=True, timeout=100)
    assert r.returncode == 7, (r.returncode, r.stderr.decode()[-2000:])
    assert time.time() - t0 < 60, "the surviving rank was not killed"


def test_deepspeed_cli_single_node_end_to_end(tmp_path):
    """bin/deepspeed -> runner -> launch -> two ranks that train a tiny model through deepspeed.initialize with ZeRO-1
    (on a CPU-only host --num_gpus is the number of local processes; the reference's single-node path dies on an
    undefined `runner`)."""
    import subprocess
    import sys
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "train.py"
    script.write_text(
        "import argparse, json, os, torch\n"
        "import deeperspeed_b200 as deepspeed\n"
        "p = argparse.ArgumentParser(); p.add_argument('--local_rank', type=int, default=-1); p.add_argument('--steps', type=int)\n"
        "p = deepspeed.add_config_arguments(p); a = p.parse_args()\n"
        "torch.manual_seed(0)\n"
        "model = torch.nn.Linear(8, 8)\n"
        "cfg = {'train_batch_size': 4, 'train_micro_batch_size_per_gpu': 2, 'steps_per_print': 100,\n"
        "       'optimizer': {'type': 'Adam', 'params': {'lr': 0.01}}, 'fp16': {'enabled': True, 'type': 'fp32'},\n"
        "       'zero_optimization': {'stage': 1}}\n"
        "engine, _, _, _ = deepspeed.initialize(args=a, model=model, model_parameters=model.parameters(), config_params=cfg)\n"
        "for _ in range(a.steps):\n"
        "    x = torch.randn(2, 8)\n"
        "    loss = engine(x).pow(2).mean(); engine.backward(loss); engine.step()\n"
        "w = torch.cat([q.detach().view(-1) for q in model.parameters()])\n"
        "import torch.distributed as dist\n"
        "ref = w.clone(); dist.broadcast(ref, src=0)\n"
        "assert torch.equal(w, ref) and engine.global_steps == a.steps\n"
        "open(os.path.join(os.path.dirname(__file__), 'done%d' % dist.get_rank()), 'w').write('ok')\n")
    env = dict(os.environ, PYTHONPATH=repo + os.pathsep + os.environ.get("PYTHONPATH", ""))
    r = subprocess.run([sys.executable, os.path.join(repo, "bin", "deepspeed"), "--num_gpus", "2", "--master_port", "29689",
                        str(script), "--steps", "3"], env=env, cwd=str(tmp_path), capture_output=True, timeout=300)
    assert r.returncode == 0, r.stderr.decode()[-3000:]
    assert (tmp_path / "done0").exists() and (tmp_path / "done1").exists()


def test_reference_helper_names_behave():
    """Small public helpers the reference exposes: the 24-bit all-reduce's decompose / reconstruct, the pipeline
    visualiser's `expand` and table, zero_to_fp32's shard arithmetic."""
    from deeperspeed_b200.runtime.comm import compressed_ar as car
    from deeperspeed_b200.runtime.pipe import pipe_visualizer as pv, schedule
    from deeperspeed_b200.utils.zero_to_fp32 import zero3_partitioned_param_info
    x = torch.tensor([0.75, -3.0, 1e-3, 120.0])
    mant, exp = car.decompose(x)
    assert mant.dtype == torch.float16 and exp.dtype == torch.int8
    assert torch.allclose(car.reconstruct(mant, exp, torch.float32), x, rtol=1e-3)
    assert car.torch2cupy(x) is x and car.compressed_all_reduce_cupy is car.compressed_all_reduce_torch
    steps = list(schedule.TrainSchedule(micro_batches=4, stages=2, stage_id=0).steps())
    labels = pv.expand(steps)
    assert labels[0] == "fwd_1" and sum(l is not None for l in labels) == 8          # 4 forward + 4 backward passes
    assert any(l and l.startswith("bwd_") for l in labels)
    full = pv.expand(steps, include_all=True)
    assert "load_batch_1" in full[0] and full[-1].endswith("step")
    table = pv.pipeline_table(2, 4)
    assert "GPU 1" in table and "Num Microbatches: 4" in table and "Non Idle Time: 16" in table
    assert zero3_partitioned_param_info(10, 4) == (2, 2) and zero3_partitioned_param_info(8, 4) == (2, 0)


def test_zero3_fused_dgrad_choice_follows_world_size(monkeypatch):
    """Unset, the K-sharded input-gradient kernel is used between two GPUs and NCCL gathers from four on; the env
    switch overrides either way."""
    from deeperspeed_b200.runtime.zero.stage3 import FP16_DeepSpeedZeroOptimizer_Stage3 as Z3

    class Probe:
        _fused_dgrad = Z3._fused_dgrad

    monkeypatch.delenv("DSB_ZERO3_FUSED_DGRAD", raising=False)
    for world, want in ((1, True), (2, True), (4, False), (8, False)):
        p = Probe()
        p.partition_count = world
        assert p._fused_dgrad is want
    p.partition_count = 8
    monkeypatch.setenv("DSB_ZERO3_FUSED_DGRAD", "1")
    assert p._fused_dgrad is True
    p.partition_count = 2
    monkeypatch.setenv("DSB_ZERO3_FUSED_DGRAD", "0")
    assert p._fused_dgrad is False


def test_version_and_checkout_attributes():
    """reference tests/unit/test_config.py test_check_version; the hash is the checkout's when there is one."""
    import re
    import subprocess
    for name in ("__git_hash__", "__git_branch__", "__version__", "__version_major__", "__version_minor__",
                 "__version_patch__"):
        assert hasattr(deepspeed, name), name
    assert deepspeed.__version__ == f"{deepspeed.__version_major__}.{deepspeed.__version_minor__}.{deepspeed.__version_patch__}"
    root = os.path.dirname(os.path.dirname(os.path.abspath(deepspeed.__file__)))
    if os.path.isdir(os.path.join(root, ".git")):
        head = subprocess.run(["git", "-C", root, "rev-parse", "--short=7", "HEAD"], capture_output=True, text=True)
        if head.returncode == 0:
            from deeperspeed_b200.version import _checkout_state
            assert _checkout_state()[0] == head.stdout.strip()      # read now: HEAD may have moved since the import
            assert re.fullmatch(r"[0-9a-f]{7}", deepspeed.__git_hash__)
    else:
        assert deepspeed.__git_branch__ == "[snapshot]"
    from deeperspeed_b200 import git_version_info as info
    assert info.version == deepspeed.__version__ and re.match(r"\d+\.\d+$", info.torch_info["version"])
    assert set(info.installed_ops) == set(info.compatible_ops) and info.installed_ops


@pytest.mark.parametrize("argv, enabled, config", [
    ([], False, None),                                               # parser extended, flags absent
    (["--deepspeed_config", "foo.json"], False, "foo.json"),         # config without the enable flag
    (["--deepspeed"], True, None),                                   # enable flag without a config
    (["--deepspeed", "--deepspeed_config", "foo.json"], True, "foo.json"),
])
def test_config_argument_combinations(argv, enabled, config):
    """reference tests/unit/test_ds_arguments.py: every combination of the two core flags next to a user flag."""
    plain = argparse.ArgumentParser()
    plain.add_argument("--num_epochs", type=int)
    ns = plain.parse_args(["--num_epochs", "2"])
    assert not hasattr(ns, "deepspeed") and not hasattr(ns, "deepspeed_config")
    with pytest.raises(SystemExit):
        plain.parse_args(["--num_epochs", "2", "--deepspeed"])      # unknown without add_config_arguments
    parser = deepspeed.add_config_arguments(plain)
    ns = parser.parse_args(["--num_epochs", "2"] + argv)
    assert ns.num_epochs == 2
    assert type(ns.deepspeed) is bool and ns.deepspeed == enabled
    assert ns.deepspeed_config == config


def test_batch_size_scheduler_reference_contract():
    """BatchSizeScheduler with the reference's constructor (bs_schedules.py:12-18): goldens produced by running the
    reference class itself - schedule keys / values and the value at selected iterations."""
    from deeperspeed_b200.runtime.bs_schedules import BatchSizeScheduler
    s = BatchSizeScheduler(final_batch_size=16, num_intervals=8, warmup_num_steps=10000)
    assert s.schedule == {0: 1, 1428: 3, 2857: 5, 4285: 7, 5714: 9, 7142: 11, 8571: 13, 10000: 16}
    assert BatchSizeScheduler(final_batch_size=1024).schedule == {0: 11, 333: 348, 666: 686, 1000: 1024}
    s = BatchSizeScheduler(final_batch_size=512, min_batch_size_multiplier=0.1, warmup_num_steps=300, num_intervals=7)
    assert s.schedule == {0: 52, 50: 128, 100: 205, 150: 282, 200: 358, 250: 435, 300: 512}
    seen = {}
    for it in range(320):
        seen[it] = s.step()
        assert s.current_batch_size == seen[it]
    assert (seen[0], seen[49], seen[50], seen[299], seen[300], seen[319]) == (52, 52, 128, 435, 512, 512)
    assert BatchSizeScheduler(final_batch_size=3, num_intervals=4, warmup_num_steps=10).schedule == {0: 1, 6: 2, 10: 3}
    restored = BatchSizeScheduler(final_batch_size=512, min_batch_size_multiplier=0.1, warmup_num_steps=300, num_intervals=7)
    restored.load_state_dict({"last_batch_iteration": 149})
    assert restored.step() == 282

    class Engine:
        sizes = []

        def set_train_batch_size(self, size):
            self.sizes.append(size)

    driven = BatchSizeScheduler(final_batch_size=8, min_batch_size_multiplier=0.5, warmup_num_steps=4, num_intervals=3,
                                deepspeed=Engine())
    for _ in range(8):
        driven.step()
    assert Engine.sizes == [4, 6, 8]                      # told once per change


def test_reference_argument_names_of_small_helpers():
    """Keyword names callers of the reference use: `model_mbeddings` (sic) of pad_to_block_size, `b` of CSRTensor.add,
    `deepspeed` of CheckOverflow, the three names of get_scalar_param, `tensor` of decompose_cupy, `return_values` of
    SynchronizedWallClockTimer.log, `num_microbatches` / `include_all` of pipeline_visualizer, and the argument order
    of `deepspeed.ops.replace_module`."""
    import inspect
    from deeperspeed_b200.ops.sparse_attention import SparseAttentionUtils
    from deeperspeed_b200.runtime.csr_tensor import CSRTensor
    from deeperspeed_b200.runtime.utils import CheckOverflow
    from deeperspeed_b200.runtime.config_utils import get_scalar_param
    from deeperspeed_b200.runtime.comm.compressed_ar import decompose_cupy
    from deeperspeed_b200.runtime.pipe.pipe_visualizer import pipeline_visualizer
    from deeperspeed_b200.utils.timer import SynchronizedWallClockTimer
    from deeperspeed_b200.module_inject import replace_transformer_layer

    def names(fn):
        return list(inspect.signature(fn).parameters)

    assert names(SparseAttentionUtils.pad_to_block_size)[-1] == "model_mbeddings"
    emb = torch.nn.Embedding(10, 4)
    ids = torch.randint(0, 10, (2, 5))
    pad, ids2, mask2, *_ = SparseAttentionUtils.pad_to_block_size(block_size=4, input_ids=ids,
                                                                 attention_mask=torch.ones(2, 5), token_type_ids=None,
                                                                 position_ids=None, inputs_embeds=None, pad_token_id=0,
                                                                 model_mbeddings=emb)
    assert pad == 3 and ids2.shape == (2, 8) and mask2.shape == (2, 8)
    assert names(CSRTensor.add) == ["self", "b"]
    assert CheckOverflow(deepspeed="engine").deepspeed == "engine"
    assert get_scalar_param(param_dict={"a": 1}, param_name="a", param_default_value=0) == 1
    mantissa, exponent = decompose_cupy(tensor=torch.tensor([0.5, 8.0]))
    assert mantissa.shape == exponent.shape == (2,)
    timers = SynchronizedWallClockTimer()
    timers("x").start(); timers("x").stop()
    values = timers.log(["x", "never_started"], return_values=True)
    assert set(values) == {"x"} and values["x"] >= 0 and timers.log(["x"], return_values=False) is None
    text = pipeline_visualizer(num_stages=2, num_microbatches=2, include_all=False)
    assert "Idle Time: 4" in text and "Non Idle Time: 8" in text
    assert names(deepspeed.ops.replace_module) == ["orig_module_impl", "model", "replacement_module_config"]
    assert names(replace_transformer_layer)[:10] == ["orig_layer_impl", "model", "micro_batch_size", "bert_config", "seed",
                                                     "preln", "fp16", "training", "huggingface", "local_rank"]
