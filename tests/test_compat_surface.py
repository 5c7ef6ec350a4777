"""The public symbols users and checkpoints of the reference touch resolve through the `deepspeed`
alias to this package."""
import importlib
import os

import deeperspeed_b200 as deepspeed

SURFACE = {
    "deepspeed": ["initialize", "add_config_arguments", "init_distributed", "DeepSpeedEngine", "PipelineEngine",
                  "PipelineModule", "checkpointing", "zero", "__version__"],
    "deepspeed.pipe": ["LayerSpec", "TiedLayerSpec", "PipelineModule"],
    "deepspeed.runtime.pipe.schedule": ["TrainSchedule", "InferenceSchedule", "PipeSchedule", "DataParallelSchedule",
                                        "LoadMicroBatch", "ForwardPass", "BackwardPass", "SendActivation",
                                        "RecvActivation", "SendGrad", "RecvGrad", "ReduceGrads", "ReduceTiedGrads",
                                        "OptimizerStep"],
    "deepspeed.runtime.pipe.topology": ["ProcessTopology", "PipeDataParallelTopology", "PipeModelDataParallelTopology",
                                        "PipelineParallelGrid"],
    "deepspeed.checkpointing": ["checkpoint", "configure", "is_configured", "reset", "get_cuda_rng_tracker",
                                "model_parallel_cuda_manual_seed", "partition_activations_in_checkpoint",
                                "set_num_layers"],
    "deepspeed.zero": ["Init", "GatheredParameters", "register_external_parameter", "ZeroParamStatus", "ZeroParamType",
                       "TiledLinear", "TiledLinearReturnBias"],
    "deepspeed.ops.adam": ["FusedAdam", "DeepSpeedCPUAdam"],
    "deepspeed.ops.lamb": ["FusedLamb"],
    "deepspeed.ops.transformer": ["DeepSpeedTransformerLayer", "DeepSpeedTransformerConfig"],
    "deepspeed.ops.sparse_attention": ["SparseSelfAttention", "BertSparseSelfAttention", "SparseAttentionUtils", "MatMul",
                                       "Softmax", "SparsityConfig", "DenseSparsityConfig", "FixedSparsityConfig",
                                       "VariableSparsityConfig", "BigBirdSparsityConfig", "BSLongformerSparsityConfig",
                                       "LocalSlidingWindowSparsityConfig"],
    "deepspeed.ops.op_builder": ["CPUAdamBuilder", "FusedAdamBuilder", "FusedLambBuilder", "TransformerBuilder",
                                 "StochasticTransformerBuilder", "SparseAttnBuilder", "UtilsBuilder", "AsyncIOBuilder",
                                 "ALL_OPS"],
    "deepspeed.ops": ["__compatible_ops__"],
    "deepspeed.git_version_info": ["version", "git_hash", "git_branch", "installed_ops", "compatible_ops"],
    "deepspeed.runtime.fp16.loss_scaler": ["LossScaler", "DynamicLossScaler"],
    "deepspeed.runtime.zero.stage1": ["FP16_DeepSpeedZeroOptimizer_Stage1"],
    "deepspeed.runtime.zero.stage2": ["FP16_DeepSpeedZeroOptimizer"],
    "deepspeed.runtime.zero.stage3": ["FP16_DeepSpeedZeroOptimizer_Stage3"],
    "deepspeed.runtime.zero.config": ["DeepSpeedZeroConfig"],
    "deepspeed.runtime.zero.partition_parameters": ["Init", "GatheredParameters", "ZeroParamStatus"],
    "deepspeed.runtime.zero.linear": ["LinearModuleForZeroStage3", "LinearFunctionForZeroStage3"],
    "deepspeed.runtime.zero.tiling": ["TiledLinear"],
    "deepspeed.runtime.fp16.fused_optimizer": ["FP16_Optimizer"],
    "deepspeed.runtime.fp16.unfused_optimizer": ["FP16_UnfusedOptimizer"],
    "deepspeed.runtime.fp16.onebit.adam": ["OnebitAdam"],
    "deepspeed.runtime.fp16.onebit.lamb": ["OnebitLamb"],
    "deepspeed.runtime.comm.nccl": ["NcclBackend"],
    "deepspeed.runtime.comm.mpi": ["MpiBackend"],
    "deepspeed.runtime.lr_schedules": ["LRRangeTest", "OneCycle", "WarmupLR", "WarmupDecayLR", "add_tuning_arguments"],
    "deepspeed.runtime.utils": ["partition_uniform", "partition_balanced", "PartitionedTensor", "CheckOverflow",
                                "get_grad_norm", "get_weight_norm", "see_memory_usage", "memory_status", "call_to_str",
                                "clip_grad_norm_", "prefix_sum_inc", "set_random_seed", "move_to_device"],
    "deepspeed.runtime.config": ["DeepSpeedConfig"],
    "deepspeed.runtime.engine": ["DeepSpeedEngine"],
    "deepspeed.runtime.pipe.engine": ["PipelineEngine"],
    "deepspeed.runtime.pipe.module": ["PipelineModule", "LayerSpec", "TiedLayerSpec"],
    "deepspeed.runtime.pipe.p2p": ["send", "recv", "init_process_groups"],
    "deepspeed.runtime.dataloader": ["DeepSpeedDataLoader", "RepeatingLoader"],
    "deepspeed.runtime.csr_tensor": ["CSRTensor"],
    "deepspeed.runtime.progressive_layer_drop": ["ProgressiveLayerDrop"],
    "deepspeed.runtime.activation_checkpointing.checkpointing": ["checkpoint", "CheckpointFunction",
                                                                 "CudaRNGStatesTracker"],
    "deepspeed.runtime.swap_tensor.partitioned_param_swapper": ["AsyncPartitionedParameterSwapper"],
    "deepspeed.runtime.swap_tensor.partitioned_optimizer_swapper": ["PartitionedOptimizerSwapper"],
    "deepspeed.runtime.swap_tensor.pipelined_optimizer_swapper": ["PipelinedOptimizerSwapper"],
    "deepspeed.utils": ["logger", "log_dist", "RepeatingLoader", "init_distributed"],
    "deepspeed.elasticity": ["compute_elastic_config", "elasticity_enabled", "ensure_immutable_elastic_config"],
    "deepspeed.profiling.flops_profiler": ["FlopsProfiler", "get_model_profile"],
    "deepspeed.module_inject": ["replace_transformer_layer", "revert_transformer_layer"],
    "deepspeed.launcher.runner": ["main", "parse_args", "fetch_hostfile", "parse_resource_filter", "encode_world_info"],
    "deepspeed.launcher.launch": ["main"],
    "deepspeed.pt.deepspeed_light": ["DeepSpeedLight"],
    "deepspeed.pt.loss_scaler": ["DynamicLossScaler"],
    "deepspeed.ops.aio": ["AsyncIOBuilder", "aio_handle"],
}


def test_reference_symbol_surface():
    deepspeed.install_deepspeed_alias()
    missing = []
    for mod, names in SURFACE.items():
        try:
            m = importlib.import_module(mod)
        except Exception as e:  # noqa: BLE001
            missing.append((mod, repr(e)[:100]))
            continue
        missing += [(mod, n) for n in names if not hasattr(m, n)]
    assert not missing, missing


def test_cli_entry_points_exist():
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for exe in ("deepspeed", "deepspeed.pt", "ds", "ds_report", "ds_elastic", "ds_ssh"):
        path = os.path.join(root, "bin", exe)
        assert os.path.isfile(path) and os.access(path, os.X_OK), path


def test_reference_module_paths_and_helper_names(tmp_path):
    """Import paths and helper names that code written against the reference may touch (found by diffing the public
    names of every reference module against this package)."""
    import importlib
    import torch
    for mod in ("constants", "runtime.config_utils", "runtime.zero.constants", "runtime.zero.offload_constants",
                "runtime.zero.offload_config", "runtime.zero.utils", "runtime.activation_checkpointing.config",
                "profiling.constants", "profiling.config", "ops.adam.multi_tensor_apply"):
        importlib.import_module("deeperspeed_b200." + mod)
    from deeperspeed_b200.constants import default_pg_timeout
    assert default_pg_timeout.total_seconds() == 30 * 60
    from deeperspeed_b200.runtime.zero.constants import ZERO_OPTIMIZATION_REDUCE_BUCKET_SIZE_DEFAULT, ZERO_OPTIMIZATION_STAGE
    assert ZERO_OPTIMIZATION_REDUCE_BUCKET_SIZE_DEFAULT == int(5e8) and ZERO_OPTIMIZATION_STAGE == "stage"
    from deeperspeed_b200.runtime.zero.offload_config import get_offload_optimizer_config, get_offload_param_config
    cfg = get_offload_optimizer_config({"offload_optimizer": {"device": "nvme", "pipeline_read": True}})
    assert cfg["device"] == "nvme" and cfg["pipeline"] is True and cfg["buffer_count"] == 4
    assert get_offload_param_config({}) is None
    from deeperspeed_b200.runtime.zero.utils import is_zero_supported_optimizer
    p = torch.nn.Parameter(torch.zeros(2))
    assert is_zero_supported_optimizer(torch.optim.Adam([p])) and not is_zero_supported_optimizer(torch.optim.SGD([p], lr=0.1))
    from deeperspeed_b200.runtime.config import DeepSpeedConfigWriter
    w = DeepSpeedConfigWriter()
    w.add_config("train_batch_size", 8)
    w.write_config(str(tmp_path / "c.json"))
    w2 = DeepSpeedConfigWriter()
    w2.load_config(str(tmp_path / "c.json"))
    assert w2.data == {"train_batch_size": 8}
    from deeperspeed_b200.runtime.constants import ValidationMode
    assert (ValidationMode.WARN, ValidationMode.IGNORE, ValidationMode.FAIL) == ("WARN", "IGNORE", "FAIL")
    from deeperspeed_b200.runtime.fp16.fused_optimizer import is_fp16_fused_supported_optimizer
    assert is_fp16_fused_supported_optimizer(torch.optim.Adam([p]))
    from deeperspeed_b200.runtime.lr_schedules import override_1cycle_params, override_warmupLR_params
    import argparse
    params = {}
    override_1cycle_params(argparse.Namespace(cycle_min_lr=0.01, cycle_max_lr=None), params)
    assert params == {"cycle_min_lr": 0.01}
    override_warmupLR_params(argparse.Namespace(warmup_num_steps=7), params)
    assert params["warmup_num_steps"] == 7
    from deeperspeed_b200.runtime.utils import GradientNoiseScale
    gns = GradientNoiseScale.__new__(GradientNoiseScale)
    gns.beta = 0.9
    raw, corrected = gns.ema(None, 2.0, 0)
    assert abs(raw - 0.2) < 1e-9 and abs(corrected - 2.0) < 1e-9
    from deeperspeed_b200.ops.adam.multi_tensor_apply import MultiTensorApply
    assert MultiTensorApply(64)(lambda chunk, flag, lists, x: (chunk, x), None, [[]], 5) == (64, 5)
    from deeperspeed_b200.ops.module_inject import revert_module, replace_module  # noqa: F401
    from deeperspeed_b200.env_report import ninja_installed
    assert isinstance(ninja_installed(), bool)
    from deeperspeed_b200.pipe import PipelineModule, LayerSpec
    assert hasattr(PipelineModule, "set_checkpoint_interval")


def test_config_free_getters_agree_with_the_config_object():
    """The reference's `get_<key>(param_dict)` free functions (runtime/config.py:45-516) exist and return what
    DeepSpeedConfig parses."""
    from deeperspeed_b200.runtime import config as c
    d = {"train_batch_size": 8, "fp16": {"enabled": True, "loss_scale": 0, "initial_scale_power": 10},
         "optimizer": {"type": "Adam", "params": {"lr": 1e-3, "max_grad_norm": 1.0}},
         "scheduler": {"type": "WarmupLR", "params": {"warmup_num_steps": 5}},
         "zero_optimization": {"stage": 2}, "gradient_clipping": 0.5, "checkpoint": {"tag_validation": "fail"},
         "tensorboard": {"enabled": True, "output_path": "/tmp/tb"}, "steps_per_print": 7}
    cfg = c.DeepSpeedConfig(None, param_dict=dict(d))
    assert c.get_fp16_enabled(d) is True and c.get_fp16_type(d) == "fp16"
    assert c.get_initial_dynamic_scale(d) == 1024 == cfg.initial_dynamic_scale
    assert c.get_dynamic_loss_scale_args(d) == cfg.dynamic_loss_scale_args
    assert c.get_optimizer_name(d) == "adam" and c.get_optimizer_params(d)["lr"] == 1e-3
    assert c.get_optimizer_gradient_clipping(d) == 1.0 and c.get_gradient_clipping(d) == 0.5
    assert c.get_scheduler_name(d) == "WarmupLR" and c.get_scheduler_params(d) == {"warmup_num_steps": 5}
    assert c.get_zero_optimization(d) is True and c.get_zero_reduce_scatter(d) is True
    assert c.get_checkpoint_tag_validation_mode(c.get_checkpoint_params(d)) == "FAIL"
    assert c.get_tensorboard_enabled(d) and c.get_tensorboard_output_path(d) == "/tmp/tb"
    assert c.get_steps_per_print(d) == 7 and c.get_train_batch_size(d) == 8
    assert c.get_pld_enabled(d) is False and c.get_amp_enabled(d) is False and c.get_sparse_gradients_enabled(d) is False
    assert c.get_sparse_fixed_config({"block": 32})["block"] == 32
    assert c.get_sparse_bigbird_config({})["num_sliding_window_blocks"] == 3


def test_subpackages_reachable_after_a_bare_import():
    """`import deepspeed; deepspeed.elasticity.compute_elastic_config(...)` works in the reference without importing
    the sub-package first (its engine / config modules pull them in); the same must hold here."""
    import subprocess
    import sys
    code = ("import deeperspeed_b200 as d\n"
            "d.elasticity.compute_elastic_config; d.elasticity.config.ElasticityError\n"
            "d.profiling.flops_profiler.FlopsProfiler; d.module_inject.replace_module\n"
            "d.ops.adam.FusedAdam; d.ops.sparse_attention.SparseSelfAttention; d.runtime.zero.Init\n"
            "d.pipe.LayerSpec; d.checkpointing.checkpoint; d.git_version_info.version; d.constants.TORCH_DISTRIBUTED_DEFAULT_PORT\n"
            "d.utils.logger; d.runtime.lr_schedules.OneCycle; d.runtime.pipe.schedule.TrainSchedule\n")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    done = subprocess.run([sys.executable, "-c", code], cwd=root, capture_output=True, text=True)
    assert done.returncode == 0, done.stderr[-800:]
