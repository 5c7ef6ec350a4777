"""The public symbols users and checkpoints of the reference touch resolve through the `deepspeed`
alias to this package."""
import importlib

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
    for exe in ("deepspeed", "ds", "ds_report", "ds_elastic", "ds_ssh"):
        path = os.path.join(root, "bin", exe)
        assert os.path.isfile(path) and os.access(path, os.X_OK), path
