"""JSON config keys and defaults (schema parity with reference deepspeed/runtime/constants.py,
deepspeed/runtime/zero/constants.py, activation_checkpointing/config.py, profiling/constants.py,
swap_tensor/constants.py, elasticity/constants.py)."""

TRAIN_BATCH_SIZE = "train_batch_size"
TRAIN_MICRO_BATCH_SIZE_PER_GPU = "train_micro_batch_size_per_gpu"
GRADIENT_ACCUMULATION_STEPS = "gradient_accumulation_steps"
STEPS_PER_PRINT = "steps_per_print"
STEPS_PER_PRINT_DEFAULT = 10
DUMP_STATE = "dump_state"
DISABLE_ALLGATHER = "disable_allgather"
FP32_ALLREDUCE = "fp32_allreduce"
PRESCALE_GRADIENTS = "prescale_gradients"
GRADIENT_PREDIVIDE_FACTOR = "gradient_predivide_factor"
SPARSE_GRADIENTS = "sparse_gradients"
GRADIENT_CLIPPING = "gradient_clipping"
ZERO_ALLOW_UNTESTED_OPTIMIZER = "zero_allow_untested_optimizer"
WALL_CLOCK_BREAKDOWN = "wall_clock_breakdown"
MEMORY_BREAKDOWN = "memory_breakdown"
VOCABULARY_SIZE = "vocabulary_size"
TENSOR_CORE_ALIGN_SIZE = 8

OPTIMIZER = "optimizer"
SCHEDULER = "scheduler"
TYPE = "type"
PARAMS = "params"
LEGACY_FUSION = "legacy_fusion"
MAX_GRAD_NORM = "max_grad_norm"

ADAM_OPTIMIZER = "adam"
ADAMW_OPTIMIZER = "adamw"
LAMB_OPTIMIZER = "lamb"
ONEBIT_ADAM_OPTIMIZER = "onebitadam"
ONEBIT_LAMB_OPTIMIZER = "onebitlamb"
DEEPSPEED_OPTIMIZERS = [ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER, ONEBIT_LAMB_OPTIMIZER]
TORCH_ADAM_PARAM = "torch_adam"
ADAM_W_MODE = "adam_w_mode"
ADAM_W_MODE_DEFAULT = True

# fp16 block (DeeperSpeed: "type" selects fp16 | bfloat16 | fp32)
FP16 = "fp16"
FP16_ENABLED = "enabled"
FP16_TYPE = "type"
FP16_TYPE_DEFAULT = "fp16"
FP16_LOSS_SCALE = "loss_scale"
FP16_INITIAL_SCALE_POWER = "initial_scale_power"
FP16_INITIAL_SCALE_POWER_DEFAULT = 32
FP16_LOSS_SCALE_WINDOW = "loss_scale_window"
FP16_LOSS_SCALE_WINDOW_DEFAULT = 1000
FP16_HYSTERESIS = "hysteresis"
FP16_HYSTERESIS_DEFAULT = 2
FP16_MIN_LOSS_SCALE = "min_loss_scale"
FP16_MIN_LOSS_SCALE_DEFAULT = 1
INITIAL_LOSS_SCALE = "init_scale"
SCALE_WINDOW = "scale_window"
DELAYED_SHIFT = "delayed_shift"
MIN_LOSS_SCALE = "min_scale"

AMP = "amp"
AMP_ENABLED = "enabled"

TENSORBOARD = "tensorboard"
TENSORBOARD_ENABLED = "enabled"
TENSORBOARD_OUTPUT_PATH = "output_path"
TENSORBOARD_JOB_NAME = "job_name"
TENSORBOARD_JOB_NAME_DEFAULT = "DeepSpeedJobName"

PROGRESSIVE_LAYER_DROP = "progressive_layer_drop"
PLD_ENABLED = "enabled"
PLD_THETA = "theta"
PLD_THETA_DEFAULT = 1.0
PLD_GAMMA = "gamma"
PLD_GAMMA_DEFAULT = 0.001

CHECKPOINT = "checkpoint"
CHECKPOINT_TAG_VALIDATION = "tag_validation"
CHECKPOINT_TAG_VALIDATION_DEFAULT = "WARN"
CHECKPOINT_TAG_VALIDATION_MODES = ["WARN", "IGNORE", "FAIL"]

SPARSE_ATTENTION = "sparse_attention"
SPARSE_DENSE_MODE = "dense"
SPARSE_FIXED_MODE = "fixed"
SPARSE_VARIABLE_MODE = "variable"
SPARSE_BIGBIRD_MODE = "bigbird"
SPARSE_BSLONGFORMER_MODE = "bslongformer"
SPARSE_MODE = "mode"
SPARSE_MODE_DEFAULT = SPARSE_FIXED_MODE
SPARSE_BLOCK = "block"
SPARSE_BLOCK_DEFAULT = 16

PIPELINE = "pipeline"

# ZeRO
ZERO_OPTIMIZATION = "zero_optimization"
ZERO_STAGE = "stage"
ZERO_OPTIMIZATION_DISABLED = 0
ZERO_OPTIMIZATION_OPTIMIZER_STATES = 1
ZERO_OPTIMIZATION_GRADIENTS = 2
ZERO_OPTIMIZATION_WEIGHTS = 3
MAX_STAGE_ZERO_OPTIMIZATION = 3
ZERO_DEFAULTS = {
    "stage": 0,
    "contiguous_gradients": False,
    "reduce_scatter": True,
    "reduce_bucket_size": int(5e8),
    "allgather_partitions": True,
    "allgather_bucket_size": int(5e8),
    "overlap_comm": False,
    "load_from_fp32_weights": True,
    "elastic_checkpoint": True,
    "cpu_offload": False,
    "cpu_offload_params": False,
    "cpu_offload_use_pin_memory": False,
    "sub_group_size": int(1e12),
    "stage3_max_live_parameters": int(1e9),
    "stage3_max_reuse_distance": int(1e9),
    "stage3_prefetch_bucket_size": int(5e7),
    "stage3_param_persistence_threshold": int(1e5),
    "stage3_gather_fp16_weights_on_model_save": False,
    # B200-native extension: fuse reduce-scatter + cast + Adam + all-gather over NVSwitch peer memory
    "fused_nvlink": "auto",
    # Stage 2 gradient memory: "sharded" keeps only this rank's reduced slices (the reference's behaviour: gradient
    # bytes / dp), "full" keeps a whole gradient arena that autograd accumulates into in place (fastest; fine while
    # the arena is small next to 180 GB of HBM), "auto" = sharded once the arena would exceed 8 GiB per rank
    "gradient_memory": "auto",
    # B200-native extension (stage 3): Linear forwards run the weight all-gather inside the GEMM kernel; shards of
    # eligible weights live on a peer-mapped heap (zero/partition_parameters.py, parallel/ag_gemm.py)
    "stage3_fused_allgather_gemm": False,
}
OFFLOAD_PARAM = "offload_param"
OFFLOAD_OPTIMIZER = "offload_optimizer"
OFFLOAD_PARAM_DEFAULTS = {"device": "cpu", "nvme_path": None, "buffer_count": 5, "buffer_size": int(1e8),
                          "max_in_cpu": int(1e9), "pin_memory": False}
OFFLOAD_OPTIMIZER_DEFAULTS = {"device": "cpu", "nvme_path": None, "buffer_count": 4, "pin_memory": False,
                              "pipeline_read": False, "pipeline_write": False, "fast_init": False}

ACT_CHKPT = "activation_checkpointing"
ACT_CHKPT_DEFAULTS = {
    "partition_activations": False,
    "contiguous_memory_optimization": False,
    "cpu_checkpointing": False,
    "number_checkpoints": None,
    "synchronize_checkpoint_boundary": False,
    "profile": False,
}

FLOPS_PROFILER = "flops_profiler"
FLOPS_PROFILER_DEFAULTS = {"enabled": False, "profile_step": 1, "module_depth": -1, "top_modules": 3, "detailed": True}

AIO = "aio"
AIO_DEFAULTS = {"block_size": 1048576, "queue_depth": 8, "thread_count": 1, "single_submit": False,
                "overlap_events": True}

ELASTICITY = "elasticity"

ROUTE_TRAIN = "train"
ROUTE_EVAL = "eval"
ROUTE_PREDICT = "predict"
ROUTE_ENCODE = "encode"

TORCH_DISTRIBUTED_DEFAULT_PORT = 29500
DEFAULT_PROCESS_GROUP_TIMEOUT_MINUTES = 30


class ValidationMode:
    """Checkpoint tag validation modes (reference runtime/constants.py:323-326)."""
    WARN = "WARN"
    IGNORE = "IGNORE"
    FAIL = "FAIL"
