"""DeepSpeedConfig: JSON/dict -> typed attributes, batch-triple solver, sanity checks.

Schema and defaults follow reference deepspeed/runtime/config.py:536-812, config_utils.py:13-76,
zero/config.py:14-177; the implementation is a small declarative reader instead of one getter per key.
"""
import copy
import json
import os

import torch

from . import constants as C
from .constants import (ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER,  # noqa: F401
                        ONEBIT_LAMB_OPTIMIZER, DEEPSPEED_OPTIMIZERS, TORCH_ADAM_PARAM, ADAM_W_MODE,
                        ADAM_W_MODE_DEFAULT)      # the reference defines the optimizer names in this module
from ..utils.logging import logger


class DeepSpeedConfigError(Exception):
    pass


def dict_raise_error_on_duplicate_keys(ordered_pairs):
    """json.load object_pairs_hook rejecting duplicate keys (reference config_utils.py:69-76)."""
    d = dict(ordered_pairs)
    if len(d) != len(ordered_pairs):
        seen, dup = set(), []
        for k, _ in ordered_pairs:
            if k in seen:
                dup.append(k)
            seen.add(k)
        raise ValueError("Duplicate keys in DeepSpeed config: {}".format(dup))
    return d


def get_scalar_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


class ScientificNotationEncoder(json.JSONEncoder):
    """Pretty printer writing large numbers as 1e+09 (reference config_utils.py:13-48)."""

    def iterencode(self, o, _one_shot=False, level=0):
        indent = self.indent if self.indent is not None else 4
        pre_close, pre = " " * level * indent, " " * (level + 1) * indent
        if isinstance(o, bool):
            return "true" if o else "false"
        if isinstance(o, (float, int)):
            return f"{o:e}" if o > 1e3 else f"{o}"
        if isinstance(o, dict):
            x = [f'\n{pre}"{k}": {self.iterencode(v, level=level + 1)}' for k, v in o.items()]
            return "{" + ", ".join(x) + f"\n{pre_close}" + "}"
        if isinstance(o, (list, tuple)):
            return f"[{', '.join(map(lambda v: self.iterencode(v, level=level + 1), o))}]"
        return "\n, ".join(super().iterencode(o, _one_shot))


class DeepSpeedConfigObject:
    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4, cls=ScientificNotationEncoder)


class DeepSpeedZeroOffloadConfig(DeepSpeedConfigObject):
    def __init__(self, d, defaults):
        for k, v in defaults.items():
            setattr(self, k, d.get(k, v) if d else v)


class DeepSpeedZeroConfig(DeepSpeedConfigObject):
    """`zero_optimization` block; the deprecated boolean form maps to stage 1 (reference zero/config.py:51-63)."""

    def __init__(self, param_dict):
        z = param_dict.get(C.ZERO_OPTIMIZATION, None)
        if isinstance(z, bool):
            logger.warning("DeepSpeedConfig: zero_optimization as a bool is deprecated; use {'stage': N}")
            z = {C.ZERO_STAGE: 1 if z else 0}
        z = z or {}
        for k, v in C.ZERO_DEFAULTS.items():
            setattr(self, k, z.get(k, v))
        for k in ("reduce_bucket_size", "allgather_bucket_size", "sub_group_size", "stage3_max_live_parameters",
                  "stage3_max_reuse_distance", "stage3_prefetch_bucket_size", "stage3_param_persistence_threshold"):
            setattr(self, k, int(getattr(self, k)))
        self.offload_param = None
        self.offload_optimizer = None
        if C.OFFLOAD_PARAM in z:
            self.offload_param = dict(C.OFFLOAD_PARAM_DEFAULTS, **(z[C.OFFLOAD_PARAM] or {}))
        elif self.cpu_offload_params:
            self.offload_param = dict(C.OFFLOAD_PARAM_DEFAULTS, pin_memory=self.cpu_offload_use_pin_memory)
        if C.OFFLOAD_OPTIMIZER in z:
            self.offload_optimizer = dict(C.OFFLOAD_OPTIMIZER_DEFAULTS, **(z[C.OFFLOAD_OPTIMIZER] or {}))
        elif self.cpu_offload:
            self.offload_optimizer = dict(C.OFFLOAD_OPTIMIZER_DEFAULTS, pin_memory=self.cpu_offload_use_pin_memory)
        if self.offload_optimizer is not None and self.offload_optimizer["device"] == "cpu":
            self.cpu_offload = True
        # reference property names
        self.param_persistence_threshold = self.stage3_param_persistence_threshold
        self.max_live_parameters = self.stage3_max_live_parameters
        self.max_reuse_distance = self.stage3_max_reuse_distance
        self.prefetch_bucket_size = self.stage3_prefetch_bucket_size
        self.gather_fp16_weights_on_model_save = self.stage3_gather_fp16_weights_on_model_save


class DeepSpeedActivationCheckpointingConfig(DeepSpeedConfigObject):
    def __init__(self, param_dict):
        a = param_dict.get(C.ACT_CHKPT, {}) or {}
        for k, v in C.ACT_CHKPT_DEFAULTS.items():
            setattr(self, k, a.get(k, v))


class DeepSpeedFlopsProfilerConfig(DeepSpeedConfigObject):
    def __init__(self, param_dict):
        a = param_dict.get(C.FLOPS_PROFILER, {}) or {}
        for k, v in C.FLOPS_PROFILER_DEFAULTS.items():
            setattr(self, k, a.get(k, v))


class DeepSpeedConfigWriter:
    """Build / load / write a config JSON programmatically (reference runtime/config.py:519-533)."""

    def __init__(self, data=None):
        self.data = data if data is not None else {}

    def add_config(self, key, value):
        self.data[key] = value

    def load_config(self, filename):
        with open(filename, "r") as f:
            self.data = json.load(f, object_pairs_hook=dict_raise_error_on_duplicate_keys)

    def write_config(self, filename):
        with open(filename, "w") as outfile:
            json.dump(self.data, outfile)


def get_aio_config(param_dict):
    a = param_dict.get(C.AIO, {}) or {}
    return {k: a.get(k, v) for k, v in C.AIO_DEFAULTS.items()}


def get_pipeline_config(param_dict):
    default = {"stages": "auto", "partition": "best", "seed_layers": False, "activation_checkpoint_interval": 0}
    cfg = dict(default)
    for k, v in (param_dict.get(C.PIPELINE, {}) or {}).items():
        cfg[k] = v
    return cfg


_SPARSE_KEYS = {
    C.SPARSE_DENSE_MODE: {},
    C.SPARSE_FIXED_MODE: {"num_local_blocks": 4, "num_global_blocks": 1, "attention": "bidirectional",
                          "horizontal_global_attention": False, "num_different_global_patterns": 1},
    C.SPARSE_VARIABLE_MODE: {"num_random_blocks": 0, "local_window_blocks": [4], "global_block_indices": [0],
                             "global_block_end_indices": None, "attention": "bidirectional",
                             "horizontal_global_attention": False},
    C.SPARSE_BIGBIRD_MODE: {"num_random_blocks": 1, "num_sliding_window_blocks": 3, "num_global_blocks": 1},
    C.SPARSE_BSLONGFORMER_MODE: {"num_sliding_window_blocks": 3, "global_block_indices": [0],
                                 "global_block_end_indices": None},
}


def get_sparse_attention(param_dict):
    """`sparse_attention` block -> dict of mode-specific keys (reference config.py:213-381)."""
    if C.SPARSE_ATTENTION not in param_dict:
        return None
    s = param_dict[C.SPARSE_ATTENTION]
    mode = s.get(C.SPARSE_MODE, C.SPARSE_MODE_DEFAULT)
    if mode not in _SPARSE_KEYS:
        raise NotImplementedError(f"Given sparsity mode, {mode}, has not been implemented yet!")
    out = {C.SPARSE_MODE: mode, C.SPARSE_BLOCK: s.get(C.SPARSE_BLOCK, C.SPARSE_BLOCK_DEFAULT),
           "different_layout_per_head": s.get("different_layout_per_head", False)}
    for k, v in _SPARSE_KEYS[mode].items():
        out[k] = s.get(k, v)
    return out


class DeepSpeedConfig:
    def __init__(self, json_file, mpu=None, param_dict=None):
        if param_dict is None:
            if isinstance(json_file, dict):
                self._param_dict = copy.deepcopy(json_file)
            else:
                with open(json_file, "r") as f:
                    self._param_dict = json.load(f, object_pairs_hook=dict_raise_error_on_duplicate_keys)
        else:
            self._param_dict = param_dict
        try:
            self.global_rank = torch.distributed.get_rank()
            self.world_size = (torch.distributed.get_world_size() if mpu is None else mpu.get_data_parallel_world_size())
        except Exception:
            self.global_rank = 0
            self.world_size = 1

        # elasticity may override the batch parameters (reference config.py:558-609)
        from ..elasticity import elasticity_enabled, compute_elastic_config, ensure_immutable_elastic_config
        from ..elasticity.constants import (ELASTICITY, IGNORE_NON_ELASTIC_BATCH_INFO,
                                            IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT)
        self.elasticity_enabled = elasticity_enabled(self._param_dict)
        if self.elasticity_enabled:
            logger.info("DeepSpeed elasticity support enabled")
            from ..version import __version__
            final_batch, valid_gpus, micro = compute_elastic_config(ds_config=self._param_dict,
                                                                    target_deepspeed_version=__version__,
                                                                    world_size=self.world_size)
            elastic_dict = self._param_dict[ELASTICITY]
            ensure_immutable_elastic_config(runtime_elastic_config_dict=elastic_dict)
            ignore = elastic_dict.get(IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT)
            if not ignore:
                for k in (C.TRAIN_BATCH_SIZE, C.TRAIN_MICRO_BATCH_SIZE_PER_GPU, C.GRADIENT_ACCUMULATION_STEPS):
                    if k in self._param_dict:
                        from ..elasticity.config import ElasticityConfigError
                        raise ElasticityConfigError(
                            f"One or more batch related parameters were found in your ds_config ({k}). These "
                            "parameters *will not be used* since elastic training is enabled; set "
                            f"{IGNORE_NON_ELASTIC_BATCH_INFO}:true to silence this error.")
            gas = final_batch // (micro * self.world_size)
            self._param_dict[C.TRAIN_BATCH_SIZE] = final_batch
            self._param_dict[C.TRAIN_MICRO_BATCH_SIZE_PER_GPU] = micro
            self._param_dict[C.GRADIENT_ACCUMULATION_STEPS] = gas

        self._initialize_params(self._param_dict)
        self._configure_train_batch_size()
        self._do_sanity_check()

    # ------------------------------------------------------------------ parsing
    def _initialize_params(self, p):
        g = get_scalar_param
        self.train_batch_size = g(p, C.TRAIN_BATCH_SIZE, None)
        self.train_micro_batch_size_per_gpu = g(p, C.TRAIN_MICRO_BATCH_SIZE_PER_GPU, None)
        self.gradient_accumulation_steps = g(p, C.GRADIENT_ACCUMULATION_STEPS, None)
        self.steps_per_print = g(p, C.STEPS_PER_PRINT, C.STEPS_PER_PRINT_DEFAULT)
        self.dump_state = g(p, C.DUMP_STATE, False)
        self.disable_allgather = g(p, C.DISABLE_ALLGATHER, False)
        self.prescale_gradients = g(p, C.PRESCALE_GRADIENTS, False)
        self.gradient_predivide_factor = g(p, C.GRADIENT_PREDIVIDE_FACTOR, 1.0)
        self.sparse_gradients_enabled = g(p, C.SPARSE_GRADIENTS, False)
        self.zero_config = DeepSpeedZeroConfig(p)
        self.zero_optimization_stage = self.zero_config.stage
        self.zero_enabled = self.zero_optimization_stage > 0
        self.activation_checkpointing_config = DeepSpeedActivationCheckpointingConfig(p)
        self.flops_profiler_config = DeepSpeedFlopsProfilerConfig(p)
        self.aio_config = get_aio_config(p)
        self.gradient_clipping = g(p, C.GRADIENT_CLIPPING, 0.0)

        fp16 = p.get(C.FP16, {}) or {}
        self.fp16_enabled = fp16.get(C.FP16_ENABLED, False)
        self.precision = self._precision(fp16) if self.fp16_enabled else torch.float32
        self.fp32_allreduce = g(p, C.FP32_ALLREDUCE, self.precision == torch.bfloat16)
        amp = p.get(C.AMP, {}) or {}
        self.amp_enabled = amp.get(C.AMP_ENABLED, False)
        self.amp_params = {k: v for k, v in amp.items() if k != C.AMP_ENABLED} if amp else False
        if self.precision == torch.bfloat16:
            # bf16 has fp32 range: loss scale is pinned to 1.0 (reference config.py:104-113)
            self.loss_scale = 1.0
        else:
            self.loss_scale = fp16.get(C.FP16_LOSS_SCALE, 0) if self.fp16_enabled else 0
        self.initial_dynamic_scale = 2 ** (fp16.get(C.FP16_INITIAL_SCALE_POWER, C.FP16_INITIAL_SCALE_POWER_DEFAULT)
                                           if self.fp16_enabled else C.FP16_INITIAL_SCALE_POWER_DEFAULT)
        self.dynamic_loss_scale_args = None
        if self.fp16_enabled and any(k in fp16 for k in (C.FP16_INITIAL_SCALE_POWER, C.FP16_LOSS_SCALE_WINDOW,
                                                         C.FP16_MIN_LOSS_SCALE, C.FP16_HYSTERESIS)):
            self.dynamic_loss_scale_args = {
                C.INITIAL_LOSS_SCALE: 2 ** fp16.get(C.FP16_INITIAL_SCALE_POWER, C.FP16_INITIAL_SCALE_POWER_DEFAULT),
                C.SCALE_WINDOW: fp16.get(C.FP16_LOSS_SCALE_WINDOW, C.FP16_LOSS_SCALE_WINDOW_DEFAULT),
                C.DELAYED_SHIFT: fp16.get(C.FP16_HYSTERESIS, C.FP16_HYSTERESIS_DEFAULT),
                C.MIN_LOSS_SCALE: fp16.get(C.FP16_MIN_LOSS_SCALE, C.FP16_MIN_LOSS_SCALE_DEFAULT),
            }

        opt = p.get(C.OPTIMIZER, None)
        self.optimizer_name = None
        self.optimizer_params = None
        self.optimizer_legacy_fusion = False
        if opt is not None:
            self.optimizer_name = opt.get(C.TYPE, None)
            if self.optimizer_name is not None and self.optimizer_name.lower() in C.DEEPSPEED_OPTIMIZERS:
                self.optimizer_name = self.optimizer_name.lower()
            self.optimizer_params = opt.get(C.PARAMS, None)
            self.optimizer_legacy_fusion = opt.get(C.LEGACY_FUSION, False)
        self.zero_allow_untested_optimizer = g(p, C.ZERO_ALLOW_UNTESTED_OPTIMIZER, False)
        sch = p.get(C.SCHEDULER, None)
        self.scheduler_name = sch.get(C.TYPE, None) if sch else None
        self.scheduler_params = sch.get(C.PARAMS, None) if sch else None

        self.wall_clock_breakdown = g(p, C.WALL_CLOCK_BREAKDOWN, False) | self.flops_profiler_config.enabled
        self.memory_breakdown = g(p, C.MEMORY_BREAKDOWN, False)
        tb = p.get(C.TENSORBOARD, {}) or {}
        self.tensorboard_enabled = tb.get(C.TENSORBOARD_ENABLED, False)
        self.tensorboard_output_path = tb.get(C.TENSORBOARD_OUTPUT_PATH, "")
        self.tensorboard_job_name = tb.get(C.TENSORBOARD_JOB_NAME, C.TENSORBOARD_JOB_NAME_DEFAULT)
        self.sparse_attention = get_sparse_attention(p)
        self.pipeline = get_pipeline_config(p)
        pld = p.get(C.PROGRESSIVE_LAYER_DROP, {}) or {}
        self.pld_enabled = pld.get(C.PLD_ENABLED, False)
        self.pld_params = {k: v for k, v in pld.items() if k != C.PLD_ENABLED} if self.pld_enabled else False
        ck = p.get(C.CHECKPOINT, {}) or {}
        mode = str(ck.get(C.CHECKPOINT_TAG_VALIDATION, C.CHECKPOINT_TAG_VALIDATION_DEFAULT)).upper()
        if mode not in C.CHECKPOINT_TAG_VALIDATION_MODES:
            raise DeepSpeedConfigError(f"Checkpoint config contains invalid tag_validation value of {mode}, "
                                       f"expecting one of {C.CHECKPOINT_TAG_VALIDATION_MODES}")
        self.checkpoint_tag_validation_enabled = mode != "IGNORE"
        self.checkpoint_tag_validation_fail = mode == "FAIL"
        self.vocabulary_size = g(p, C.VOCABULARY_SIZE, None)

    @staticmethod
    def _precision(fp16):
        t = str(fp16.get(C.FP16_TYPE, C.FP16_TYPE_DEFAULT)).lower()
        if t in ("fp16", "float16", "half"):
            return torch.float16
        if t in ("bfloat16", "bf16"):
            return torch.bfloat16
        if t in ("fp32", "float32", "float"):
            return torch.float32
        raise DeepSpeedConfigError(f"fp16.type must be fp16 | bfloat16 | fp32, got {t}")

    # ------------------------------------------------------------------ batch triple
    def _batch_assertion(self):
        tb, mb, gas = self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps
        assert tb > 0, f"Train batch size: {tb} has to be greater than 0"
        assert mb > 0, f"Micro batch size per gpu: {mb} has to be greater than 0"
        assert gas > 0, f"Gradient accumulation steps: {gas} has to be greater than 0"
        assert tb == mb * gas * self.world_size, (
            f"Check batch related parameters. train_batch_size is not equal to micro_batch_per_gpu * "
            f"gradient_acc_step * world_size {tb} != {mb} * {gas} * {self.world_size}")

    def _set_batch_related_parameters(self):
        tb, mb, gas, ws = (self.train_batch_size, self.train_micro_batch_size_per_gpu,
                           self.gradient_accumulation_steps, self.world_size)
        if tb is not None and mb is not None and gas is not None:
            return
        if tb is not None and mb is not None:
            self.gradient_accumulation_steps = tb // mb // ws
        elif tb is not None and gas is not None:
            self.train_micro_batch_size_per_gpu = tb // ws // gas
        elif mb is not None and gas is not None:
            self.train_batch_size = mb * gas * ws
        elif tb is not None:
            self.gradient_accumulation_steps = 1
            self.train_micro_batch_size_per_gpu = tb // ws
        elif mb is not None:
            self.train_batch_size = mb * ws
            self.gradient_accumulation_steps = 1
        else:
            assert False, "Either train_batch_size or micro_batch_per_gpu needs to be provided"

    def _configure_train_batch_size(self):
        self._set_batch_related_parameters()
        self._batch_assertion()

    # ------------------------------------------------------------------ checks
    def _do_sanity_check(self):
        assert self.train_micro_batch_size_per_gpu, f"DeepSpeedConfig: {C.TRAIN_MICRO_BATCH_SIZE_PER_GPU} is not defined"
        assert self.gradient_accumulation_steps, f"DeepSpeedConfig: {C.GRADIENT_ACCUMULATION_STEPS} is not defined"
        if self.zero_enabled:
            assert self.fp16_enabled, "DeepSpeedConfig: ZeRO is only supported if fp16 is enabled"
            assert self.zero_optimization_stage <= C.MAX_STAGE_ZERO_OPTIMIZATION, (
                f"DeepSpeedConfig: Maximum supported ZeRO stage is {C.MAX_STAGE_ZERO_OPTIMIZATION}")
        if self.vocabulary_size and self.vocabulary_size % C.TENSOR_CORE_ALIGN_SIZE != 0:
            logger.warning(f"DeepSpeedConfig: vocabulary size {self.vocabulary_size} is not aligned to "
                           f"{C.TENSOR_CORE_ALIGN_SIZE}, may import tensor core utilization.")
        if (self.optimizer_params is not None and C.MAX_GRAD_NORM in self.optimizer_params
                and self.optimizer_params[C.MAX_GRAD_NORM] > 0):
            if self.fp16_enabled:
                logger.warning(f"DeepSpeedConfig: In FP16 mode, DeepSpeed will pass {C.MAX_GRAD_NORM} to FP16 wrapper")
            else:
                logger.warning(f"DeepSpeedConfig: In FP32 mode, {C.MAX_GRAD_NORM} is not permitted; setting to zero")
                self.optimizer_params[C.MAX_GRAD_NORM] = 0.0

    def print(self, name):
        logger.info(f"{name}:")
        for arg in sorted(vars(self)):
            if arg != "_param_dict":
                logger.info("  {} {} {}".format(arg, "." * (29 - len(arg)), getattr(self, arg)))
        logger.info("  json = {}".format(json.dumps(self._param_dict, sort_keys=True, indent=4,
                                                    cls=ScientificNotationEncoder, separators=(",", ":"))))


# ---- module-level getters of the reference (runtime/config.py:45-516) ------------------------------------------------
# The reference parses every key with a free function `get_<name>(param_dict)`; here one parser fills DeepSpeedConfig
# and the free functions read the attribute it produced, so both spellings always agree.
def _parsed(param_dict):
    cfg = DeepSpeedConfig.__new__(DeepSpeedConfig)
    cfg._param_dict = param_dict
    cfg._initialize_params(param_dict)
    return cfg


def _getter(attr, post=None):
    def get(param_dict):
        v = getattr(_parsed(param_dict), attr)
        return post(v) if post else v
    get.__name__ = "get_" + attr
    return get


get_pld_enabled = _getter("pld_enabled")
get_pld_params = _getter("pld_params")
get_amp_enabled = _getter("amp_enabled")
get_amp_params = _getter("amp_params")
get_fp16_enabled = _getter("fp16_enabled")
get_fp16_type = _getter("precision", lambda p: {torch.float16: "fp16", torch.bfloat16: "bfloat16"}.get(p, "fp32"))
get_loss_scale = _getter("loss_scale")
get_initial_dynamic_scale = _getter("initial_dynamic_scale")
get_dynamic_loss_scale_args = _getter("dynamic_loss_scale_args")
get_gradient_accumulation_steps = _getter("gradient_accumulation_steps")
get_sparse_gradients_enabled = _getter("sparse_gradients_enabled")
get_zero_optimization = _getter("zero_enabled")
get_zero_reduce_scatter = _getter("zero_config", lambda z: z.reduce_scatter)
get_allreduce_always_fp32 = _getter("fp32_allreduce")
get_prescale_gradients = _getter("prescale_gradients")
get_gradient_predivide_factor = _getter("gradient_predivide_factor")
get_steps_per_print = _getter("steps_per_print")
get_disable_allgather = _getter("disable_allgather")
get_dump_state = _getter("dump_state")
get_gradient_clipping = _getter("gradient_clipping")
get_optimizer_name = _getter("optimizer_name")
get_optimizer_params = _getter("optimizer_params")
get_optimizer_legacy_fusion = _getter("optimizer_legacy_fusion")
get_zero_allow_untested_optimizer = _getter("zero_allow_untested_optimizer")
get_scheduler_name = _getter("scheduler_name")
get_scheduler_params = _getter("scheduler_params")
get_train_batch_size = _getter("train_batch_size")
get_train_micro_batch_size_per_gpu = _getter("train_micro_batch_size_per_gpu")
get_wall_clock_breakdown = _getter("wall_clock_breakdown")
get_memory_breakdown = _getter("memory_breakdown")
get_tensorboard_enabled = _getter("tensorboard_enabled")
get_tensorboard_output_path = _getter("tensorboard_output_path")
get_tensorboard_job_name = _getter("tensorboard_job_name")


def get_optimizer_gradient_clipping(param_dict):
    params = get_optimizer_params(param_dict)
    return params.get("max_grad_norm", None) if params else None


def get_checkpoint_params(param_dict):
    return param_dict.get(C.CHECKPOINT, {})


def get_checkpoint_tag_validation_mode(checkpoint_params):
    mode = str(checkpoint_params.get(C.CHECKPOINT_TAG_VALIDATION, C.CHECKPOINT_TAG_VALIDATION_DEFAULT)).upper()
    if mode not in C.CHECKPOINT_TAG_VALIDATION_MODES:
        raise DeepSpeedConfigError(f"Checkpoint config contains invalid tag_validation value of {mode}, "
                                   f"expecting one of {C.CHECKPOINT_TAG_VALIDATION_MODES}")
    return mode


def get_sparse_attention_mode(param_dict):
    return param_dict.get(C.SPARSE_MODE, C.SPARSE_MODE_DEFAULT)


def get_sparse_attention_type(param_dict):
    return param_dict.get("attention", "bidirectional")


def _sparse_mode_config(mode):
    def get(sparsity):
        out = {C.SPARSE_MODE: mode, C.SPARSE_BLOCK: sparsity.get(C.SPARSE_BLOCK, C.SPARSE_BLOCK_DEFAULT)}
        if mode != C.SPARSE_DENSE_MODE:
            out["different_layout_per_head"] = sparsity.get("different_layout_per_head", False)
        for k, v in _SPARSE_KEYS[mode].items():
            out[k] = sparsity.get(k, v)
        return out
    return get


get_sparse_dense_config = _sparse_mode_config(C.SPARSE_DENSE_MODE)
get_sparse_fixed_config = _sparse_mode_config(C.SPARSE_FIXED_MODE)
get_sparse_variable_config = _sparse_mode_config(C.SPARSE_VARIABLE_MODE)
get_sparse_bigbird_config = _sparse_mode_config(C.SPARSE_BIGBIRD_MODE)
get_sparse_bslongformer_config = _sparse_mode_config(C.SPARSE_BSLONGFORMER_MODE)
