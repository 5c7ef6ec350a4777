"""Builds the optimizer stack of an engine: basic optimizer -> (fp16 / bf16 master-weight wrapper | ZeRO stage 1-3).

Selection rules per reference deepspeed/runtime/engine.py:560-760 (optimizer names and their parameters in the JSON
config, which wrapper goes with which optimizer, the ZeRO constructor options). Written as tables: one entry per
optimizer name, one per ZeRO stage, each a small function of the engine's configuration accessors.
"""
import torch

from ..utils.logging import logger, log_dist
from . import constants as C


# ----------------------------------------------------------------------------------------------- basic optimizers
def _adam(engine, params, opts, name):
    use_torch = opts.pop(C.TORCH_ADAM_PARAM, False)
    # "adamw" always means decoupled weight decay; "adam" follows adam_w_mode (default True, as in the reference)
    decoupled = name == C.ADAMW_OPTIMIZER or opts.pop(C.ADAM_W_MODE, C.ADAM_W_MODE_DEFAULT)
    opts.pop(C.ADAM_W_MODE, None)
    if use_torch:
        return (torch.optim.AdamW if decoupled else torch.optim.Adam)(params, **opts)
    if engine.zero_cpu_offload():
        from ..ops.adam import DeepSpeedCPUAdam
        return DeepSpeedCPUAdam(params, **opts, adamw_mode=decoupled)
    from ..ops.adam import FusedAdam
    return FusedAdam(params, **opts, adam_w_mode=decoupled)


def _lamb(engine, params, opts, name):
    from ..ops.lamb import FusedLamb
    return FusedLamb(params, **opts)


def _onebit(engine, params, opts, name):
    if name == C.ONEBIT_ADAM_OPTIMIZER:
        from .fp16.onebit.adam import OnebitAdam as cls
        label = "1-bit Adam"
    else:
        from .fp16.onebit.lamb import OnebitLamb as cls
        label = "1-bit Lamb"
    if not engine.fp16_enabled():
        logger.warning(f"Currently the convergence of {label} is only verified under FP16")
    return cls(params, engine, **opts)


_BUILDERS = {C.ADAM_OPTIMIZER: _adam, C.ADAMW_OPTIMIZER: _adam, C.LAMB_OPTIMIZER: _lamb,
             C.ONEBIT_ADAM_OPTIMIZER: _onebit, C.ONEBIT_LAMB_OPTIMIZER: _onebit}


def build_basic_optimizer(engine, model_parameters):
    """The optimizer named in the config (`optimizer.type` / `optimizer.params`)."""
    opts = dict(engine.optimizer_params() or {})
    if "max_grad_norm" in opts:
        raise ValueError("'max_grad_norm' is not supported as an optimizer parameter, please switch to using the "
                         "deepspeed parameter 'gradient_clipping' see: "
                         "https://www.deepspeed.ai/docs/config-json/#gradient-clipping for more details")
    name = engine.optimizer_name()
    build = _BUILDERS.get(name)
    if build is not None:
        return build(engine, model_parameters, opts, name)
    return getattr(torch.optim, name)(model_parameters, **opts)      # any torch.optim class by name


# ----------------------------------------------------------------------------------------------- precision wrapper
def wrap_for_reduced_precision(engine, optimizer):
    """fp32 master weights + loss scaling for fp16 / bf16 training without ZeRO. Adam-family optimizers get the flat
    fused wrapper; everything else (LAMB's per-tensor trust ratio, client optimizers) keeps per-parameter masters."""
    from ..ops.adam import FusedAdam
    from .fp16.fused_optimizer import FP16_Optimizer
    from .fp16.unfused_optimizer import FP16_UnfusedOptimizer
    bf16 = engine.precision() == torch.bfloat16           # bf16 has fp32's range: no loss scaling
    dynamic = engine.dynamic_loss_scale() and not bf16
    static_scale = 1.0 if bf16 else engine.loss_scale()
    shared = dict(mpu=engine.mpu, clip_grad=engine.gradient_clipping())
    if isinstance(optimizer, FusedAdam) or engine.optimizer_name() == C.ONEBIT_ADAM_OPTIMIZER:
        flat = dict(shared, fused_adam_legacy=engine.optimizer_legacy_fusion(),
                    dp_process_group=engine.data_parallel_group, fp32_allreduce=engine.allreduce_always_fp32())
        if dynamic:
            log_dist("Creating fp16 optimizer with dynamic loss scale", ranks=[0])
            return FP16_Optimizer(optimizer, dynamic_loss_scale=True, timers=None,
                                  initial_dynamic_scale=engine.initial_dynamic_scale(),
                                  dynamic_loss_args=engine.dynamic_loss_scale_args(), **flat)
        log_dist("Creating fp16 optimizer with static loss scale: {}".format(engine.loss_scale()), ranks=[0])
        return FP16_Optimizer(optimizer, static_loss_scale=static_scale, **flat)
    log_dist("Creating fp16 unfused optimizer with dynamic loss scale", ranks=[0])
    return FP16_UnfusedOptimizer(optimizer, static_loss_scale=static_scale, dynamic_loss_scale=dynamic,
                                 dynamic_loss_args=engine.dynamic_loss_scale_args(),
                                 fused_lamb_legacy=engine.optimizer_name() == C.LAMB_OPTIMIZER, **shared)


# ----------------------------------------------------------------------------------------------- ZeRO
def _scaling_and_groups(e):
    return dict(static_loss_scale=e.loss_scale(), dynamic_loss_scale=e.dynamic_loss_scale(),
                dynamic_loss_args=e.dynamic_loss_scale_args(), clip_grad=e.gradient_clipping(),
                dp_process_group=e.data_parallel_group, mpu=e.mpu, precision=e.precision(),
                postscale_gradients=e.postscale_gradients(), gradient_predivide_factor=e.gradient_predivide_factor())


def _fp32_reduce(e):
    return e.allreduce_always_fp32() and e.precision() != torch.bfloat16


def _zero1(e, optimizer):
    assert e.zero_reduce_scatter(), "Stage 1 only supports reduce scatter mode"
    from .zero.stage1 import FP16_DeepSpeedZeroOptimizer_Stage1
    return FP16_DeepSpeedZeroOptimizer_Stage1(
        optimizer, all_gather_partitions=e.zero_allgather_partitions(), allgather_size=e.zero_allgather_bucket_size(),
        max_elements_per_comm=e.zero_reduce_bucket_size(), fp32_allreduce=_fp32_reduce(e),
        fused_nvlink=e.zero_fused_nvlink(), elastic_checkpoint=e.zero_elastic_checkpoint(), **_scaling_and_groups(e))


def _zero2(e, optimizer):
    from .zero.stage2 import FP16_DeepSpeedZeroOptimizer
    return FP16_DeepSpeedZeroOptimizer(
        optimizer, timers=e.timers, contiguous_gradients=e.zero_contiguous_gradients(),
        reduce_bucket_size=e.zero_reduce_bucket_size(), allgather_bucket_size=e.zero_allgather_bucket_size(),
        reduce_scatter=e.zero_reduce_scatter(), overlap_comm=e.zero_overlap_comm(), cpu_offload=e.zero_cpu_offload(),
        pin_memory=True, allreduce_always_fp32=_fp32_reduce(e),
        gradient_accumulation_steps=e.gradient_accumulation_steps(), fused_nvlink=e.zero_fused_nvlink(),
        elastic_checkpoint=e.zero_elastic_checkpoint(), shard_gradients=e.zero_shard_gradient_memory(),
        **_scaling_and_groups(e))


def _zero3(e, optimizer):
    from .zero.stage3 import FP16_DeepSpeedZeroOptimizer_Stage3
    return FP16_DeepSpeedZeroOptimizer_Stage3(
        e.module, optimizer, timers=e.timers, contiguous_gradients=e.zero_contiguous_gradients(),
        reduce_bucket_size=e.zero_reduce_bucket_size(), prefetch_bucket_size=e.zero_prefetch_bucket_size(),
        max_reuse_distance=e.zero_max_reuse_distance(), max_live_parameters=e.zero_max_live_parameters(),
        param_persistence_threshold=e.zero_param_persistence_threshold(), reduce_scatter=e.zero_reduce_scatter(),
        overlap_comm=e.zero_overlap_comm(), offload_optimizer_config=e.zero_offload_optimizer(),
        offload_param_config=e.zero_offload_param(), sub_group_size=e.zero_sub_group_size(),
        gradient_accumulation_steps=e.gradient_accumulation_steps(), aio_config=e.aio_config(),
        fused_allgather_gemm=bool(e._config.zero_config.stage3_fused_allgather_gemm) or None, **_scaling_and_groups(e))


_ZERO_STAGES = {C.ZERO_OPTIMIZATION_OPTIMIZER_STATES: _zero1, C.ZERO_OPTIMIZATION_GRADIENTS: _zero2,
                C.ZERO_OPTIMIZATION_WEIGHTS: _zero3}


def wrap_for_zero(engine, optimizer):
    stage = engine.zero_optimization_stage()
    log_dist("Creating fp16 ZeRO stage {} optimizer".format(stage), ranks=[0])
    if stage not in _ZERO_STAGES:
        raise NotImplementedError("ZeRO stage {} not implemented".format(stage))
    return _ZERO_STAGES[stage](engine, optimizer)
