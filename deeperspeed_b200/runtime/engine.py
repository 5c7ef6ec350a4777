"""DeepSpeedEngine: wraps an nn.Module with distributed setup, precision handling, optimizer / scheduler wiring
(ZeRO 1/2/3, FP16 fused/unfused, AMP-less fp32), gradient accumulation, reduction, timers and checkpointing.

Public behaviour per reference deepspeed/runtime/engine.py:102-1915. Differences that matter on B200:
  * device-agnostic (CUDA+NCCL or CPU+gloo), so the whole engine runs in CI without a GPU;
  * gradients live in flat arenas (runtime/flat.py): reductions are in-place collectives on arena buckets;
  * overflow / norm / unscale / clip / cast are fused into device kernels with one host read per step;
  * timers use CUDA events (no stream serialisation).
"""
import hashlib
import os
from collections import OrderedDict
from shutil import copyfile

import torch
import torch.distributed as dist
from torch.nn.modules import Module

from ..utils.logging import logger, log_dist
from ..utils.timer import SynchronizedWallClockTimer, ThroughputTimer
from ..utils.distributed import init_distributed
from . import constants as C
from .constants import (ADAM_OPTIMIZER, ADAMW_OPTIMIZER, LAMB_OPTIMIZER, ONEBIT_ADAM_OPTIMIZER,  # noqa: F401
                        ONEBIT_LAMB_OPTIMIZER, DEEPSPEED_OPTIMIZERS, ROUTE_TRAIN, ROUTE_EVAL, ROUTE_PREDICT)
from .config import DeepSpeedConfig
from .csr_tensor import CSRTensor
from .dataloader import DeepSpeedDataLoader
from .fp16.fused_optimizer import FP16_Optimizer
from .fp16.unfused_optimizer import FP16_UnfusedOptimizer
from .progressive_layer_drop import ProgressiveLayerDrop
from . import lr_schedules
from . import utils as ds_utils
from . import optimizer_factory
from .engine_train import TrainStepMixin, _get_global_rank, is_zero_supported_optimizer, split_half_float_double_csr
from .engine_checkpoint import CheckpointMixin

MEMORY_OPT_ALLREDUCE_SIZE = 500000000
DeepSpeedOptimizerCallable = None

try:
    from torch.utils.tensorboard import SummaryWriter as _SummaryWriter
except Exception:  # pragma: no cover
    _SummaryWriter = None


from .zero.utils import _initialize_parameter_parallel_groups  # noqa: E402  (same helper, one implementation)


def print_configuration(args, name):
    """Every attribute of `args`, one per line, names padded with dots to a common column."""
    logger.info(f"{name}:")
    for key in sorted(vars(args)):
        logger.info("  {} {} {}".format(key, "." * (29 - len(key)), getattr(args, key)))


# what a fresh engine knows before any configuration is read: attribute -> initial value
_INITIAL_STATE = dict(
    data_parallel_group=None, global_steps=0, global_samples=0, micro_steps=0, skipped_steps=0, gradient_average=True,
    warn_unscaled_loss=True, loaded_checkpoint_mp_world_size=None, loaded_checkpoint_dp_world_size=None,
    enable_backward_allreduce=True, progressive_layer_drop=None, store_gradients=False, store_gradients_cpu=False,
    stored_gradients=None, optimizer=None, lr_scheduler=None, save_non_zero_checkpoint=False,
    save_zero_checkpoint=False)


class _CallableInt(int):
    def __call__(self):
        return int(self)


class DeepSpeedEngine(TrainStepMixin, CheckpointMixin, Module):
    r"""DeepSpeed engine for training."""

    def __init__(self, args, model, optimizer=None, model_parameters=None, training_data=None, lr_scheduler=None,
                 mpu=None, dist_init_required=None, collate_fn=None, config_params=None, dont_change_device=False):
        super().__init__()
        # ---- what the caller handed in
        self.client_optimizer, self.client_model_parameters = optimizer, model_parameters
        self.client_lr_scheduler, self.training_data, self.collate_fn = lr_scheduler, training_data, collate_fn
        self.mpu, self.config_params, self.dont_change_device = mpu, config_params, dont_change_device
        for attr, value in _INITIAL_STATE.items():
            setattr(self, attr, value)
        self.layer_outputs, self._hook_handles = {}, []
        self.dist_backend = "nccl" if torch.cuda.is_available() else "gloo"

        # ---- process group, configuration, device
        if dist_init_required is None:
            dist_init_required = not dist.is_initialized()
        if dist_init_required:
            init_distributed(dist_backend=self.dist_backend)
        else:
            assert dist.is_initialized() is True, "Torch distributed not initialized. Please set dist_init_required " \
                "to True or initialize before calling deepspeed.initialize()"
        self._do_args_sanity_check(args)
        self._configure_with_arguments(args, mpu)
        self._do_sanity_check()
        self._set_distributed_vars()
        if self.tensorboard_enabled() and self.global_rank == 0:
            self.summary_writer = self.get_summary_writer()

        # ---- model, timers, data
        self._configure_distributed_model(model)
        self.timers = SynchronizedWallClockTimer()
        self.tput_timer = ThroughputTimer(batch_size=self.train_micro_batch_size_per_gpu(),
                                          num_workers=self.dp_world_size, steps_per_output=self.steps_per_print(),
                                          monitor_memory=False)
        self.training_dataloader = self.deepspeed_io(training_data) if training_data else None

        # ---- optimizer stack and schedule
        if model_parameters or optimizer:
            self._configure_optimizer(optimizer, model_parameters)
            self._configure_lr_scheduler(lr_scheduler)
            self._report_progress(0)

        # ---- embedding gradients that travel as CSR, checkpoint roles, progressive layer drop
        self.csr_tensor_module_names = set()
        if self.sparse_gradients_enabled():
            for name, module in self.module.named_modules():
                if isinstance(module, torch.nn.Embedding):
                    self.csr_tensor_module_names.add(name + ".weight")
                    logger.info("Will convert {} to sparse (csr) tensor during training".format(name))
        self._configure_checkpointing(dist_init_required)
        if self.pld_enabled():
            self.progressive_layer_drop = self._configure_progressive_layer_drop()
        if self.global_rank == 0:
            self._config.print("DeepSpeedEngine configuration")
            if self.dump_state():
                print_configuration(self, "DeepSpeedEngine")

    # ------------------------------------------------------------------ DeeperSpeed extras
    def register_forward_hook(self, layers_to_hook, layer_name_pattern="transformerlayer"):
        """DeeperSpeed extra (reference runtime/engine.py:226-254): forward hooks on every module whose class name
        matches `layer_name_pattern`; after a forward, `self.layer_outputs[key]` holds the (CPU) outputs of the layers
        selected by `layers_to_hook` (a list of `layer_number`s, or "all"). key = module.layer_number when the
        module has one, else its position among the matching modules. `self.layers_to_hook` can be changed between
        calls without re-registering (PipelineEngine.train_batch / eval_batch / inference_batch do that)."""
        import re
        for h in self._hook_handles:
            h.remove()
        self._hook_handles = []
        self.layer_outputs = {}
        self.layers_to_hook = layers_to_hook
        pattern = re.compile(layer_name_pattern, re.IGNORECASE)
        idx = 0
        for name, module in self.module.named_modules():
            if pattern.search(module.__class__.__name__):
                def _hook(mod, inp, out, _idx=idx):
                    key = getattr(mod, "layer_number", _idx)
                    want = self.layers_to_hook
                    if not (want == "all" or (want and int(key) in want)):
                        return
                    outs = out if isinstance(out, (tuple, list)) else (out,)
                    self.layer_outputs[key] = [o.detach().cpu() if torch.is_tensor(o) else o for o in outs]
                self._hook_handles.append(module.register_forward_hook(_hook))
                idx += 1

    def swap_tensor_config(self):
        return getattr(self._config, "swap_tensor_config", None)

    def autotune_gemms(self, tokens_per_micro_batch, iters=7, verbose=False):
        """Time every GEMM kernel variant (single-CTA / CTA-pair / tail-wave K split) for the linear layers of this
        rank's module at `tokens_per_micro_batch` rows and let the native dispatcher use the winners. The reference
        does the equivalent cuBLAS algorithm search when a transformer layer is created (gemm_test.h:27-293).
        No-op without a GPU. Returns {(M, N, K, trans_a, trans_b): (variant, [ms])}."""
        from ..ops import gemm_autotune
        dtype = torch.bfloat16 if self.precision() == "bfloat16" else torch.float16
        return gemm_autotune.tune_module(self.module, int(tokens_per_micro_batch), dtype=dtype, iters=iters,
                                         verbose=verbose)

    def set_train_batch_size(self, train_batch_size):
        """Adjust the global batch size by changing gradient accumulation (micro-batch and dp stay fixed)."""
        if train_batch_size % (self.train_micro_batch_size_per_gpu() * self.dp_world_size) != 0:
            raise ValueError("Train batch size must be divisible by micro-batch data parallelism")
        new_gas = train_batch_size // (self.train_micro_batch_size_per_gpu() * self.dp_world_size)
        self._config.train_batch_size = train_batch_size
        self._config.gradient_accumulation_steps = new_gas

    # ------------------------------------------------------------------ config accessors
    def get_batch_info(self):
        """(train_batch_size, train_micro_batch_size_per_gpu, gradient_accumulation_steps) as the docstring of the
        reference promises (engine.py:256-267). The reference actually returns the three accessor METHODS; the values
        returned here can still be called, so code written against either reading works."""
        return tuple(_CallableInt(v()) for v in (self.train_batch_size, self.train_micro_batch_size_per_gpu,
                                                 self.gradient_accumulation_steps))

    def pld_theta(self):
        return self.pld_params()[C.PLD_THETA]

    def pld_gamma(self):
        return self.pld_params()[C.PLD_GAMMA]

    def get_summary_writer(self, name="DeepSpeedJobName", base=os.path.join(os.path.expanduser("~"), "tensorboard")):
        """TensorBoard writer under <output_path>/<job_name>, or (no output path configured)
        ~/tensorboard/<cluster job id>/logs/<job name>."""
        if _SummaryWriter is None:
            logger.warning("tensorboard is not installed; tensorboard logging disabled")
            self._config.tensorboard_enabled = False
            return None
        job = self.tensorboard_job_name()
        if self.tensorboard_output_path():
            log_dir = os.path.join(self.tensorboard_output_path(), job)
        else:
            cluster_job = next((os.environ[k] for k in ("DLWS_JOB_ID", "DLTS_JOB_ID") if k in os.environ),
                               "unknown-job-id")
            log_dir = os.path.join(base, cluster_job, "logs", job or name)
        os.makedirs(log_dir, exist_ok=True)
        return _SummaryWriter(log_dir=log_dir)

    def optimizer_name(self):
        return self.client_optimizer.__class__.__name__ if self.client_optimizer else self._config.optimizer_name

    def zero_cpu_offload(self):
        return self._config.zero_config.offload_optimizer is not None

    def zero_optimization_partition_gradients(self):
        return self.zero_optimization_stage() >= C.ZERO_OPTIMIZATION_GRADIENTS

    def zero_optimization_partition_weights(self):
        return self.zero_optimization_stage() >= C.ZERO_OPTIMIZATION_WEIGHTS

    def zero_fused_nvlink(self):
        v = self._config.zero_config.fused_nvlink
        if v == "auto":
            return torch.cuda.is_available() and self.dp_world_size > 1 and os.environ.get("DSB_FUSED_NVLINK", "1") != "0"
        return bool(v)

    def zero_shard_gradient_memory(self):
        """Stage 2: keep only the owned gradient slices (see ZERO_DEFAULTS["gradient_memory"])."""
        mode = str(self._config.zero_config.gradient_memory).lower()
        if mode not in ("auto", "full", "sharded"):
            raise ValueError(f"zero_optimization.gradient_memory must be auto, full or sharded, got {mode!r}")
        if mode != "auto":
            return mode == "sharded"
        arena_bytes = sum(p.numel() * p.element_size() for p in self.module.parameters() if p.requires_grad)
        return self.dp_world_size > 1 and arena_bytes > (8 << 30)

    def postscale_gradients(self):
        return not self._config.prescale_gradients

    def dynamic_loss_scale(self):
        return self._config.loss_scale == 0

    # ------------------------------------------------------------------ setup
    def _configure_lr_scheduler(self, client_lr_scheduler):
        """The scheduler named in the config wins over the one passed to initialize()."""
        configured = self._scheduler_from_config(self.optimizer)
        self.lr_scheduler = configured or client_lr_scheduler
        if self.global_rank == 0:
            logger.info(f"DeepSpeed using configured LR scheduler = {self.scheduler_name()}" if configured
                        else "DeepSpeed using client LR scheduler")
        log_dist(f"DeepSpeed LR Scheduler = {self.lr_scheduler}", ranks=[0])

    def _configure_checkpointing(self, dist_init_required):
        """Who writes what: module states come from the first data-parallel replica (every replica under ZeRO-3, where
        each holds different shards); every ZeRO rank writes its own optimizer shard."""
        replica = self.mpu.get_data_parallel_rank() if self.mpu else self.global_rank
        self.save_non_zero_checkpoint = replica == 0 or self.zero_optimization_partition_weights()
        if self.zero_optimization():
            shard_group = getattr(self.optimizer, "dp_process_group", None)
            shard_rank = dist.get_rank(group=shard_group) if shard_group is not None else replica
            self.save_zero_checkpoint = shard_rank == replica

    def _scheduler_from_config(self, optimizer):
        """Instantiate `scheduler.type`: one of ours (runtime/lr_schedules.py) or any torch.optim.lr_scheduler class."""
        kind = self.scheduler_name()
        if kind is None:
            return None
        ours = getattr(lr_schedules, kind, None)
        if ours is not None:
            return ours(optimizer, **self.scheduler_params())
        assert hasattr(torch.optim.lr_scheduler, kind), f"DeepSpeed does not recognize LR scheduler {kind}"
        # torch schedulers insist on a torch.optim.Optimizer: hand them the innermost one
        return getattr(torch.optim.lr_scheduler, kind)(lr_schedules.get_torch_optimizer(optimizer),
                                                       **self.scheduler_params())

    def _set_distributed_vars(self):
        if torch.cuda.is_available():
            if self.local_rank >= 0:
                torch.cuda.set_device(self.local_rank)
            self.device = torch.device("cuda", self.local_rank) if self.local_rank >= 0 else torch.device("cuda")
        else:
            self.device = torch.device("cpu")
        self.world_size, self.global_rank = dist.get_world_size(), dist.get_rank()

    def _configure_with_arguments(self, args, mpu):
        """LOCAL_RANK from the launcher's environment (falling back to args.local_rank) and the parsed configuration."""
        fallback = getattr(args, "local_rank", 0) if args is not None else 0
        self.local_rank = int(os.environ.get("LOCAL_RANK", fallback))
        if hasattr(args, "local_rank"):
            args.local_rank = self.local_rank
        config_file = None
        if args is not None:
            config_file = getattr(args, "deepspeed_config", None) or getattr(args, "deepscale_config", None)
        self._config = DeepSpeedConfig(config_file, mpu, param_dict=self.config_params)

    def _do_args_sanity_check(self, args):
        """`--deepscale_config` is the pre-rename spelling of `--deepspeed_config`; without `config_params` a config
        file must be named and exist."""
        legacy = getattr(args, "deepscale_config", None)
        if legacy is not None:
            logger.warning("************ --deepscale_config is deprecated, please use --deepspeed_config ************")
            assert getattr(args, "deepspeed_config", None) is None, \
                "Not sure how to proceed, we were given both a deepscale_config and deepspeed_config"
            args.deepspeed_config = legacy
        if self.config_params is not None:
            return
        path = getattr(args, "deepspeed_config", None) if args is not None else None
        assert path is not None, "DeepSpeed requires --deepspeed_config to specify configuration file"
        assert os.path.isfile(path), "DeepSpeed configuration file: {} is not an existing file".format(path)

    def _is_supported_optimizer(self, optimizer_name):
        return optimizer_name in C.DEEPSPEED_OPTIMIZERS or getattr(torch.optim, optimizer_name, None) is not None

    def _do_sanity_check(self):
        name = self.optimizer_name()
        if not self.client_optimizer and name is not None:
            assert self._is_supported_optimizer(name), "{} is not a supported DeepSpeed Optimizer".format(name)
        if name in (C.LAMB_OPTIMIZER, C.ONEBIT_LAMB_OPTIMIZER):
            assert self.dynamic_loss_scale() or self.precision() != torch.float16, \
                "DeepSpeed {} optimizer requires dynamic loss scaling".format(name)

    def _broadcast_model(self):
        """Every data-parallel replica starts from rank 0's weights (ZeRO-3 shards that are not materialised here
        are skipped: they were scattered from one source when they were partitioned)."""
        from .zero.partition_parameters import ZeroParamStatus
        for p in self.module.parameters():
            if hasattr(p, "ds_status") and p.ds_status is not ZeroParamStatus.AVAILABLE:
                continue
            if torch.is_tensor(p):
                dist.broadcast(p.data, self.broadcast_src_rank, group=self.data_parallel_group)

    def _configure_distributed_model(self, model):
        self.module = model
        if self.fp16_enabled() and self.precision() != torch.float32:
            self.module.to(self.precision())
        if not self.dont_change_device:
            self.module.to(self.device)
        if self.mpu is not None:
            group = self.mpu.get_data_parallel_group()
            self.data_parallel_group = group
            self.dp_world_size = self.mpu.get_data_parallel_world_size()
            self.mp_world_size = self.mpu.get_model_parallel_world_size()
            self.broadcast_src_rank = _get_global_rank(group, 0)
        else:
            self.data_parallel_group = _initialize_parameter_parallel_groups()
            self.dp_world_size, self.mp_world_size, self.broadcast_src_rank = dist.get_world_size(), 1, 0
        if not self.amp_enabled():
            self._broadcast_model()

    # ------------------------------------------------------------------ optimizer wiring (runtime/optimizer_factory.py)
    def _configure_optimizer(self, client_optimizer, model_parameters):
        """basic optimizer (client's, or built from the config) -> ZeRO stage | reduced-precision wrapper | as is."""
        if client_optimizer is None:
            basic = self._configure_basic_optimizer(model_parameters)
            origin = "DeepSpeed Optimizer param name {}".format(self.optimizer_name())
        else:
            if callable(client_optimizer) and not isinstance(client_optimizer, torch.optim.Optimizer):
                client_optimizer = client_optimizer(model_parameters)
            client_optimizer.param_groups[:] = [g for g in client_optimizer.param_groups if len(g["params"]) != 0]
            basic, origin = client_optimizer, "client Optimizer"
        if self.global_rank == 0:
            logger.info(f"Using {origin} as basic optimizer")
            logger.info("DeepSpeed Basic Optimizer = {}".format(basic.__class__.__name__))
        self.basic_optimizer = basic      # the inner optimizer, whatever wrapper ends up in self.optimizer

        if self.zero_optimization():
            assert not self.amp_enabled(), "Amp and ZeRO are not currently compatible, please use (legacy) fp16 mode " \
                "which performs similar to amp opt_mode=O2"
            if not is_zero_supported_optimizer(basic):
                assert self.zero_allow_untested_optimizer(), 'You are using an untested ZeRO Optimizer. Please add ' \
                    '<"zero_allow_untested_optimizer": true> in the configuration file to use it.'
                if self.global_rank == 0:
                    logger.warning("**** You are using ZeRO with an untested optimizer, proceed with caution *****")
            self.optimizer = self._configure_zero_optimizer(basic)
        elif self.amp_enabled():
            assert not (self.fp16_enabled() and self.precision() != torch.float32), \
                "Cannot enable both amp with (legacy) fp16 mode"
            from .amp import NativeAmp
            self.amp = NativeAmp(self.amp_params(), self.device)      # torch.autocast + GradScaler (no apex needed)
            logger.info(f"Initializing AMP with these params: {self.amp_params()}")
            self.optimizer = basic
            self._broadcast_model()      # deferred in _configure_distributed_model when amp is on, as in the reference
        elif self.fp16_enabled() and self.precision() != torch.float32:
            self.optimizer = self._configure_fp16_optimizer(basic)
        else:
            self.optimizer = basic
        log_dist("DeepSpeed Final Optimizer = {}".format(self.optimizer_name()), ranks=[0])

    def _configure_basic_optimizer(self, model_parameters):
        return optimizer_factory.build_basic_optimizer(self, model_parameters)

    def _configure_fp16_optimizer(self, optimizer):
        return optimizer_factory.wrap_for_reduced_precision(self, optimizer)

    def _configure_zero_optimizer(self, optimizer):
        return optimizer_factory.wrap_for_zero(self, optimizer)

    def _configure_progressive_layer_drop(self):
        return ProgressiveLayerDrop(theta=self.pld_theta(), gamma=self.pld_gamma())

    # ------------------------------------------------------------------ data
    def deepspeed_io(self, dataset, batch_size=None, route=C.ROUTE_TRAIN, pin_memory=True, data_sampler=None,
                     collate_fn=None, num_local_io_workers=None):
        """A DeepSpeedDataLoader over `dataset` sharded by data-parallel rank (training), or read sequentially
        (evaluation / prediction routes); training loaders feed the throughput timer."""
        if not isinstance(dataset, torch.utils.data.Dataset):
            raise ValueError("Training data must be a torch Dataset")
        if data_sampler is None and route in (C.ROUTE_PREDICT, C.ROUTE_EVAL):
            data_sampler = torch.utils.data.SequentialSampler(dataset)
        if self.mpu is not None:
            replicas, my_replica = self.mpu.get_data_parallel_world_size(), self.mpu.get_data_parallel_rank()
        else:
            replicas, my_replica = self.dp_world_size, dist.get_rank(group=self.data_parallel_group)
        return DeepSpeedDataLoader(dataset=dataset, batch_size=batch_size or self.train_micro_batch_size_per_gpu(),
                                   pin_memory=pin_memory, collate_fn=collate_fn or self.collate_fn,
                                   local_rank=self.local_rank,
                                   tput_timer=self.tput_timer if route == C.ROUTE_TRAIN else None,
                                   num_local_io_workers=num_local_io_workers, data_sampler=data_sampler,
                                   data_parallel_world_size=replicas, data_parallel_rank=my_replica)

    def train(self, mode=True):
        self.warn_unscaled_loss = True
        self.module.train(mode)

    def eval(self):
        self.warn_unscaled_loss = True
        self.module.train(False)

    def _scale_loss(self, prescaled_loss):
        """loss / gradient_accumulation_steps for a tensor or every tensor of a tuple / list (other types pass through
        with one warning)."""
        gas = self.gradient_accumulation_steps()
        if isinstance(prescaled_loss, torch.Tensor):
            return prescaled_loss / gas
        if isinstance(prescaled_loss, (tuple, list)):
            return [l / gas if isinstance(l, torch.Tensor) else l for l in prescaled_loss]
        if self.warn_unscaled_loss:
            logger.warning(f"DeepSpeed unable to scale loss because of type: {type(prescaled_loss)}")
            self.warn_unscaled_loss = False
        return prescaled_loss


# ----------------------------------------------------------------------------------------------- generated accessors
# The reference exposes one zero-argument method per configuration value (engine.train_batch_size(), ...). Here they are
# generated from a table: (method name, attribute path below engine._config).
_CONFIG_ACCESSORS = (
    ("checkpoint_tag_validation_enabled", "checkpoint_tag_validation_enabled"),
    ("checkpoint_tag_validation_fail", "checkpoint_tag_validation_fail"),
    ("elasticity_enabled", "elasticity_enabled"), ("pld_enabled", "pld_enabled"), ("pld_params", "pld_params"),
    ("tensorboard_enabled", "tensorboard_enabled"), ("tensorboard_output_path", "tensorboard_output_path"),
    ("tensorboard_job_name", "tensorboard_job_name"), ("wall_clock_breakdown", "wall_clock_breakdown"),
    ("flops_profiler_enabled", "flops_profiler_config.enabled"),
    ("flops_profiler_profile_step", "flops_profiler_config.profile_step"),
    ("flops_profiler_module_depth", "flops_profiler_config.module_depth"),
    ("flops_profiler_top_modules", "flops_profiler_config.top_modules"),
    ("flops_profiler_detailed", "flops_profiler_config.detailed"), ("memory_breakdown", "memory_breakdown"),
    ("sparse_gradients_enabled", "sparse_gradients_enabled"), ("train_batch_size", "train_batch_size"),
    ("train_micro_batch_size_per_gpu", "train_micro_batch_size_per_gpu"), ("optimizer_params", "optimizer_params"),
    ("optimizer_legacy_fusion", "optimizer_legacy_fusion"), ("scheduler_name", "scheduler_name"),
    ("scheduler_params", "scheduler_params"), ("zero_optimization", "zero_enabled"),
    ("zero_allow_untested_optimizer", "zero_allow_untested_optimizer"),
    ("zero_reduce_scatter", "zero_config.reduce_scatter"), ("zero_overlap_comm", "zero_config.overlap_comm"),
    ("zero_offload_optimizer", "zero_config.offload_optimizer"),
    ("zero_offload_param", "zero_config.offload_param"), ("zero_sub_group_size", "zero_config.sub_group_size"),
    ("zero_optimization_stage", "zero_optimization_stage"),
    ("zero_reduce_bucket_size", "zero_config.reduce_bucket_size"),
    ("zero_allgather_bucket_size", "zero_config.allgather_bucket_size"),
    ("zero_contiguous_gradients", "zero_config.contiguous_gradients"),
    ("zero_load_from_fp32_weights", "zero_config.load_from_fp32_weights"),
    ("zero_elastic_checkpoint", "zero_config.elastic_checkpoint"),
    ("zero_max_live_parameters", "zero_config.max_live_parameters"),
    ("zero_max_reuse_distance", "zero_config.max_reuse_distance"),
    ("zero_prefetch_bucket_size", "zero_config.prefetch_bucket_size"),
    ("zero_param_persistence_threshold", "zero_config.param_persistence_threshold"),
    ("zero_gather_fp16_weights_on_model_save", "zero_config.gather_fp16_weights_on_model_save"),
    ("fp16_enabled", "fp16_enabled"), ("precision", "precision"), ("amp_enabled", "amp_enabled"),
    ("amp_params", "amp_params"), ("loss_scale", "loss_scale"),
    ("gradient_accumulation_steps", "gradient_accumulation_steps"), ("allreduce_always_fp32", "fp32_allreduce"),
    ("gradient_predivide_factor", "gradient_predivide_factor"), ("steps_per_print", "steps_per_print"),
    ("zero_allgather_partitions", "zero_config.allgather_partitions"), ("dump_state", "dump_state"),
    ("gradient_clipping", "gradient_clipping"), ("initial_dynamic_scale", "initial_dynamic_scale"),
    ("dynamic_loss_scale_args", "dynamic_loss_scale_args"), ("aio_config", "aio_config"),
)


def _make_config_accessor(path):
    parts = path.split(".")

    def accessor(self):
        value = self._config
        for part in parts:
            value = getattr(value, part)
        return value

    return accessor


for _name, _path in _CONFIG_ACCESSORS:
    _fn = _make_config_accessor(_path)
    _fn.__name__ = _name
    _fn.__doc__ = f"Configuration value `{_path}`."
    setattr(DeepSpeedEngine, _name, _fn)
del _name, _path, _fn
