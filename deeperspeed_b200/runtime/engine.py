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


def _initialize_parameter_parallel_groups(parameter_parallel_size=None):
    data_parallel_size = int(dist.get_world_size())
    if parameter_parallel_size is None:
        parameter_parallel_size = int(data_parallel_size)
    assert data_parallel_size % parameter_parallel_size == 0, "world size should be divisible by parameter parallel size"
    rank = dist.get_rank()
    my_group = None
    for i in range(dist.get_world_size() // parameter_parallel_size):
        ranks = range(i * parameter_parallel_size, (i + 1) * parameter_parallel_size)
        group = torch.distributed.new_group(ranks)
        if rank in ranks:
            my_group = group
    return my_group


def print_configuration(args, name):
    logger.info("{}:".format(name))
    for arg in sorted(vars(args)):
        dots = "." * (29 - len(arg))
        logger.info("  {} {} {}".format(arg, dots, getattr(args, arg)))


class DeepSpeedEngine(TrainStepMixin, CheckpointMixin, Module):
    r"""DeepSpeed engine for training."""

    def __init__(self, args, model, optimizer=None, model_parameters=None, training_data=None, lr_scheduler=None,
                 mpu=None, dist_init_required=None, collate_fn=None, config_params=None, dont_change_device=False):
        super().__init__()
        self.dont_change_device = dont_change_device
        self.client_optimizer = optimizer
        self.client_model_parameters = model_parameters
        self.client_lr_scheduler = lr_scheduler
        self.training_data = training_data
        self.collate_fn = collate_fn
        self.mpu = mpu
        self.data_parallel_group = None
        self.global_steps = 0
        self.global_samples = 0
        self.micro_steps = 0
        self.skipped_steps = 0
        self.gradient_average = True
        self.warn_unscaled_loss = True
        self.config_params = config_params
        self.loaded_checkpoint_mp_world_size = None
        self.loaded_checkpoint_dp_world_size = None
        self.enable_backward_allreduce = True
        self.progressive_layer_drop = None
        self.dist_backend = "nccl" if torch.cuda.is_available() else "gloo"
        self.store_gradients = False
        self.store_gradients_cpu = False
        self.stored_gradients = None
        self.layer_outputs = {}
        self._hook_handles = []

        if dist_init_required is None:
            dist_init_required = not dist.is_initialized()
        if dist_init_required is False:
            assert dist.is_initialized() is True, "Torch distributed not initialized. Please set dist_init_required to True or initialize before calling deepspeed.initialize()"
        else:
            init_distributed(dist_backend=self.dist_backend)

        self._do_args_sanity_check(args)
        self._configure_with_arguments(args, mpu)
        self._do_sanity_check()
        self._set_distributed_vars()

        if self.tensorboard_enabled() and self.global_rank == 0:
            self.summary_writer = self.get_summary_writer()

        self._configure_distributed_model(model)

        self.timers = SynchronizedWallClockTimer()
        self.tput_timer = ThroughputTimer(batch_size=self.train_micro_batch_size_per_gpu(),
                                          num_workers=self.dp_world_size, steps_per_output=self.steps_per_print(),
                                          monitor_memory=False)
        if training_data:
            self.training_dataloader = self.deepspeed_io(training_data)
        else:
            self.training_dataloader = None

        self.optimizer = None
        self.lr_scheduler = None
        if model_parameters or optimizer:
            self._configure_optimizer(optimizer, model_parameters)
            self._configure_lr_scheduler(lr_scheduler)
            self._report_progress(0)

        self.csr_tensor_module_names = set()
        if self.sparse_gradients_enabled():
            for name, module in self.module.named_modules():
                if isinstance(module, torch.nn.Embedding):
                    self.csr_tensor_module_names.add(name + ".weight")
                    logger.info("Will convert {} to sparse (csr) tensor during training".format(name))

        self.save_non_zero_checkpoint = False
        self.save_zero_checkpoint = False
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
        return self.train_batch_size, self.train_micro_batch_size_per_gpu, self.gradient_accumulation_steps

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

    def postscale_gradients(self):
        return not self._config.prescale_gradients

    def dynamic_loss_scale(self):
        return self._config.loss_scale == 0

    # ------------------------------------------------------------------ setup
    def _configure_lr_scheduler(self, client_lr_scheduler):
        lr_scheduler = self._scheduler_from_config(self.optimizer)
        if lr_scheduler:
            if self.global_rank == 0:
                logger.info(f"DeepSpeed using configured LR scheduler = {self.scheduler_name()}")
            self.lr_scheduler = lr_scheduler
        else:
            if self.global_rank == 0:
                logger.info("DeepSpeed using client LR scheduler")
            self.lr_scheduler = client_lr_scheduler
        log_dist(f"DeepSpeed LR Scheduler = {self.lr_scheduler}", ranks=[0])

    def _configure_checkpointing(self, dist_init_required):
        dp_rank = self.global_rank
        if self.mpu:
            dp_rank = self.mpu.get_data_parallel_rank()
        # only the first data-parallel replica writes module states; ZeRO ranks each write their optimizer shard
        self.save_non_zero_checkpoint = (dp_rank == 0) or self.zero_optimization_partition_weights()
        if self.zero_optimization():
            param_rank = dist.get_rank(group=self.optimizer.dp_process_group) if hasattr(self.optimizer, "dp_process_group") else dp_rank
            self.save_zero_checkpoint = (param_rank == dp_rank)

    def _scheduler_from_config(self, optimizer):
        scheduler_name = self.scheduler_name()
        if scheduler_name is not None:
            if hasattr(lr_schedules, scheduler_name):
                scheduler = getattr(lr_schedules, scheduler_name)
            else:
                assert hasattr(torch.optim.lr_scheduler, scheduler_name), f"DeepSpeed does not recognize LR scheduler {scheduler_name}"
                scheduler = getattr(torch.optim.lr_scheduler, scheduler_name)
                optimizer = lr_schedules.get_torch_optimizer(optimizer)
            return scheduler(optimizer, **self.scheduler_params())
        return None

    def _set_distributed_vars(self):
        if self.local_rank >= 0 and torch.cuda.is_available():
            torch.cuda.set_device(self.local_rank)
            self.device = torch.device("cuda", self.local_rank)
        elif torch.cuda.is_available():
            self.device = torch.device("cuda")
        else:
            self.device = torch.device("cpu")
        self.world_size = dist.get_world_size()
        self.global_rank = dist.get_rank()

    def _configure_with_arguments(self, args, mpu):
        self.local_rank = int(os.environ.get("LOCAL_RANK", getattr(args, "local_rank", 0) if args is not None else 0))
        if hasattr(args, "local_rank"):
            args.local_rank = self.local_rank
        config_file = None
        if args is not None:
            config_file = getattr(args, "deepspeed_config", None)
            if config_file is None and getattr(args, "deepscale_config", None) is not None:
                config_file = args.deepscale_config
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
        if not self.client_optimizer:
            if self.optimizer_name() is not None:
                assert self._is_supported_optimizer(self.optimizer_name()), \
                    "{} is not a supported DeepSpeed Optimizer".format(self.optimizer_name())
        if self.optimizer_name() in (C.LAMB_OPTIMIZER, C.ONEBIT_LAMB_OPTIMIZER):
            assert self.dynamic_loss_scale() or self.precision() != torch.float16, \
                "DeepSpeed {} optimizer requires dynamic loss scaling".format(self.optimizer_name())

    def _broadcast_model(self):
        def is_replicated(p):
            if hasattr(p, "ds_status"):
                from .zero.partition_parameters import ZeroParamStatus
                if p.ds_status is not ZeroParamStatus.AVAILABLE:
                    return False
            return True

        for p in self.module.parameters():
            if torch.is_tensor(p) and is_replicated(p):
                dist.broadcast(p.data, self.broadcast_src_rank, group=self.data_parallel_group)

    def _configure_distributed_model(self, model):
        self.module = model
        if self.fp16_enabled() and self.precision() != torch.float32:
            self.module.to(self.precision())
        if not self.dont_change_device:
            self.module.to(self.device)
        if self.mpu is None:
            self.data_parallel_group = _initialize_parameter_parallel_groups()
            self.dp_world_size = dist.get_world_size()
            self.mp_world_size = 1
            self.broadcast_src_rank = 0
        else:
            self.data_parallel_group = self.mpu.get_data_parallel_group()
            self.dp_world_size = self.mpu.get_data_parallel_world_size()
            self.mp_world_size = self.mpu.get_model_parallel_world_size()
            self.broadcast_src_rank = _get_global_rank(self.mpu.get_data_parallel_group(), 0)
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
            raise RuntimeError("apex AMP is not available in this environment; use the fp16 block (fp16 or bfloat16) "
                               "which gives the O2-style master-weight regime")
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
