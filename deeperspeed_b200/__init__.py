"""deeperspeed_b200 — a B200-native (sm_100a) training engine with the capabilities and public API of
MLRichter/DeeperSpeed (a DeepSpeed 0.3.15 fork). `import deeperspeed_b200 as deepspeed` is a drop-in; calling
`deeperspeed_b200.install_deepspeed_alias()` additionally makes `import deepspeed.*` resolve here (needed to unpickle
reference checkpoints, which store `deepspeed.runtime.fp16.loss_scaler.*` objects).

Public API per reference deepspeed/__init__.py:52-212."""
import sys
import types

from .version import __version__, __version_major__, __version_minor__, __version_patch__, git_hash, git_branch
__git_hash__, __git_branch__ = git_hash, git_branch
version = __version__
from . import ops, elasticity, profiling, module_inject, constants, pt  # reachable as attributes after a bare import
from .runtime.engine import DeepSpeedEngine
from .runtime.engine import ADAM_OPTIMIZER, LAMB_OPTIMIZER  # noqa: F401
from .runtime.pipe.engine import PipelineEngine
from .runtime.lr_schedules import add_tuning_arguments
from .runtime.config import DeepSpeedConfig, DeepSpeedConfigError
from .runtime.activation_checkpointing import checkpointing
from .ops.transformer import DeepSpeedTransformerLayer, DeepSpeedTransformerConfig
from .utils import log_dist, logger
from .utils.distributed import init_distributed
from .runtime import zero
from .pipe import PipelineModule


def initialize(args=None, model=None, optimizer=None, model_parameters=None, training_data=None, lr_scheduler=None,
               mpu=None, dist_init_required=None, collate_fn=None, config_params=None):
    """Initialize the engine. Returns (engine, optimizer, training_dataloader, lr_scheduler)."""
    log_dist("DeepSpeed info: version={}, git-hash={}, git-branch={}".format(__version__, git_hash, git_branch), ranks=[0])
    assert model is not None, "deepspeed.initialize requires a model"
    if not isinstance(model, PipelineModule):
        engine = DeepSpeedEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                 training_data=training_data, lr_scheduler=lr_scheduler, mpu=mpu,
                                 dist_init_required=dist_init_required, collate_fn=collate_fn,
                                 config_params=config_params)
    else:
        assert mpu is None, "mpu must be None with pipeline parallelism"
        engine = PipelineEngine(args=args, model=model, optimizer=optimizer, model_parameters=model_parameters,
                                training_data=training_data, lr_scheduler=lr_scheduler, mpu=model.mpu(),
                                dist_init_required=dist_init_required, collate_fn=collate_fn,
                                config_params=config_params)
    return engine, engine.optimizer, engine.training_dataloader, engine.lr_scheduler


def _add_core_arguments(parser):
    group = parser.add_argument_group("DeepSpeed", "DeepSpeed configurations")
    group.add_argument("--deepspeed", default=False, action="store_true",
                       help="Enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepspeed_config", default=None, type=str, help="DeepSpeed json configuration file.")
    group.add_argument("--deepscale", default=False, action="store_true",
                       help="Deprecated enable DeepSpeed (helper flag for user code, no impact on DeepSpeed backend)")
    group.add_argument("--deepscale_config", default=None, type=str, help="Deprecated DeepSpeed json configuration file.")
    group.add_argument("--deepspeed_mpi", default=False, action="store_true",
                       help="Run via MPI, this will attempt to discover the necessary variables to initialize torch "
                            "distributed from the MPI environment")
    return parser


def add_config_arguments(parser):
    """Update the argument parser to enable parsing of DeepSpeed command line arguments."""
    return _add_core_arguments(parser)


def install_deepspeed_alias(force=False):
    """Register `deepspeed` (and its reference sub-module paths) as aliases of this package."""
    from . import compat
    compat.install(force=force)
