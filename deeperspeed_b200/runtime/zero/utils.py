"""ZeRO helpers under the reference's import path (names per deepspeed/runtime/zero/utils.py:8-55)."""
import torch
import torch.distributed as dist

from ...ops.adam import DeepSpeedCPUAdam, FusedAdam
from ...utils.logging import logger


def _initialize_parameter_parallel_groups(parameter_parallel_size=None):
    """Cut the world into blocks of `parameter_parallel_size` consecutive ranks (default: one block) and return the
    process group of the block this rank is in. Collective: every rank creates every block's group."""
    world, me = dist.get_world_size(), dist.get_rank()
    size = parameter_parallel_size or world
    logger.info("data_parallel_size: %s, parameter_parallel_size: %s", world, size)
    assert world % size == 0, "world size should be divisible by parameter parallel size"
    groups = [dist.new_group(range(first, first + size)) for first in range(0, world, size)]
    return groups[me // size]


def _supported_optimizer_types():
    kinds = [torch.optim.Adam, torch.optim.AdamW, FusedAdam, DeepSpeedCPUAdam]
    try:                      # apex is optional
        import apex
        kinds.append(apex.optimizers.FusedAdam)
    except ImportError:
        pass
    return kinds


ZERO_SUPPORTED_OPTIMIZERS = _supported_optimizer_types()


def is_zero_supported_optimizer(optimizer):
    """Exact type match (a subclass may change the update rule ZeRO's sharded step relies on)."""
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(f"Checking ZeRO support for optimizer={optimizer.__class__.__name__} type={type(optimizer)}")
    return type(optimizer) in ZERO_SUPPORTED_OPTIMIZERS
