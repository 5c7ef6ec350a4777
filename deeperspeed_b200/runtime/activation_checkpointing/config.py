"""Activation-checkpointing sub-config under the reference's import path
(deepspeed/runtime/activation_checkpointing/config.py:27-103)."""
from ..config import DeepSpeedActivationCheckpointingConfig  # noqa: F401
from ..constants import ACT_CHKPT, ACT_CHKPT_DEFAULTS  # noqa: F401

ACT_CHKPT_PARTITION_ACTIVATIONS = "partition_activations"
ACT_CHKPT_NUMBER_CHECKPOINTS = "number_checkpoints"
ACT_CHKPT_CONTIGUOUS_MEMORY_OPTIMIZATION = "contiguous_memory_optimization"
ACT_CHKPT_SYNCHRONIZE_CHECKPOINT_BOUNDARY = "synchronize_checkpoint_boundary"
ACT_CHKPT_PROFILE = "profile"
ACT_CHKPT_CPU_CHECKPOINTING = "cpu_checkpointing"
ACT_CHKPT_DEFAULT = dict(ACT_CHKPT_DEFAULTS)
