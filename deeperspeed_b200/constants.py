"""Package-level constants under the reference's import path (deepspeed/constants.py:10-16)."""
from datetime import timedelta

from .runtime.constants import TORCH_DISTRIBUTED_DEFAULT_PORT, DEFAULT_PROCESS_GROUP_TIMEOUT_MINUTES

# Default process group wide timeout, if applicable (only for the nccl backend with blocking wait / async error
# handling enabled).
default_pg_timeout = timedelta(minutes=DEFAULT_PROCESS_GROUP_TIMEOUT_MINUTES)
