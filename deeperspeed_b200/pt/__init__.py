"""Legacy `deepspeed.pt.*` aliases (reference deepspeed/__init__.py:39-49)."""
import sys
from ..runtime import config as deepspeed_config
from ..runtime.fp16 import loss_scaler
from ..runtime import utils as deepspeed_utils
sys.modules[__name__ + ".deepspeed_config"] = deepspeed_config
sys.modules[__name__ + ".loss_scaler"] = loss_scaler
sys.modules[__name__ + ".deepspeed_utils"] = deepspeed_utils
