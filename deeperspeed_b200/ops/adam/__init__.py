from .fused_adam import FusedAdam
from .cpu_adam import DeepSpeedCPUAdam
