"""Reference module path of the ZeRO config (deepspeed/runtime/zero/config.py); the class is defined with the rest of
the configuration system in runtime/config.py."""
from ..config import DeepSpeedZeroConfig  # noqa: F401
