"""Reference module path of the ZeRO config (deepspeed/runtime/zero/config.py); the class is defined with the rest of
the configuration system in runtime/config.py."""
from ..config import DeepSpeedZeroConfig  # noqa: F401
from .constants import *  # noqa: F401,F403  (stage numbers and key names are imported from here too)
