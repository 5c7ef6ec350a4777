"""Legacy module path (deepspeed/pt/deepspeed_light.py) kept for old scripts."""
from ..runtime.engine import DeepSpeedEngine as DeepSpeedLight  # noqa: F401
from ..runtime.engine import DeepSpeedEngine  # noqa: F401
