"""`deepspeed.ops.module_inject` compatibility alias (reference deepspeed/ops/module_inject.py)."""
from ..module_inject.replace_module import replace_transformer_layer, revert_transformer_layer, replace_module  # noqa: F401
