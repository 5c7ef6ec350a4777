"""Native-op front ends. The same names are reachable after a bare `import deepspeed` as in the reference
(deepspeed/ops/__init__.py): `ops.adam`, `ops.lamb`, `ops.sparse_attention`, `ops.transformer`, the transformer layer
classes and `replace_module`."""
from . import adam, lamb, sparse_attention, transformer
from .transformer import DeepSpeedTransformerLayer, DeepSpeedTransformerConfig
from .module_inject import replace_module

from ..git_version_info import compatible_ops as __compatible_ops__  # noqa: E402,F401
