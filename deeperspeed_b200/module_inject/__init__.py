from .replace_module import replace_transformer_layer, revert_transformer_layer, replace_module
from .inject import module_inject
