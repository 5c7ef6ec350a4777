"""`deepspeed.ops.module_inject` compatibility alias (reference deepspeed/ops/module_inject.py)."""
from ..module_inject.replace_module import replace_transformer_layer, revert_transformer_layer, replace_module  # noqa: F401


def revert_module(orig_module_impl, model, orig_module_config, replacement_module_config):
    """Put the client's original layers back in place of DeepSpeedTransformerLayer (reference ops/module_inject.py:160-182)."""
    return revert_transformer_layer(orig_module_impl, model, orig_module_config,
                                    preln=bool(getattr(replacement_module_config, "pre_layer_norm", False)))
