"""`deepspeed.ops.module_inject` compatibility alias (reference deepspeed/ops/module_inject.py)."""
from ..module_inject.replace_module import (replace_transformer_layer, revert_transformer_layer,  # noqa: F401
                                            replace_layers_with_config)
from .transformer import DeepSpeedTransformerConfig


def replace_module(orig_module_impl, model, replacement_module_config):
    """Replace the client's layers of class `orig_module_impl` by DeepSpeedTransformerLayer built from
    `replacement_module_config` (reference ops/module_inject.py:72-88)."""
    assert isinstance(replacement_module_config, DeepSpeedTransformerConfig), \
        "Only DeepSpeedTransformerConfig is currently supported as replacement config"
    return replace_layers_with_config(orig_module_impl, model, replacement_module_config)


def revert_module(orig_module_impl, model, orig_module_config, replacement_module_config):
    """Put the client's original layers back in place of DeepSpeedTransformerLayer (reference ops/module_inject.py:160-182)."""
    return revert_transformer_layer(orig_module_impl, model, orig_module_config,
                                    preln=bool(getattr(replacement_module_config, "pre_layer_norm", False)))
