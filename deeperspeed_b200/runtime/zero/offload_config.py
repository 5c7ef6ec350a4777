"""Offload sub-configs under the reference's import path (deepspeed/runtime/zero/offload_config.py:29-66): plain
dicts of the `offload_param` / `offload_optimizer` blocks with defaults filled in."""
from ..constants import OFFLOAD_PARAM, OFFLOAD_OPTIMIZER, OFFLOAD_PARAM_DEFAULTS, OFFLOAD_OPTIMIZER_DEFAULTS


def get_offload_param_config(param_dict):
    block = param_dict.get(OFFLOAD_PARAM)
    return dict(OFFLOAD_PARAM_DEFAULTS, **block) if block is not None else None


def get_default_offload_param_config():
    return dict(OFFLOAD_PARAM_DEFAULTS)


def get_offload_optimizer_config(param_dict):
    block = param_dict.get(OFFLOAD_OPTIMIZER)
    if block is None:
        return None
    cfg = dict(OFFLOAD_OPTIMIZER_DEFAULTS, **block)
    cfg["pipeline"] = bool(cfg["pipeline_read"] or cfg["pipeline_write"])
    return cfg


def get_default_offload_optimizer_config():
    cfg = dict(OFFLOAD_OPTIMIZER_DEFAULTS)
    cfg["pipeline"] = False
    return cfg
