"""Config helpers under the reference's import path (deepspeed/runtime/config_utils.py:13-76); the implementations live
in runtime/config.py."""
from .config import (ScientificNotationEncoder, DeepSpeedConfigObject, get_scalar_param,  # noqa: F401
                     dict_raise_error_on_duplicate_keys)


def get_list_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)


def get_dict_param(param_dict, param_name, param_default_value):
    return param_dict.get(param_name, param_default_value)
