from .partition_parameters import (ZeroParamType, ZeroParamStatus, Init, GatheredParameters,
                                   register_external_parameter)
from .tiling import TiledLinear, TiledLinearReturnBias
