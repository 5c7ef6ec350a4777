"""ElasticityConfig: the validated `elasticity` block of a config (keys and messages per reference
deepspeed/elasticity/config.py:30-110), parsed with a table of (attribute, key, default) instead of a cascade."""
import json

from .constants import *


class ElasticityError(Exception):
    """Base exception for all elasticity related errors"""


class ElasticityConfigError(ElasticityError):
    """Elasticity configuration error"""


class ElasticityIncompatibleWorldSize(ElasticityError):
    """Attempting to run a world size that is incompatible with a given elastic config"""


# attribute, config key, default; the first two are mandatory once elasticity is enabled
_FIELDS = (
    ("max_acceptable_batch_size", MAX_ACCEPTABLE_BATCH_SIZE, MAX_ACCEPTABLE_BATCH_SIZE_DEFAULT),
    ("micro_batches", MICRO_BATCHES, MICRO_BATCHES_DEFAULT),
    ("min_gpus", MIN_GPUS, MIN_GPUS_DEFAULT),
    ("max_gpus", MAX_GPUS, MAX_GPUS_DEFAULT),
    ("min_time", MIN_TIME, MIN_TIME_DEFAULT),
    ("version", VERSION, VERSION_DEFAULT),
    ("prefer_larger_batch_size", PREFER_LARGER_BATCH, PREFER_LARGER_BATCH_DEFAULT),
    ("ignore_non_elastic_batch_info", IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT),
)
_MANDATORY = (MAX_ACCEPTABLE_BATCH_SIZE, MICRO_BATCHES)


class ElasticityConfig:
    def __init__(self, param_dict):
        self.enabled = param_dict.get(ENABLED, ENABLED_DEFAULT)
        for attr, key, default in _FIELDS:
            if self.enabled and key in _MANDATORY and key not in param_dict:
                raise ElasticityConfigError(f"Elasticity config missing {key}")
            setattr(self, attr, param_dict.get(key, default))
        self._validate()

    def _validate(self):
        mb = self.micro_batches
        if not isinstance(mb, list):
            raise ElasticityConfigError(f"Elasticity expected value of {MICRO_BATCHES} to be a list of micro batches, "
                                        f"instead is: {type(mb)}, containing: {mb}")
        if any(not isinstance(m, int) for m in mb):
            raise ElasticityConfigError(f"Elasticity expected {MICRO_BATCHES} to only contain a list of integers, "
                                        f"instead contains: f{mb}")
        if any(m <= 0 for m in mb):
            raise ElasticityConfigError(f"Elasticity expected {MICRO_BATCHES} to only contain positive integers, "
                                        f"instead contains: f{mb}")
        if min(self.min_gpus, self.max_gpus) < 1:
            raise ElasticityConfigError(f"Elasticity min/max gpus must be > 0, given min_gpus: {self.min_gpus}, "
                                        f"max_gpus: {self.max_gpus}")
        if self.max_gpus < self.min_gpus:
            raise ElasticityConfigError(f"Elasticity min_gpus cannot be greater than max_gpus, given min_gpus: "
                                        f"{self.min_gpus}, max_gpus: {self.max_gpus}")
        if self.min_time < 0:
            raise ElasticityConfigError(f"Elasticity min time needs to be >= 0: given {self.min_time}")

    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4)
