"""Elastic batch-size solver v0.1 (reference deepspeed/elasticity/elasticity.py:19-334): pick the train batch size
<= max_train_batch_size that is compatible with the largest number of GPU counts."""
import json
import os
import re
from functools import reduce
from math import gcd

from .config import ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize
from .constants import (ELASTICITY, ENABLED, ENABLED_DEFAULT, LATEST_ELASTICITY_VERSION, MINIMUM_DEEPSPEED_VERSION,
                        DEEPSPEED_ELASTICITY_CONFIG)
from ..utils.logging import logger

# Highly composite numbers: good batch-size multipliers because they have many divisors (many valid GPU counts).
HCN_LIST = [1, 2, 4, 6, 12, 24, 36, 48, 60, 120, 180, 240, 360, 720, 840, 1260, 1680, 2520, 5040, 7560, 10080, 15120,
            20160, 25200, 27720, 45360, 50400, 55440, 83160, 110880, 166320, 221760, 277200, 332640, 498960, 554400,
            665280, 720720]


def _lcm(a, b):
    return a * b // gcd(a, b)


def get_candidate_batch_sizes(base_list, max_acceptable_batch_size):
    candidates = set()
    for base in base_list:
        if base >= max_acceptable_batch_size:
            candidates.add(base)
        else:
            value = max_acceptable_batch_size // base
            index = next((i for i, h in enumerate(HCN_LIST) if h > value), len(HCN_LIST))
            candidates.add(HCN_LIST[index - 1] * base)
    return sorted(candidates)


def get_valid_gpus(batch_size, micro_batches, min_valid_gpus, max_valid_gpus):
    valid = set()
    for micro_batch in micro_batches:
        if batch_size % micro_batch == 0:
            max_gpus = batch_size // micro_batch
            if min_valid_gpus <= max_gpus <= max_valid_gpus:
                valid.add(max_gpus)
            for i in range(1, max_gpus // 2 + 1):
                if max_gpus % i == 0 and min_valid_gpus <= i <= max_valid_gpus:
                    valid.add(i)
    return sorted(valid)


def get_best_candidates(candidate_batch_sizes, micro_batches, min_gpus, max_gpus, prefer_larger):
    max_valid_gpus, valid_gpus, final_batch_size = 0, None, int(min(micro_batches))
    for batch_size in candidate_batch_sizes:
        current = get_valid_gpus(batch_size, micro_batches, min_gpus, max_gpus)
        better = len(current) > max_valid_gpus or (len(current) == max_valid_gpus and (
            (prefer_larger and batch_size > final_batch_size) or (not prefer_larger and batch_size < final_batch_size)))
        if better:
            max_valid_gpus, valid_gpus, final_batch_size = len(current), current, batch_size
    return final_batch_size, valid_gpus


def _get_compatible_gpus_v01(micro_batches, max_acceptable_batch_size, min_gpus=None, max_gpus=None, prefer_larger=True):
    min_gpus = min_gpus or 1
    max_gpus = max_gpus or max_acceptable_batch_size // min(micro_batches)
    if not all(mb <= max_acceptable_batch_size for mb in micro_batches):
        raise ValueError(f"All micro batches must be less than or equal to max_acceptable_batch_size: "
                         f"{max_acceptable_batch_size}")
    lcm = reduce(_lcm, micro_batches)
    base_list = list(micro_batches) + [lcm]
    candidates = get_candidate_batch_sizes(base_list, max_acceptable_batch_size)
    return get_best_candidates(candidates, micro_batches, min_gpus, max_gpus, prefer_larger)


def _parse_version(version_str):
    m = re.search(r"^(\d+)\.(\d+)\.(\d+)", version_str)
    if m:
        return int(m.group(1)), int(m.group(2)), int(m.group(3))
    m = re.search(r"^(\d+)\.(\d+)", version_str)
    assert m is not None, f"Unable to parse version number, expecting major.minor[.patch] format but received {version_str}"
    return int(m.group(1)), int(m.group(2)), 0


def _compatible_ds_version_check(target_deepspeed_version):
    min_version = _parse_version(MINIMUM_DEEPSPEED_VERSION)
    trg_version = _parse_version(target_deepspeed_version)
    err = (f"Target deepspeed version of {target_deepspeed_version} is not compatible with minimum version "
           f"{MINIMUM_DEEPSPEED_VERSION} supporting elasticity.")
    if trg_version < min_version:
        raise ElasticityError(err)
    return True


def elasticity_enabled(ds_config):
    if ELASTICITY not in ds_config:
        return False
    return ds_config[ELASTICITY].get(ENABLED, ENABLED_DEFAULT)


def ensure_immutable_elastic_config(runtime_elastic_config_dict):
    """The scheduler passes its view of the elastic config via env; the runtime view must match it."""
    if DEEPSPEED_ELASTICITY_CONFIG in os.environ:
        scheduler = ElasticityConfig(json.loads(os.environ[DEEPSPEED_ELASTICITY_CONFIG]))
        runtime = ElasticityConfig(runtime_elastic_config_dict)
        for attr, label in (("max_acceptable_batch_size", "max_acceptable_batch_size"), ("micro_batches", "micro_batches"),
                            ("version", "version")):
            if getattr(runtime, attr) != getattr(scheduler, attr):
                raise ElasticityConfigError(f"Elastic config '{label}={getattr(scheduler, attr)}' seen by resource "
                                            f"scheduler does not match config passed to runtime "
                                            f"{label}={getattr(runtime, attr)}")
    else:
        logger.warning("Unable to find DEEPSPEED_ELASTICITY_CONFIG environment variable, cannot guarantee resource "
                       "scheduler will scale this job using compatible GPU counts.")


def compute_elastic_config(ds_config, target_deepspeed_version, world_size=0):
    """Returns (final_batch_size, valid_gpus[, micro_batch_size if world_size > 0])."""
    if not isinstance(ds_config, dict):
        raise ValueError(f"Expected ds_config to be a dictionary but received a {type(ds_config)}, containing: {ds_config}")
    if ELASTICITY not in ds_config:
        raise ElasticityConfigError(f"'{ELASTICITY}' is missing from config json, please add it if running an "
                                    "elastic training job.")
    elastic_config_dict = ds_config[ELASTICITY]
    if not elastic_config_dict.get(ENABLED, ENABLED_DEFAULT):
        raise ElasticityConfigError("Elasticity is disabled, please enable it ('enabled':true) if running an elastic "
                                    "training job.")
    elastic_config = ElasticityConfig(elastic_config_dict)
    if float(elastic_config.version) > LATEST_ELASTICITY_VERSION:
        raise ElasticityConfigError(f"Attempting to run elasticity version {elastic_config.version} but runtime only "
                                    f"supports up to {LATEST_ELASTICITY_VERSION}")
    if not _compatible_ds_version_check(target_deepspeed_version):
        raise ElasticityError("Unable to run elasticity on target deepspeed version")
    if float(elastic_config.version) == 0.1:
        final_batch_size, valid_gpus = _get_compatible_gpus_v01(
            micro_batches=elastic_config.micro_batches, max_acceptable_batch_size=elastic_config.max_acceptable_batch_size,
            min_gpus=elastic_config.min_gpus, max_gpus=elastic_config.max_gpus,
            prefer_larger=elastic_config.prefer_larger_batch_size)
        final_batch_size = int(final_batch_size)
    else:
        raise NotImplementedError(f"Unable to find elastic logic for version: {elastic_config.version}")
    if world_size > 0:
        if world_size not in valid_gpus:
            raise ElasticityIncompatibleWorldSize(f"World size ({world_size}) is not valid with the current list of "
                                                  f"valid GPU counts: {valid_gpus}")
        micro_batch_size = None
        for mbsz in sorted(set(elastic_config.micro_batches), reverse=True):
            if final_batch_size // world_size % mbsz == 0:
                micro_batch_size = mbsz
                break
        assert micro_batch_size is not None, (f"Unable to find divisible micro batch size world_size={world_size}, "
                                              f"final_batch_size={final_batch_size}, micro_batches={elastic_config.micro_batches}.")
        return final_batch_size, valid_gpus, micro_batch_size
    return final_batch_size, valid_gpus
