#!/usr/bin/env python
"""Offline recovery: stitch the fp32 master shards of a ZeRO checkpoint into one consolidated fp32 state_dict.

  python zero_to_fp32.py <checkpoint_dir> <output_file>

Handles stage 1 (`local_sub_partitions_of_fp32_groups`, bucket-major / rank-minor), stage 2
(`single_partition_of_fp32_groups`) and stage 3 (`fp32_flat_groups`, per-parameter shards) using the `param_shapes`
stored next to the optimizer state. Same role as reference deepspeed/utils/zero_to_fp32.py:21-151. Self-contained
on purpose: it is copied into every checkpoint directory."""
import argparse
import glob
import os
import pickle
import re
import types
from collections import OrderedDict

import torch


class _TolerantUnpickler(pickle.Unpickler):
    """The optimizer files pickle a few framework objects (loss scalers). This script only needs the tensors, so a
    class that cannot be imported (the framework is not installed where the checkpoint was moved to) becomes an empty
    stand-in instead of an ImportError."""

    def find_class(self, mod_name, name):
        try:
            return super().find_class(mod_name, name)
        except (ImportError, AttributeError):
            def _setstate(self, state):
                if isinstance(state, dict):
                    self.__dict__.update(state)
            return type(name, (), {"__module__": mod_name, "__setstate__": _setstate})


_tolerant_pickle = types.SimpleNamespace(Unpickler=_TolerantUnpickler, load=pickle.load, loads=pickle.loads,
                                         __name__="pickle")


def _load(path):
    return torch.load(path, map_location="cpu", weights_only=False, pickle_module=_tolerant_pickle)


def _rank_key(path):
    m = re.search(r"zero_pp_rank_(\d+)_mp_rank_(\d+)", path)
    return (int(m.group(2)), int(m.group(1))) if m else (0, 0)


def get_optim_files(checkpoint_dir):
    files = sorted(glob.glob(os.path.join(checkpoint_dir, "*_optim_states.pt")), key=_rank_key)
    if not files:
        raise FileNotFoundError(f"can't find '*_optim_states.pt' files in directory '{checkpoint_dir}'")
    return files


def _align8(n):
    return (n + 7) // 8 * 8


def _take_params(flat, param_shapes, state_dict, aligned):
    """Cut `flat` into the named parameters; `aligned` = arena layout with 8-element aligned parameter starts."""
    off = 0
    for name, shape in param_shapes.items():
        numel = 1
        for d in shape:
            numel *= d
        state_dict[name] = flat.narrow(0, off, numel).view(*shape).clone()
        off += numel
        if aligned:
            off = _align8(off)
    return off


_FP32_KEY = {1: "local_sub_partitions_of_fp32_groups", 2: "single_partition_of_fp32_groups", 3: "fp32_flat_groups"}


def parse_optim_states(files):
    """(zero stage, world size, param shapes, this checkpoint's fp32 groups per rank) from the per-rank optimizer
    files (reference utils/zero_to_fp32.py:36-60; stage 1 is understood here as well)."""
    sds = [_load(f) for f in files]
    first = sds[0]["optimizer_state_dict"]
    if "zero_stage" not in first:
        raise ValueError("non zero checkpoint")
    stage = first["zero_stage"]
    if stage not in _FP32_KEY:
        raise ValueError(f"unknown zero stage {stage}")
    groups = [sd["optimizer_state_dict"][_FP32_KEY[stage]] for sd in sds]
    return stage, first["partition_count"], sds[0]["param_shapes"], groups


def zero3_partitioned_param_info(unpartitioned_numel, world_size):
    """(elements of a parameter's shard when the parameter divides evenly, padding of the last shard otherwise)."""
    remainder = unpartitioned_numel % world_size
    return unpartitioned_numel // world_size, (world_size - remainder) if remainder else 0


def convert_zero_chkpt_to_fp32_consolid_state_dict(checkpoint_dir, output_file=None):
    files = get_optim_files(checkpoint_dir)
    sds = [_load(f) for f in files]
    osd = [sd["optimizer_state_dict"] for sd in sds]
    stage = osd[0]["zero_stage"]
    world = osd[0]["partition_count"]
    param_shapes = sds[0]["param_shapes"]
    state_dict = OrderedDict()
    total = sum(int(torch.tensor(list(s)).prod()) if len(s) else 1 for s in param_shapes.values())
    total_aligned = sum(_align8(int(torch.tensor(list(s)).prod()) if len(s) else 1) for s in param_shapes.values())
    if stage in (1, 2):
        key = "local_sub_partitions_of_fp32_groups" if stage == 1 else "single_partition_of_fp32_groups"
        ngroups = len(osd[0][key])
        flats = []
        for gi in range(ngroups):
            per_rank = [o[key][gi] for o in osd[:world]]
            per_rank = [[p] if torch.is_tensor(p) else list(p) for p in per_rank]
            pieces = []
            for k in range(len(per_rank[0])):
                for r in range(len(per_rank)):
                    pieces.append(per_rank[r][k].float().view(-1))
            flats.append(torch.cat(pieces))
        flat = torch.cat(flats) if len(flats) > 1 else flats[0]
        # one param group in model order is the common case; with several groups the shapes are consumed in order
        aligned = flat.numel() >= total_aligned and total_aligned != total
        _take_params(flat, param_shapes, state_dict, aligned)
    elif stage == 3:
        flat_groups = [torch.cat([g.float().view(-1) for g in o["fp32_flat_groups"]]) for o in osd[:world]]
        offset = 0
        for name, shape in param_shapes.items():
            numel = 1
            for d in shape:
                numel *= d
            part = (numel + world - 1) // world
            full = torch.cat([fg.narrow(0, offset, part) for fg in flat_groups])
            state_dict[name] = full.narrow(0, 0, numel).view(*shape).clone()
            offset += part
    else:
        raise ValueError(f"unknown zero stage {stage}")
    if output_file is not None:
        print(f"Saving fp32 state dict to {output_file} (total_numel = {total})")
        torch.save(state_dict, output_file)
    return state_dict


if __name__ == "__main__":
    parser = argparse.ArgumentParser()
    parser.add_argument("checkpoint_dir", type=str, help="path to the deepspeed checkpoint folder, e.g. path/global_step1")
    parser.add_argument("output_file", type=str, help="path to the pytorch fp32 state_dict output file")
    args = parser.parse_args()
    convert_zero_chkpt_to_fp32_consolid_state_dict(args.checkpoint_dir, args.output_file)
