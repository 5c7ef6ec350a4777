"""PipelineModule / LayerSpec / TiedLayerSpec.

Public behaviour per reference deepspeed/runtime/pipe/module.py:23-589: a model is a flat list of layer specs; each
rank builds only the layers of its stage; partitioning by `uniform`, `parameters` or `type:<regex>`; tied layers are
grouped across stages with an initial broadcast and a gradient all-reduce; one checkpoint file per layer."""
import glob
import os
import re as regex
from collections import defaultdict
from functools import partial

import torch
import torch.distributed as dist
import torch.nn as nn

from ...utils.logging import logger
from .. import utils as ds_utils
from ..activation_checkpointing import checkpointing
from .topology import PipeDataParallelTopology, PipelineParallelGrid


class PipelineError(Exception):
    """Errors related to the use of deepspeed.PipelineModule"""


class LayerSpec:
    """Deferred construction of a layer: stores the class and its arguments so that only the owning stage
    ever allocates the parameters (a 1.3B..100B+ model never materialises on one device)."""

    def __init__(self, typename, *module_args, **module_kwargs):
        self.typename = typename
        self.module_args = module_args
        self.module_kwargs = module_kwargs
        if not issubclass(typename, nn.Module):
            raise RuntimeError("LayerSpec only supports torch.nn.Module types.")
        self.global_rank = dist.get_rank() if dist.is_initialized() else -1

    def __repr__(self):
        return ds_utils.call_to_str(self.typename.__name__, *self.module_args, **self.module_kwargs)

    def build(self, log=False):
        if log:
            logger.info(f"RANK={self.global_rank} building {repr(self)}")
        return self.typename(*self.module_args, **self.module_kwargs)


class TiedLayerSpec(LayerSpec):
    def __init__(self, key, typename, *module_args, forward_fn=None, tied_weight_attr="weight", **module_kwargs):
        super().__init__(typename, *module_args, **module_kwargs)
        self.key = key
        self.forward_fn = forward_fn
        self.tied_weight_attr = tied_weight_attr


class PipelineModule(nn.Module):
    def __init__(self, layers, num_stages=None, topology=None, loss_fn=None, seed_layers=False, seed_fn=None,
                 base_seed=1234, partition_method="parameters", activation_checkpoint_interval=0,
                 activation_checkpoint_func=checkpointing.checkpoint, checkpointable_layers=None):
        super().__init__()
        if num_stages is None and topology is None:
            raise RuntimeError("must provide num_stages or topology")
        self.micro_offset = 0
        self.loss_fn = loss_fn
        self.checkpointable_layers = checkpointable_layers
        if checkpointable_layers is not None:
            assert isinstance(checkpointable_layers, list), "param `checkpointable_layers` must be type of list."
        self.seed_layers = seed_layers
        self.seed_fn = seed_fn
        self.base_seed = base_seed
        if dist.is_initialized() and dist.get_rank() == 0:
            try:
                logger.info(f"SEED_LAYERS={self.seed_layers} BASE_SEED={self.base_seed} SEED_FN={self.seed_fn}")
            except Exception:
                logger.info(f"SEED_LAYERS={self.seed_layers} BASE_SEED={self.base_seed}")

        if not dist.is_initialized():
            from ...utils.distributed import init_distributed
            init_distributed(verbose=False)
        self.world_group = dist.new_group(ranks=range(dist.get_world_size()))
        self.global_rank = dist.get_rank(group=self.world_group)
        self.world_size = dist.get_world_size(group=self.world_group)
        self.local_rank = int(os.environ.get("LOCAL_RANK", 0))

        if topology:
            self._topo = topology
            self.num_stages = self._topo.get_dim("pipe")
        else:
            self.num_stages = num_stages
            if self.world_size % self.num_stages != 0:
                raise RuntimeError(f"num_stages ({self.num_stages}) must divide distributed world size ({self.world_size})")
            dp = self.world_size // num_stages
            topology = PipeDataParallelTopology(num_pp=num_stages, num_dp=dp)
            self._topo = topology

        self._grid = PipelineParallelGrid(process_group=self.world_group, topology=self._topo)
        self.stage_id = self._topo.get_coord(self.global_rank).pipe

        self._layer_specs = list(layers)
        self._num_layers = len(self._layer_specs)
        self._local_start = 0
        self._local_stop = None
        self._partition_layers(method=partition_method)

        self.forward_funcs = []
        self.tied_modules = nn.ModuleDict()
        self.tied_weight_attrs = {}
        self._build()
        if torch.cuda.is_available():
            self.to(f"cuda:{self.local_rank}")

        self.tied_comms = self._index_tied_modules()
        self._synchronize_tied_weights()

        self.activation_checkpoint_interval = activation_checkpoint_interval
        self.activation_checkpoint_func = activation_checkpoint_func

    # ------------------------------------------------------------------ construction
    def _build(self):
        specs = self._layer_specs
        for local_idx, layer in enumerate(specs[self._local_start:self._local_stop]):
            layer_idx = local_idx + self._local_start
            if self.seed_layers:
                if self.seed_fn:
                    self.seed_fn(self.base_seed + layer_idx)
                else:
                    ds_utils.set_random_seed(self.base_seed + layer_idx)
            if isinstance(layer, PipelineModule):
                raise NotImplementedError("RECURSIVE BUILD NOT YET IMPLEMENTED")
            elif isinstance(layer, nn.Module):
                name = str(layer_idx)
                self.forward_funcs.append(layer)
                self.add_module(name, layer)
            elif isinstance(layer, TiedLayerSpec):
                if layer.key not in self.tied_modules:
                    self.tied_modules[layer.key] = layer.build()
                    self.tied_weight_attrs[layer.key] = layer.tied_weight_attr
                if layer.forward_fn is None:
                    self.forward_funcs.append(self.tied_modules[layer.key])
                else:
                    self.forward_funcs.append(partial(layer.forward_fn, self.tied_modules[layer.key]))
            elif isinstance(layer, LayerSpec):
                module = layer.build()
                name = str(layer_idx)
                self.forward_funcs.append(module)
                self.add_module(name, module)
            else:
                self.forward_funcs.append(layer)
        # everything in a pipeline is "model parallel": unique to this stage, not replicated across it
        for p in self.parameters():
            p.model_parallel = True

    def _count_layer_params(self):
        param_counts = [0] * len(self._layer_specs)
        for idx, layer in enumerate(self._layer_specs):
            if isinstance(layer, LayerSpec):
                l = layer.build()
                params = filter(lambda p: p.requires_grad, l.parameters())
                param_counts[idx] = sum(p.numel() for p in params)
            elif isinstance(layer, nn.Module):
                params = filter(lambda p: p.requires_grad, layer.parameters())
                param_counts[idx] = sum(p.numel() for p in params)
        return param_counts

    def _find_layer_type(self, layername):
        idxs = []
        typeregex = regex.compile(layername, regex.IGNORECASE)
        for idx, layer in enumerate(self._layer_specs):
            name = None
            if isinstance(layer, LayerSpec):
                name = layer.typename.__name__
            elif isinstance(layer, nn.Module):
                name = layer.__class__.__name__
            else:
                try:
                    name = layer.__name__
                except AttributeError:
                    continue
            if typeregex.search(name):
                idxs.append(idx)
        if len(idxs) == 0:
            raise RuntimeError(f"Partitioning '{layername}' found no valid layers to partition.")
        return idxs

    def _partition_layers(self, method="uniform"):
        num_stages = self._topo.get_dim("pipe")
        stage_id = self._topo.get_coord(self.global_rank).pipe
        if self.global_rank == 0:
            logger.info(f"Partitioning pipeline stages with method {method}")
        method = method.lower()
        if method == "uniform":
            self.parts = ds_utils.partition_uniform(num_items=len(self._layer_specs), num_parts=num_stages)
        elif method == "parameters":
            # meta-device construction keeps counting cheap for huge specs
            try:
                with torch.device("meta"):
                    param_counts = self._count_layer_params()
            except Exception:
                param_counts = self._count_layer_params()
            self.parts = ds_utils.partition_balanced(weights=param_counts, num_parts=num_stages)
        elif method.startswith("type:"):
            layertype = method.split(":")[1]
            binary_weights = [0] * len(self._layer_specs)
            for idx in self._find_layer_type(layertype):
                binary_weights[idx] = 1
            self.parts = ds_utils.partition_balanced(weights=binary_weights, num_parts=num_stages)
        elif method == "profile":
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        else:
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        if self.global_rank == 0:
            for stage in range(num_stages):
                start, stop = self.parts[stage], self.parts[stage + 1]
                print(f"stage={stage} layers={stop - start}")
                for idx, layer in enumerate(self._layer_specs[start:stop]):
                    name = str(layer)
                    if isinstance(layer, LayerSpec):
                        name = layer.typename.__name__
                    if isinstance(layer, nn.Module):
                        name = layer.__class__.__name__
                    else:
                        try:
                            name = layer.__name__
                        except AttributeError:
                            pass
                    print(f"    {idx + start:2d}: {name}")
            if self.loss_fn:
                try:
                    print(f"  loss: {self.loss_fn.__name__}")
                except AttributeError:
                    print(f"  loss: {self.loss_fn.__class__.__name__}")
        self._set_bounds(start=self.parts[stage_id], stop=self.parts[stage_id + 1])

    def _set_bounds(self, start=None, stop=None):
        self._local_start = start
        self._local_stop = stop

    # ------------------------------------------------------------------ forward
    def set_checkpoint_interval(self, interval):
        assert interval >= 0
        self.activation_checkpoint_interval = interval

    def forward(self, forward_input):
        # micro_offset keeps layer seeds distinct across micro-batches (dropout) when seed_layers is on
        self.micro_offset += 1

        def exec_range_func(start, end):
            local_micro_offset = self.micro_offset + 1

            def exec_func(*inputs):
                if len(inputs) == 1:
                    inputs = inputs[0]
                for idx, layer in enumerate(self.forward_funcs[start:end]):
                    self.curr_layer = idx + self._local_start
                    if self.seed_layers:
                        new_seed = (self.base_seed * local_micro_offset) + self.curr_layer
                        if self.seed_fn:
                            self.seed_fn(new_seed)
                        else:
                            ds_utils.set_random_seed(new_seed)
                    inputs = layer(inputs)
                return inputs

            return exec_func

        if self.activation_checkpoint_interval == 0:
            func = exec_range_func(0, len(self.forward_funcs))
            x = func(forward_input)
        else:
            num_layers = len(self.forward_funcs)
            x = forward_input
            for start_idx in range(0, num_layers, self.activation_checkpoint_interval):
                end_idx = min(start_idx + self.activation_checkpoint_interval, num_layers)
                funcs = self.forward_funcs[start_idx:end_idx]
                if not isinstance(x, tuple):
                    x = (x,)
                # a chunk fed only by tensors that carry no gradient (token ids into an embedding) cannot be
                # re-entered by autograd through a checkpoint: its parameters would silently get no gradient (the
                # reference raises in backward here), so it runs un-checkpointed
                differentiable_input = any(torch.is_tensor(t) and t.requires_grad for t in x)
                if self._is_checkpointable(funcs) and differentiable_input:
                    x = self.activation_checkpoint_func(exec_range_func(start_idx, end_idx), *x)
                else:
                    x = exec_range_func(start_idx, end_idx)(*x)
        return x

    def _is_checkpointable(self, funcs):
        if self.__class__.__name__ == "GPT2ModelPipe":
            return all("ParallelTransformerLayerPipe" in f.__class__.__name__ for f in funcs)
        if self.checkpointable_layers is not None:
            return all(f.__class__.__name__ in self.checkpointable_layers for f in funcs)
        params = [f.parameters() for f in funcs if isinstance(f, torch.nn.Module)]
        return any(len(list(p)) > 0 for p in params)

    # ------------------------------------------------------------------ tied weights
    def allreduce_tied_weight_gradients(self):
        for key, comm in self.tied_comms.items():
            weight = getattr(self.tied_modules[key], comm["weight_attr"])
            if weight.grad is not None:
                dist.all_reduce(weight.grad, group=comm["group"])

    def _synchronize_tied_weights(self):
        for key, comm in self.tied_comms.items():
            dist.broadcast(getattr(comm["module"], comm["weight_attr"]), src=min(comm["ranks"]), group=comm["group"])

    def _index_tied_modules(self):
        """Build the communication structures for tied modules (collective: all ranks create all groups)."""
        tied_comms = {}
        if self._topo.get_dim("pipe") == 1:
            return tied_comms
        specs = self._layer_specs
        tie_keys = set(s.key for s in specs if isinstance(s, TiedLayerSpec))
        for key in sorted(tie_keys):
            tied_layers = [idx for idx, layer in enumerate(specs) if isinstance(layer, TiedLayerSpec) and layer.key == key]
            tied_stages = set(self.stage_owner(idx) for idx in tied_layers)
            for dp in range(self._grid.data_parallel_size):
                for mp in range(self._grid.get_slice_parallel_world_size()):
                    tied_ranks = []
                    for s in sorted(tied_stages):
                        if self._grid.get_slice_parallel_world_size() > 1:
                            tied_ranks.append(self._grid.stage_to_global(stage_id=s, data=dp, model=mp))
                        else:
                            tied_ranks.append(self._grid.stage_to_global(stage_id=s, data=dp))
                    group = dist.new_group(ranks=tied_ranks)
                    if self.global_rank in tied_ranks:
                        assert key in self.tied_modules
                        tied_comms[key] = {"ranks": tied_ranks, "group": group,
                                           "weight_attr": self.tied_weight_attrs[key],
                                           "module": self.tied_modules[key]}
                        # only the lowest rank of the group "owns" the parameter for norm purposes
                        if self.global_rank != tied_ranks[0]:
                            for p in self.tied_modules[key].parameters():
                                p.model_parallel = False
        return tied_comms

    # ------------------------------------------------------------------ queries
    def partitions(self):
        return self.parts

    def stage_owner(self, layer_idx):
        assert 0 <= layer_idx < self._num_layers
        for stage in range(self._topo.get_dim("pipe")):
            if self.parts[stage] <= layer_idx < self.parts[stage + 1]:
                return stage
        raise RuntimeError(f"Layer {layer_idx} not owned? parts={self.parts}")

    def topology(self):
        return self._topo

    def mpu(self):
        return self._grid

    def num_pipeline_stages(self):
        return self._topo.get_dim("pipe")

    # ------------------------------------------------------------------ checkpoints (one file per layer)
    def ckpt_prefix(self, checkpoints_path, tag):
        rank_name = "module"
        omit_dims = frozenset(["data"])
        axes = [a for a in self._grid._topo.get_axis_names() if a not in omit_dims]
        for dim in axes:
            rank = getattr(self._grid._topo.get_coord(rank=self.global_rank), dim)
            rank_name += f"-{dim}_{rank:02d}"
        return os.path.join(checkpoints_path, str(tag), rank_name)

    def ckpt_layer_path(self, ckpt_dir, local_layer_idx):
        idx = local_layer_idx + self._local_start
        layer_ckpt_path = os.path.join(ckpt_dir, f"layer_{idx:02d}")
        rank_repr = self._grid._topo.get_rank_repr(rank=self.global_rank)
        if rank_repr != "":
            layer_ckpt_path += f"-{rank_repr}"
        layer_ckpt_path += "-model_states.pt"
        return layer_ckpt_path

    def ckpt_layer_path_list(self, ckpt_dir, local_layer_idx):
        idx = local_layer_idx + self._local_start
        layer_ckpt_path = os.path.join(ckpt_dir, f"layer_{idx:02d}-")
        layer_ckpt_path += "*model_states.pt"
        ckpt_files = glob.glob(layer_ckpt_path)
        ckpt_files.sort()
        return ckpt_files

    def save_state_dict(self, save_dir):
        if self._grid.data_parallel_id != 0:
            return
        os.makedirs(save_dir, exist_ok=True)
        for idx, layer in enumerate(self.forward_funcs):
            model_ckpt_path = self.ckpt_layer_path(save_dir, idx)
            if not hasattr(layer, "state_dict"):
                continue
            # move to CPU so loading never needs the saving device
            orig = layer.state_dict()
            final = type(orig)({k: v.clone().cpu() if torch.is_tensor(v) else v for k, v in orig.items()})
            torch.save(final, model_ckpt_path)

    def load_state_dir(self, load_dir, strict=True):
        for idx, layer in enumerate(self.forward_funcs):
            if not hasattr(layer, "load_state_dict"):
                continue
            model_ckpt_path = self.ckpt_layer_path(load_dir, idx)
            layer.load_state_dict(torch.load(model_ckpt_path, map_location=lambda storage, loc: storage,
                                             weights_only=False), strict=strict)
        self._synchronize_tied_weights()
