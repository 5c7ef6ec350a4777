"""PipelineModule / LayerSpec / TiedLayerSpec.

Contract of reference deepspeed/runtime/pipe/module.py:23-589: a model is a flat list of layer specs; a rank builds only
the layers of its stage; stages are cut by `uniform`, `parameters` or `type:<regex>`; tied layers are grouped across
stages (initial broadcast, gradient all-reduce); checkpoints are one file per layer.

Own additions: `partition_method="compute"` balances an estimate of the forward + backward GEMM work instead of raw
parameter counts (an embedding lookup costs nothing, an output projection costs as much as its weight says), layer
costs are measured on the meta device so that cutting a 100B-parameter spec list allocates nothing, and the forward
pass runs pre-built segments instead of re-deriving the checkpoint chunks on every call.
"""
import glob
import os
import re as regex
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
    """A layer that is described now and built later, by the stage that owns it: the class and its arguments are
    stored, so a model far larger than one device's memory can be listed on every rank."""

    def __init__(self, typename, *module_args, **module_kwargs):
        if not issubclass(typename, nn.Module):
            raise RuntimeError("LayerSpec only supports torch.nn.Module types.")
        self.typename = typename
        self.module_args = module_args
        self.module_kwargs = module_kwargs
        self.global_rank = dist.get_rank() if dist.is_initialized() else -1

    def __repr__(self):
        return ds_utils.call_to_str(self.typename.__name__, *self.module_args, **self.module_kwargs)

    def build(self, log=False):
        if log:
            logger.info(f"RANK={self.global_rank} building {repr(self)}")
        return self.typename(*self.module_args, **self.module_kwargs)


class TiedLayerSpec(LayerSpec):
    """A LayerSpec whose module is shared by every spec with the same `key` (e.g. input embedding and output
    projection); `forward_fn(module, x)` lets the reuses compute something different with the shared weights."""

    def __init__(self, key, typename, *module_args, forward_fn=None, tied_weight_attr="weight", **module_kwargs):
        super().__init__(typename, *module_args, **module_kwargs)
        self.key = key
        self.forward_fn = forward_fn
        self.tied_weight_attr = tied_weight_attr


def _spec_name(spec):
    """Class / function name of a spec, or None if it has none (used by `type:` partitioning and the stage report)."""
    if isinstance(spec, LayerSpec):
        return spec.typename.__name__
    if isinstance(spec, nn.Module):
        return spec.__class__.__name__
    return getattr(spec, "__name__", None)


def _trainable_numel(module):
    return sum(p.numel() for p in module.parameters() if p.requires_grad)


def _spec_weights(specs, compute):
    """One weight per spec: trainable parameter count, or (compute=True) parameters that take part in GEMMs. A class
    may override the estimate with an attribute `pipe_compute_weight` (number, or callable(module) -> number);
    embedding tables count as zero because a lookup moves rows instead of multiplying them."""
    def weigh(module, cls):
        if compute:
            override = getattr(cls, "pipe_compute_weight", None)
            if override is not None:
                return float(override(module) if callable(override) else override)
            lookup = sum(_trainable_numel(m) for m in module.modules() if isinstance(m, (nn.Embedding, nn.EmbeddingBag)))
            return _trainable_numel(module) - lookup
        return _trainable_numel(module)

    def measure():
        out = []
        for spec in specs:
            if isinstance(spec, LayerSpec):
                out.append(weigh(spec.build(), spec.typename))
            elif isinstance(spec, nn.Module):
                out.append(weigh(spec, spec.__class__))
            else:
                out.append(0)
        return out

    try:
        with torch.device("meta"):      # shapes only: nothing is allocated or initialised
            return measure()
    except Exception:
        return measure()


class _Segment:
    """A run of consecutive local layers executed as one unit (the granule of activation checkpointing)."""

    def __init__(self, owner, start, stop):
        self.owner, self.start, self.stop = owner, start, stop

    def __call__(self, *inputs):
        o = self.owner
        x = inputs[0] if len(inputs) == 1 else inputs
        micro = o.micro_offset + 1
        for i in range(self.start, self.stop):
            o.curr_layer = i + o._local_start
            if o.seed_layers:
                # a different seed per layer and micro-batch (dropout) that a recompute reproduces
                seed = o.base_seed * micro + o.curr_layer
                (o.seed_fn or ds_utils.set_random_seed)(seed)
            x = o.forward_funcs[i](x)
        return x


class PipelineModule(nn.Module):
    """Sequential model split over pipeline stages. `layers` is a list of modules, LayerSpecs or plain callables; only
    the layers of this rank's stage are built."""

    def __init__(self, layers, num_stages=None, topology=None, loss_fn=None, seed_layers=False, seed_fn=None,
                 base_seed=1234, partition_method="parameters", activation_checkpoint_interval=0,
                 activation_checkpoint_func=checkpointing.checkpoint, checkpointable_layers=None):
        super().__init__()
        if num_stages is None and topology is None:
            raise RuntimeError("must provide num_stages or topology")
        if checkpointable_layers is not None:
            assert isinstance(checkpointable_layers, list), "param `checkpointable_layers` must be type of list."
        self.micro_offset = 0
        self.loss_fn = loss_fn
        self.checkpointable_layers = checkpointable_layers
        self.seed_layers, self.seed_fn, self.base_seed = seed_layers, seed_fn, base_seed

        if not dist.is_initialized():
            from ...utils.distributed import init_distributed
            init_distributed(verbose=False)
        self.world_group = dist.new_group(ranks=range(dist.get_world_size()))
        self.global_rank = dist.get_rank(group=self.world_group)
        self.world_size = dist.get_world_size(group=self.world_group)
        self.local_rank = int(os.environ.get("LOCAL_RANK", 0))
        if self.global_rank == 0:
            logger.info(f"SEED_LAYERS={seed_layers} BASE_SEED={base_seed} SEED_FN={getattr(seed_fn, '__name__', seed_fn)}")

        if topology is None:
            if self.world_size % num_stages != 0:
                raise RuntimeError(f"num_stages ({num_stages}) must divide distributed world size ({self.world_size})")
            topology = PipeDataParallelTopology(num_pp=num_stages, num_dp=self.world_size // num_stages)
        self._topo = topology
        self.num_stages = topology.get_dim("pipe")
        self._grid = PipelineParallelGrid(process_group=self.world_group, topology=topology)
        self.stage_id = topology.get_coord(self.global_rank).pipe

        self._layer_specs = list(layers)
        self._num_layers = len(self._layer_specs)
        self._local_start, self._local_stop = 0, None
        self._partition_layers(method=partition_method)

        self.forward_funcs = []
        self.tied_modules = nn.ModuleDict()
        self.tied_weight_attrs = {}
        self._build()
        if torch.cuda.is_available():
            self.to(f"cuda:{self.local_rank}")

        self.tied_comms = self._index_tied_modules()
        self._synchronize_tied_weights()

        self._segments_for = None
        self.activation_checkpoint_interval = activation_checkpoint_interval
        self.activation_checkpoint_func = activation_checkpoint_func

    # ------------------------------------------------------------------ construction
    def _build(self):
        for idx in range(self._local_start, self._local_stop):
            spec = self._layer_specs[idx]
            if self.seed_layers:
                (self.seed_fn or ds_utils.set_random_seed)(self.base_seed + idx)
            if isinstance(spec, PipelineModule):
                raise NotImplementedError("RECURSIVE BUILD NOT YET IMPLEMENTED")
            if isinstance(spec, TiedLayerSpec):
                if spec.key not in self.tied_modules:
                    self.tied_modules[spec.key] = spec.build()
                    self.tied_weight_attrs[spec.key] = spec.tied_weight_attr
                shared = self.tied_modules[spec.key]
                self.forward_funcs.append(shared if spec.forward_fn is None else partial(spec.forward_fn, shared))
            elif isinstance(spec, (LayerSpec, nn.Module)):
                module = spec.build() if isinstance(spec, LayerSpec) else spec
                self.add_module(str(idx), module)
                self.forward_funcs.append(module)
            else:
                self.forward_funcs.append(spec)      # a plain callable
        # nothing in a pipeline stage is replicated over the "model" group: norms and overflow checks must count it
        for p in self.parameters():
            p.model_parallel = True

    def _count_layer_params(self):
        return _spec_weights(self._layer_specs, compute=False)

    def _find_layer_type(self, layername):
        pattern = regex.compile(layername, regex.IGNORECASE)
        hits = [i for i, spec in enumerate(self._layer_specs)
                if _spec_name(spec) is not None and pattern.search(_spec_name(spec))]
        if not hits:
            raise RuntimeError(f"Partitioning '{layername}' found no valid layers to partition.")
        return hits

    def _partition_layers(self, method="uniform"):
        stages = self._topo.get_dim("pipe")
        if self.global_rank == 0:
            logger.info(f"Partitioning pipeline stages with method {method}")
        method = method.lower()
        n = len(self._layer_specs)
        if method == "uniform":
            self.parts = ds_utils.partition_uniform(num_items=n, num_parts=stages)
        elif method in ("parameters", "compute"):
            weights = _spec_weights(self._layer_specs, compute=(method == "compute"))
            self.parts = ds_utils.partition_balanced(weights=weights, num_parts=stages)
        elif method.startswith("type:"):
            marked = set(self._find_layer_type(method.split(":", 1)[1]))
            self.parts = ds_utils.partition_balanced(weights=[int(i in marked) for i in range(n)], num_parts=stages)
        else:
            raise NotImplementedError(f"Partitioning method {method} not implemented.")
        if self.global_rank == 0:
            self._print_partition(stages)
        self._set_bounds(start=self.parts[self.stage_id], stop=self.parts[self.stage_id + 1])

    def _print_partition(self, stages):
        for stage in range(stages):
            lo, hi = self.parts[stage], self.parts[stage + 1]
            print(f"stage={stage} layers={hi - lo}")
            for i in range(lo, hi):
                print(f"    {i:2d}: {_spec_name(self._layer_specs[i]) or self._layer_specs[i]}")
        if self.loss_fn:
            print(f"  loss: {getattr(self.loss_fn, '__name__', self.loss_fn.__class__.__name__)}")

    def _set_bounds(self, start=None, stop=None):
        self._local_start = start
        self._local_stop = stop

    # ------------------------------------------------------------------ forward
    def set_checkpoint_interval(self, interval):
        assert interval >= 0
        self.activation_checkpoint_interval = interval

    def _segments(self):
        """The local layers cut into runs of `activation_checkpoint_interval` (one run if checkpointing is off),
        rebuilt only when the interval or the layer list changes."""
        key = (self.activation_checkpoint_interval, len(self.forward_funcs))
        if self._segments_for != key:
            n, step = len(self.forward_funcs), self.activation_checkpoint_interval or max(len(self.forward_funcs), 1)
            self._segment_list = [_Segment(self, lo, min(lo + step, n)) for lo in range(0, max(n, 1), step)]
            self._segment_ckpt = [self.activation_checkpoint_interval > 0 and
                                  self._is_checkpointable(self.forward_funcs[s.start:s.stop]) for s in self._segment_list]
            self._segments_for = key
        return zip(self._segment_list, self._segment_ckpt)

    def forward(self, forward_input):
        self.micro_offset += 1      # keeps layer seeds distinct across micro-batches when seed_layers is on
        x = forward_input
        for segment, may_checkpoint in self._segments():
            if not may_checkpoint:
                x = segment(*x) if isinstance(x, tuple) else segment(x)
                continue
            args = x if isinstance(x, tuple) else (x,)
            # a run fed only by tensors without gradient (token ids into an embedding) cannot be re-entered by autograd
            # through a checkpoint - its parameters would silently get no gradient - so it runs un-checkpointed
            if any(torch.is_tensor(t) and t.requires_grad for t in args):
                x = self.activation_checkpoint_func(segment, *args)
            else:
                x = segment(*args)
        return x

    def _is_checkpointable(self, funcs):
        if self.__class__.__name__ == "GPT2ModelPipe":
            return all("ParallelTransformerLayerPipe" in f.__class__.__name__ for f in funcs)
        if self.checkpointable_layers is not None:
            return all(f.__class__.__name__ in self.checkpointable_layers for f in funcs)
        return any(isinstance(f, nn.Module) and next(f.parameters(), None) is not None for f in funcs)

    # ------------------------------------------------------------------ tied weights
    def allreduce_tied_weight_gradients(self):
        """Sum the gradients of every tied weight over the stages that hold a copy."""
        for key, comm in self.tied_comms.items():
            grad = getattr(self.tied_modules[key], comm["weight_attr"]).grad
            if grad is not None:
                dist.all_reduce(grad, group=comm["group"])

    def _synchronize_tied_weights(self):
        for comm in self.tied_comms.values():
            dist.broadcast(getattr(comm["module"], comm["weight_attr"]), src=min(comm["ranks"]), group=comm["group"])

    def _index_tied_modules(self):
        """{key: ranks, group, weight_attr, module} for the tied modules on this rank. Group creation is collective:
        every rank creates the group of every (key, data, model) coordinate, in the same order."""
        comms = {}
        if self.num_stages == 1:
            return comms
        owners = {}
        for idx, spec in enumerate(self._layer_specs):
            if isinstance(spec, TiedLayerSpec):
                owners.setdefault(spec.key, set()).add(self.stage_owner(idx))
        tp = self._grid.get_slice_parallel_world_size()
        for key in sorted(owners):
            for dp in range(self._grid.data_parallel_size):
                for mp in range(tp):
                    where = {"data": dp, "model": mp} if tp > 1 else {"data": dp}
                    ranks = [self._grid.stage_to_global(stage_id=s, **where) for s in sorted(owners[key])]
                    group = dist.new_group(ranks=ranks)
                    if self.global_rank not in ranks:
                        continue
                    assert key in self.tied_modules
                    comms[key] = {"ranks": ranks, "group": group, "weight_attr": self.tied_weight_attrs[key],
                                  "module": self.tied_modules[key]}
                    if self.global_rank != ranks[0]:
                        # for gradient norms the shared weight is counted once, by the lowest rank of its group
                        for p in self.tied_modules[key].parameters():
                            p.model_parallel = False
        return comms

    # ------------------------------------------------------------------ queries
    def partitions(self):
        return self.parts

    def stage_owner(self, layer_idx):
        assert 0 <= layer_idx < self._num_layers
        for stage in range(self.num_stages):
            if self.parts[stage] <= layer_idx < self.parts[stage + 1]:
                return stage
        raise RuntimeError(f"Layer {layer_idx} not owned? parts={self.parts}")

    def topology(self):
        return self._topo

    def mpu(self):
        return self._grid

    def num_pipeline_stages(self):
        return self.num_stages

    # ------------------------------------------------------------------ checkpoints: one file per layer
    def ckpt_prefix(self, checkpoints_path, tag):
        coord = self._topo.get_coord(rank=self.global_rank)
        name = "module" + "".join(f"-{axis}_{getattr(coord, axis):02d}"
                                  for axis in self._topo.get_axis_names() if axis != "data")
        return os.path.join(checkpoints_path, str(tag), name)

    def ckpt_layer_path(self, ckpt_dir, local_layer_idx):
        """layer_<global index>[-model_<tp rank>]-model_states.pt"""
        stem = f"layer_{local_layer_idx + self._local_start:02d}"
        tp_tag = self._topo.get_rank_repr(rank=self.global_rank)
        return os.path.join(ckpt_dir, (f"{stem}-{tp_tag}" if tp_tag else stem) + "-model_states.pt")

    def ckpt_layer_path_list(self, ckpt_dir, local_layer_idx):
        stem = f"layer_{local_layer_idx + self._local_start:02d}-"
        return sorted(glob.glob(os.path.join(ckpt_dir, stem + "*model_states.pt")))

    def save_state_dict(self, save_dir):
        if self._grid.data_parallel_id != 0:      # replicas hold identical layers
            return
        os.makedirs(save_dir, exist_ok=True)
        for i, layer in enumerate(self.forward_funcs):
            if not hasattr(layer, "state_dict"):
                continue
            state = layer.state_dict()
            # CPU copies, so that loading never depends on the device the checkpoint was written from
            torch.save(type(state)((k, v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in state.items()),
                       self.ckpt_layer_path(save_dir, i))

    def load_state_dir(self, load_dir, strict=True):
        for i, layer in enumerate(self.forward_funcs):
            if hasattr(layer, "load_state_dict"):
                state = torch.load(self.ckpt_layer_path(load_dir, i), map_location="cpu", weights_only=False)
                layer.load_state_dict(state, strict=strict)
        self._synchronize_tied_weights()
