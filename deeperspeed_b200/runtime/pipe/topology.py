"""Cartesian process topology and the pipeline / data / model process-group grid (the `mpu` interface).

Public behaviour per reference deepspeed/runtime/pipe/topology.py:13-468 (axis names, rank order, group membership,
the Megatron-style accessors). Mechanism: a topology is a mixed-radix number system. Axis k has stride
prod(dims[k+1:]), a coordinate is the digit vector of the rank, and every query (rank of a coordinate, the ranks of an
axis group, a filtered slice) is stride arithmetic; nothing is enumerated up front.
"""
from collections import namedtuple
from itertools import product

import torch.distributed as dist


class ProcessTopology:
    """n-dimensional cartesian coordinates <-> linear ranks, row-major (the last axis varies fastest)."""

    def __init__(self, axes, dims):
        self.axes = list(axes)
        self.dims = list(dims)
        self.ProcessCoord = namedtuple("ProcessCoord", axes)
        self._strides = [1] * len(self.dims)
        for k in range(len(self.dims) - 2, -1, -1):
            self._strides[k] = self._strides[k + 1] * self.dims[k + 1]
        self._size = self._strides[0] * self.dims[0] if self.dims else 1

    # ---- digits <-> rank
    def _digits(self, rank):
        return [(rank // s) % d for s, d in zip(self._strides, self.dims)]

    def _rank_of(self, digits):
        return sum(x * s for x, s in zip(digits, self._strides))

    def _axis(self, axis):
        return self.axes.index(axis)

    @property
    def mapping(self):
        """{coordinate: rank} for every rank (built on demand; kept because callers print and iterate it)."""
        return {self.get_coord(r): r for r in range(self._size)}

    def get_rank(self, **coord_kwargs):
        if len(coord_kwargs) != len(self.axes):
            raise ValueError("get_rank() does not support slices. Use filter_match())")
        digits = []
        for axis, dim in zip(self.axes, self.dims):
            assert axis in coord_kwargs and 0 <= coord_kwargs[axis] < dim, f"key {coord_kwargs} invalid"
            digits.append(coord_kwargs[axis])
        return self._rank_of(digits)

    def get_axis_names(self):
        return self.axes

    def get_rank_repr(self, rank, omit_axes=("data", "pipe"), inner_sep="_", outer_sep="-"):
        """e.g. 'model_00' (checkpoint file names are derived from this)."""
        digits = self._digits(rank)
        skip = frozenset(omit_axes)
        return outer_sep.join(f"{a}{inner_sep}{x:02d}" for a, x in zip(self.axes, digits) if a not in skip)

    def get_dim(self, axis):
        return self.dims[self._axis(axis)] if axis in self.axes else 0

    def get_coord(self, rank):
        if not 0 <= rank < self._size:
            raise ValueError(f"rank {rank} not found in topology.")
        return self.ProcessCoord(*self._digits(rank))

    def get_axis_comm_lists(self, axis):
        """Groups of ranks that differ only along `axis`, ordered by their first member."""
        if axis not in self.axes:
            return []
        k = self._axis(axis)
        stride, dim = self._strides[k], self.dims[k]
        firsts = [r for r in range(self._size) if (r // stride) % dim == 0]
        return [[first + i * stride for i in range(dim)] for first in firsts]

    def filter_match(self, **filter_kwargs):
        """All ranks whose coordinates agree with the given axis values."""
        fixed = {self._axis(a): v for a, v in filter_kwargs.items()}
        free = [range(d) if k not in fixed else (fixed[k],) for k, d in enumerate(self.dims)]
        return [self._rank_of(digits) for digits in product(*free)]

    def get_axis_list(self, axis, idx):
        return self.filter_match(**{axis: idx})

    def world_size(self):
        return self._size

    def __str__(self):
        return str(self.mapping)


def _prime_factors(N):
    """Prime factors of N in non-decreasing order."""
    if N <= 0:
        raise ValueError("Values must be strictly positive.")
    out, p = [], 2
    while N > 1:
        if N % p == 0:
            out.append(p)
            N //= p
        else:
            p += 1
    return out


class PipeDataParallelTopology(ProcessTopology):
    """pipe x data: the data-parallel replicas of a stage have adjacent ranks, so gradient reductions stay among
    neighbouring GPUs while only activations cross between stages."""

    def __init__(self, num_pp, num_dp):
        super().__init__(axes=["pipe", "data"], dims=[num_pp, num_dp])


class PipeModelDataParallelTopology(ProcessTopology):
    """3-D pipe x data x model (tensor-parallel ranks innermost)."""

    def __init__(self, num_pp, num_mp, num_dp):
        super().__init__(axes=["pipe", "data", "model"], dims=[num_pp, num_dp, num_mp])


class PipelineParallelGrid:
    """Every process group a pipeline engine needs, plus the Megatron-style `mpu` accessors.

    Groups: data-parallel (same pipe / model coordinates), pipe (same data / model), "ds model" (all ranks that share
    my data coordinate; the checkpoint's mp_rank is my index in it) and slice (tensor-parallel). `p2p_groups` lists
    every rank with its next-stage neighbour (wrapping around); the communicators for those pairs are created by
    runtime/pipe/p2p.py, one per direction."""

    def __init__(self, topology=None, process_group=None):
        self.global_rank = dist.get_rank()
        self.world_size = dist.get_world_size()
        self._topo = topology if topology is not None else self._default_topology(self.world_size)
        topo = self._topo
        self.data_parallel_size = max(topo.get_dim("data"), 1)
        self.pipe_parallel_size = max(topo.get_dim("pipe"), 1)
        self.model_parallel_size = max(topo.get_dim("model"), 1)
        self.slice_parallel_size = self.model_parallel_size
        assert self._is_grid_valid(), "Invalid Grid"

        me = topo.get_coord(self.global_rank)
        self.stage_id = me.pipe
        self.data_parallel_id = me.data
        self.is_first_stage = self.stage_id == 0
        self.is_last_stage = self.stage_id == self.pipe_parallel_size - 1

        # group creation is collective: every rank walks the same lists in the same order
        ds_model_lists = [sorted(topo.get_axis_list("data", dp)) for dp in range(self.data_parallel_size)]
        ranks, self.ds_model_proc_group = self._join(ds_model_lists)
        self.ds_model_world_size = len(ranks)
        self.ds_model_rank = ranks.index(self.global_rank)

        self.dp_groups = topo.get_axis_comm_lists("data")
        self.dp_group, self.dp_proc_group = self._join(self.dp_groups)

        self.p2p_groups = self._build_p2p_groups()

        self.pipe_groups = topo.get_axis_comm_lists("pipe")
        self.pp_group, self.pp_proc_group = self._join(self.pipe_groups)

        if "model" in topo.get_axis_names():
            self.model_groups = topo.get_axis_comm_lists("model")
        else:
            self.model_groups = [[r] for r in range(self.world_size)]
        self.mp_group = []
        if self.world_size > 1 or "model" in topo.get_axis_names():
            self.slice_group, self.slice_proc_group = self._join(self.model_groups)
        else:
            self.slice_group, self.slice_proc_group = [self.global_rank], None

    @staticmethod
    def _default_topology(world_size):
        """No topology given: deal the prime factors of the world size alternately to the pipe and data axes."""
        num_pp = num_dp = 1
        for i, prime in enumerate(_prime_factors(world_size)):
            if i % 2 == 0:
                num_pp *= prime
            else:
                num_dp *= prime
        return PipeDataParallelTopology(num_dp=num_dp, num_pp=num_pp)

    def _join(self, rank_lists):
        """Create one process group per list; return (ranks, group) of the list this rank belongs to."""
        mine = (None, None)
        for ranks in rank_lists:
            group = dist.new_group(ranks=ranks)
            if self.global_rank in ranks:
                mine = (ranks, group)
        assert mine[0] is not None
        return mine

    def get_stage_id(self):
        return self._topo.get_coord(rank=self.global_rank).pipe

    def get_data_parallel_id(self):
        return self._topo.get_coord(rank=self.global_rank).data

    def _build_p2p_groups(self):
        """[rank, next-stage rank] for every rank, the last stage wrapping around to the first."""
        topo = self._topo
        if "pipe" not in topo.get_axis_names():
            return [[r, r] for r in range(self.world_size)]
        k = topo._axis("pipe")
        stride, dim = topo._strides[k], topo.dims[k]
        pairs = []
        for rank in range(self.world_size):
            stage = (rank // stride) % dim
            pairs.append([rank, rank + (((stage + 1) % dim) - stage) * stride])
        return pairs

    def _is_grid_valid(self):
        return self._topo.world_size() == dist.get_world_size()

    def stage_to_global(self, stage_id, **kwargs):
        """Global rank of the process with my coordinates except pipe = stage_id (and any overrides in kwargs)."""
        coord = self._topo.get_coord(self.global_rank)._replace(pipe=stage_id, **kwargs)
        return self._topo.get_rank(**coord._asdict())

    def topology(self):
        return self._topo

    # ---- mpu interface
    def get_global_rank(self):
        return self.global_rank

    def get_pipe_parallel_rank(self):
        return self.stage_id

    def get_pipe_parallel_world_size(self):
        return self.pipe_parallel_size

    def get_pipe_parallel_group(self):
        return self.pp_proc_group

    def get_data_parallel_rank(self):
        return self.data_parallel_id

    def get_data_parallel_world_size(self):
        return self.data_parallel_size

    def get_data_parallel_group(self):
        return self.dp_proc_group

    def get_data_parallel_src_rank(self):
        return self.dp_group[0]

    def get_model_parallel_rank(self):
        return self.ds_model_rank

    def get_model_parallel_world_size(self):
        return self.ds_model_world_size

    def get_model_parallel_group(self):
        return self.ds_model_proc_group

    def get_slice_parallel_rank(self):
        return getattr(self._topo.get_coord(self.global_rank), "model", 0)

    def get_slice_parallel_world_size(self):
        return self.slice_parallel_size

    def get_slice_parallel_group(self):
        return self.slice_proc_group
