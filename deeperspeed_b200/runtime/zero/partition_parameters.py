"""ZeRO-3 parameter partitioning: `zero.Init`, `GatheredParameters`, external-parameter registration.

API per reference deepspeed/runtime/zero/partition_parameters.py:56-1130. Every parameter created inside
`zero.Init` becomes a "ds param": it keeps only `ds_tensor` (its 1/dp shard, zero-padded to a multiple of dp) and the
methods all_gather()/partition()/... ; `param.data` is a 1-element placeholder while the parameter is NOT_AVAILABLE.

Mechanism (torch 2.x): while the context is active every nn.Module subclass `__init__` is wrapped so that, right
after a sub-module finishes constructing, its directly-owned parameters are cast, moved and scattered — the
reference additionally monkey-patches torch.Tensor.__new__/torch.empty (partition_parameters.py:183-261); here
construction happens under `torch.device(<target>)` instead, which covers the same factory calls.
"""
import functools
import os
from enum import Enum

import torch
import torch.distributed as dist

from ...utils.logging import logger
from ...utils.distributed import init_distributed

param_count = 0
_orig_inits = {}
partitioned_param_data_shape = [1]      # what `param.data` looks like while the parameter is partitioned


class ZeroParamType(Enum):
    NORMAL = 1       # same as regular pytorch parameters
    PARTITIONED = 2  # parameters are partitioned across data parallel process
    REMOTE = 3       # parameters are hosted on a remote device (cpu / nvme)


class ZeroParamStatus(Enum):
    AVAILABLE = 1      # parameters are fully present and ready for use on all processes
    NOT_AVAILABLE = 2  # parameters are partitioned, only the local shard exists
    INFLIGHT = 3       # an all-gather of this parameter has been issued and not yet waited on


def print_rank_0(message, debug=False, force=False):
    if (debug or force) and (not dist.is_initialized() or dist.get_rank() == 0):
        print(message)


def is_zero_param(parameter):
    return hasattr(parameter, "ds_id")


def _init_external_params(module):
    if not hasattr(module, "_external_params"):
        module._external_params = {}

        def external_parameters(self):
            if not hasattr(self, "_external_params"):
                self._external_params = {}
            return self._external_params.items()

        def all_parameters(self):
            import itertools
            return itertools.chain(self.named_parameters(self, recurse=False), external_parameters(self))

        module.ds_external_parameters = external_parameters.__get__(module)
        module.all_parameters = all_parameters.__get__(module)


def register_external_parameter(module, parameter):
    """Tell ZeRO-3 that `module`'s forward uses `parameter` owned by another module, so it is gathered with it."""
    if not isinstance(parameter, torch.nn.Parameter):
        raise RuntimeError("Parameter is not a torch.nn.Parameter")
    _init_external_params(module)
    key = id(parameter)
    module._external_params[key] = parameter


def unregister_external_parameter(module, parameter):
    if not isinstance(parameter, torch.nn.Parameter):
        raise RuntimeError("Parameter is not a torch.nn.Parameter")
    if not hasattr(module, "_external_params") or id(parameter) not in module._external_params:
        raise RuntimeError("Parameter is not a registered external parameter of module.")
    del module._external_params[id(parameter)]


class ZeROOrderedDict(dict):
    """`module._parameters` replacement that auto-registers external parameters: reading a partitioned parameter
    through a module that is currently in forward gathers it and records it as external to the parent."""

    def __init__(self, parent_module, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self._parent_module = parent_module
        self._in_forward = False

    def __getitem__(self, key):
        param = super().__getitem__(key)
        if param is None:
            return param
        if hasattr(param, "ds_status") and param.ds_status == ZeroParamStatus.NOT_AVAILABLE:
            if self._parent_module._parameters._in_forward:
                if getattr(param, "ds_fused_ag", False) and key == "weight" \
                        and "forward" in self._parent_module.__dict__:
                    return param   # its Linear gathers it inside the GEMM kernel (stage3: _fused_linear_forward)
                register_external_parameter(FWD_MODULE_STACK[-1], param)
                param.all_gather()
        return param


FWD_MODULE_STACK = []


def inject_parameters(module, cls):
    for m in module.modules():
        if cls is ZeROOrderedDict:
            new_param = cls(parent_module=m)
        else:
            new_param = cls()
        for key, param in m._parameters.items():
            new_param[key] = param
        m._parameters = new_param


def _all_subclasses(cls):
    seen, stack = set(), [cls]
    while stack:
        c = stack.pop()
        for sub in c.__subclasses__():
            if sub not in seen:
                seen.add(sub)
                stack.append(sub)
    return seen


class InsertPostInitMethodToModuleSubClasses:
    """Context manager base: wraps every nn.Module subclass __init__ with a post-init callback."""

    def __init__(self, enabled=True, mem_efficient_linear=True):
        self.mem_efficient_linear = mem_efficient_linear
        self.enabled = enabled

    def __enter__(self):
        if not self.enabled:
            return

        def partition_after(f):
            @functools.wraps(f)
            def wrapper(module, *args, **kwargs):
                # only the outermost __init__ frame of an object post-processes it
                depth_attr = "_ds_init_depth"
                module.__dict__[depth_attr] = module.__dict__.get(depth_attr, 0) + 1
                try:
                    f(module, *args, **kwargs)
                finally:
                    module.__dict__[depth_attr] -= 1
                if module.__dict__[depth_attr] == 0:
                    self._post_init_method(module)
            return wrapper

        def _enable_class(cls):
            if "__init__" in cls.__dict__ and cls not in _orig_inits:
                _orig_inits[cls] = cls.__dict__["__init__"]
                cls.__init__ = partition_after(cls.__dict__["__init__"])

        def _init_subclass(cls, **kwargs):
            _enable_class(cls)

        for subclass in _all_subclasses(torch.nn.modules.module.Module):
            _enable_class(subclass)
        self._old_init_subclass = torch.nn.modules.module.Module.__dict__.get("__init_subclass__", None)
        torch.nn.modules.module.Module.__init_subclass__ = classmethod(_init_subclass)
        self._device_ctx = None
        target = getattr(self, "_construct_device", None)
        if target is not None:
            self._device_ctx = torch.device(target)
            self._device_ctx.__enter__()
        if self.mem_efficient_linear:
            from .linear import LinearFunctionForZeroStage3
            self.linear_bk = torch.nn.functional.linear
            torch.nn.functional.linear = LinearFunctionForZeroStage3.apply_compat

    def __exit__(self, exc_type, exc_value, traceback):
        if not self.enabled:
            return False
        if self._device_ctx is not None:
            self._device_ctx.__exit__(exc_type, exc_value, traceback)
        for cls, orig in list(_orig_inits.items()):
            cls.__init__ = orig
        _orig_inits.clear()
        if self._old_init_subclass is None:
            try:
                del torch.nn.modules.module.Module.__init_subclass__
            except AttributeError:
                pass
        else:
            torch.nn.modules.module.Module.__init_subclass__ = self._old_init_subclass
        if self.mem_efficient_linear:
            torch.nn.functional.linear = self.linear_bk
        return False

    def _post_init_method(self, module):
        pass


class Init(InsertPostInitMethodToModuleSubClasses):
    param_id = 0

    def __init__(self, module=None, data_parallel_group=None, mem_efficient_linear=True, remote_device=None,
                 pin_memory=False, deepspeed_config=None, param_dict=None, enabled=True, fused_allgather_gemm=None):
        """Partition parameters as they are constructed so models larger than one device can be built.
        `remote_device`: None/"cuda" keeps shards on the GPU, "cpu" (optionally pinned) or "nvme" host them remotely.
        `fused_allgather_gemm` (default: env DSB_ZERO3_FUSED_AG=1): shards of eligible 2-D weights live on a symmetric
        (peer-mapped) heap and nn.Linear forwards run the all-gather INSIDE the GEMM kernel (`parallel/ag_gemm.py`)
        instead of all_gather + linear; backward still gathers the weight the regular way."""
        super().__init__(enabled=enabled, mem_efficient_linear=mem_efficient_linear)
        if not dist.is_initialized():
            init_distributed()
            assert dist.is_initialized(), "Parameters cannot be scattered without initializing torch.distributed"
        self.ds_process_group = data_parallel_group if data_parallel_group is not None else dist.group.WORLD
        self.rank = dist.get_rank(group=self.ds_process_group)
        self.world_size = dist.get_world_size(group=self.ds_process_group)
        if torch.cuda.is_available():
            self.local_device = torch.device("cuda:{}".format(int(os.environ.get("LOCAL_RANK", 0))))
            torch.cuda.set_device(self.local_device)
        else:
            self.local_device = torch.device("cpu")
        self._construct_device = self.local_device if self.local_device.type == "cuda" else None
        self.remote_device = self.local_device if remote_device in (None, "cuda", "none") else remote_device
        self.pin_memory = pin_memory if (self.remote_device == "cpu") else False
        self.dtype = torch.half if torch.cuda.is_available() else torch.float32
        if deepspeed_config is not None or param_dict is not None:
            from ..config import DeepSpeedConfig
            cfg = DeepSpeedConfig(deepspeed_config, param_dict=param_dict)
            if cfg.fp16_enabled:
                self.dtype = cfg.precision
            off = cfg.zero_config.offload_param
            if off is not None and remote_device is None:
                self.remote_device = off["device"]
                self.pin_memory = off.get("pin_memory", False)
        self.param_swapper = None
        if fused_allgather_gemm is None:
            fused_allgather_gemm = os.environ.get("DSB_ZERO3_FUSED_AG", "0") == "1"
        self.fused_allgather_gemm = bool(fused_allgather_gemm) and self.local_device.type == "cuda" \
            and 1 < self.world_size <= 8 and self.remote_device == self.local_device
        self._ag_heaps = []
        # already constructed module: convert in place
        if module is not None:
            assert isinstance(module, torch.nn.Module)
            for param in module.parameters(recurse=True):
                if is_zero_param(param):
                    continue
                self._convert_to_deepspeed_param(param)
                param.partition()

    def _post_init_method(self, module):
        print_rank_0(f"Converting Params in {module.__class__.__name__}", force=False)
        for name, param in module.named_parameters(recurse=False):
            if not is_zero_param(param):
                self._convert_to_deepspeed_param(param)
                param.data = param.data.to(self.local_device)
                if param.is_floating_point():
                    param.data = param.data.to(self.dtype)
                param.partition()
                global param_count
                param_count += param.ds_numel

    def _convert_to_deepspeed_param(self, param):
        param.ds_param_type = ZeroParamType.PARTITIONED
        param.ds_status = ZeroParamStatus.AVAILABLE
        param.ds_shape = param.shape
        param.ds_numel = param.numel()
        param.ds_tensor = None
        param.ds_active_sub_modules = 0
        param.ds_persist = False
        param.ds_fused_ag = False
        param.ds_id = Init.param_id
        Init.param_id += 1
        param.ds_process_group = self.ds_process_group
        param.ds_world = self.world_size
        param.ds_rank = self.rank
        init = self

        def all_gather(param_list=None, async_op=False, hierarchy=0):
            cls = param
            if param_list is None:
                param_list = [cls]
            return init._all_gather(param_list, async_op=async_op, hierarchy=hierarchy)

        def partition(param_list=None, hierarchy=0, has_been_updated=False):
            cls = param
            if param_list is None:
                param_list = [cls]
            init._partition(param_list, has_been_updated=has_been_updated)

        def reduce_gradients_at_owner(param_list=None, hierarchy=0):
            cls = param
            if param_list is None:
                param_list = [cls]
            init._reduce_scatter_gradients(param_list)

        def partition_gradients(param_list=None, partition_buffers=None, hierarchy=0, accumulate=False):
            cls = param
            if param_list is None:
                param_list = [cls]
            if isinstance(partition_buffers, torch.Tensor):
                partition_buffers = [partition_buffers]
            init._partition_gradients(param_list, partition_buffers=partition_buffers, accumulate=accumulate)

        def aligned_size():
            return init._aligned_size(param)

        def padding_size():
            return init._padding_size(param)

        def partitioned_size():
            return init._partitioned_size(param)

        param.all_gather = all_gather
        param.partition = partition
        param.reduce_gradients_at_owner = reduce_gradients_at_owner
        param.partition_gradients = partition_gradients
        param.aligned_size = aligned_size
        param.padding_size = padding_size
        param.partitioned_size = partitioned_size

    # ---- symmetric-heap shards for the fused all-gather + GEMM forward
    AG_HEAP_CHUNK_BYTES = 256 << 20

    def _fused_ag_eligible(self, param):
        if not self.fused_allgather_gemm or len(param.ds_shape) != 2:
            return False
        rows, cols = param.ds_shape
        return (param.dtype in (torch.float16, torch.bfloat16) and rows % (256 * self.world_size) == 0
                and cols % 8 == 0)

    def _ag_alloc(self, numel, dtype):
        """Carve a shard out of the peer-mapped heap (collective: every rank allocates the same sequence)."""
        from ...parallel.symm import SymmetricHeap
        nbytes = (numel * 2 + 255) // 256 * 256
        heap = self._ag_heaps[-1] if self._ag_heaps else None
        if heap is None or heap._cursor + nbytes > heap.nbytes:
            heap = SymmetricHeap(max(self.AG_HEAP_CHUNK_BYTES, nbytes + 4096), group=self.ds_process_group)
            self._ag_heaps.append(heap)
        t, off = heap.alloc(numel, dtype)
        return t, heap.peer_ptrs(off)

    # ---- size helpers (per-parameter: pad = (-numel) mod dp; reference partition_parameters.py:547-555)
    def _aligned_size(self, param):
        return param.ds_numel + self._padding_size(param)

    def _padding_size(self, param):
        remainder = param.ds_numel % self.world_size
        return (self.world_size - remainder) if remainder else 0

    def _partitioned_size(self, param):
        return self._aligned_size(param) // self.world_size

    # ---- gather / scatter
    def _all_gather(self, param_list, async_op=False, hierarchy=None):
        handles = []
        to_gather = []
        for param in param_list:
            if param.ds_status == ZeroParamStatus.NOT_AVAILABLE:
                if async_op:
                    handles.append(self._allgather_param(param, async_op=True))
                    param.ds_status = ZeroParamStatus.INFLIGHT
                else:
                    to_gather.append(param)
        if not async_op:
            for p in to_gather:
                self._allgather_param(p, async_op=False)
                p.ds_status = ZeroParamStatus.AVAILABLE
            return None
        return handles

    def _shard_on_device(self, param):
        t = param.ds_tensor
        if t.device != self.local_device:
            t = t.to(self.local_device, non_blocking=True)
        return t

    def _allgather_param(self, param, async_op=False):
        swap = getattr(param, "ds_swap", None)      # shard on NVMe (ZeRO-3 optimizer's arena swapper): bring it in
        if swap is None:
            return self._allgather_resident(param, async_op)
        swapper, arena = swap
        swapper.acquire(arena)
        try:
            shard = self._shard_on_device(param)
            if shard.device.type == "cpu":
                # stays in host memory (CPU run): the pool buffer may be recycled before an asynchronous gather ends
                shard = shard.clone()
            return self._allgather_resident(param, async_op, shard)
        finally:
            swapper.release(arena)      # device copies out of the buffer are fenced by the swapper's events

    def _allgather_resident(self, param, async_op=False, shard=None):
        part = self._partitioned_size(param)
        if shard is None:
            shard = self._shard_on_device(param)
        flat = torch.empty(part * self.world_size, dtype=shard.dtype, device=self.local_device)
        handle = None
        if self.world_size > 1:
            handle = dist.all_gather_into_tensor(flat, shard.contiguous(), group=self.ds_process_group, async_op=async_op)
        else:
            flat.copy_(shard)
        param.data = flat.narrow(0, 0, param.ds_numel).view(param.ds_shape).data
        return handle

    def _partition(self, param_list, force=False, has_been_updated=False):
        for param in param_list:
            self._partition_param(param, has_been_updated=has_been_updated)
            param.ds_status = ZeroParamStatus.NOT_AVAILABLE

    def _partition_param(self, param, buffer=None, has_been_updated=False):
        assert param.ds_status is not ZeroParamStatus.INFLIGHT, f"{param.ds_id} cannot partition a param in flight"
        if param.ds_status is not ZeroParamStatus.AVAILABLE:
            return
        if param.ds_tensor is not None and not has_been_updated:
            # shard is current: just drop the gathered copy
            param.data = torch.ones(partitioned_param_data_shape, dtype=param.dtype, device=self.local_device).data
            return
        swap = getattr(param, "ds_swap", None)
        if swap is not None:            # shard on NVMe: bring its arena in, update it below, store it again
            swap[0].acquire(swap[1])
        try:
            self._write_shard(param)
            refreshed = getattr(param, "ds_shard_updated", None)
            if refreshed is not None:       # the ZeRO-3 optimizer keeps an fp32 master of this shard
                refreshed(param)
        finally:
            if swap is not None:
                swap[0].write_back(swap[1])
                swap[0].release(swap[1])
        param.data = torch.ones(partitioned_param_data_shape, dtype=param.dtype, device=self.local_device).data

    def _write_shard(self, param):
        """Copy this rank's slice of the gathered `param.data` into its shard (created on first use)."""
        part = self._partitioned_size(param)
        start = part * self.rank
        one_dim = param.data.contiguous().view(-1)
        if param.ds_tensor is None and self._fused_ag_eligible(param):
            import types
            shard, ptrs = self._ag_alloc(part, param.dtype)
            param.ds_tensor = shard
            param.ds_tensor.requires_grad = False
            param.ds_tensor.ds_numel = part
            param.ds_fused_ag = True
            param.ds_heap_shard = shard.view(-1)   # survives the optimizer re-homing ds_tensor.data into its flat buffer
            param.ds_sharded = types.SimpleNamespace(world=self.world_size, rank=self.rank, N=param.ds_shape[0],
                                                     K=param.ds_shape[1], peer_ptrs=ptrs)
        if param.ds_tensor is None:
            dev = self.local_device if not isinstance(self.remote_device, str) else torch.device("cpu")
            shard = torch.zeros(part, dtype=param.dtype, device=dev)
            if self.pin_memory and dev.type == "cpu" and torch.cuda.is_available():
                shard = shard.pin_memory()
            param.ds_tensor = shard
            param.ds_tensor.requires_grad = False
            param.ds_tensor.ds_numel = part
        end = min(start + part, param.ds_numel)
        if start < param.ds_numel:
            param.ds_tensor.narrow(0, 0, end - start).copy_(one_dim.narrow(0, start, end - start))

    # ---- gradients
    def _reduce_scatter_gradients(self, param_list):
        for param in param_list:
            self._reduce_scatter_gradient(param)

    def _reduce_scatter_gradient(self, param):
        """Average the full gradient over dp and keep only this rank's slice in param.grad (flat, padded)."""
        part = self._partitioned_size(param)
        full = torch.zeros(part * self.world_size, dtype=param.grad.dtype, device=param.grad.device)
        full.narrow(0, 0, param.ds_numel).copy_(param.grad.contiguous().view(-1))
        out = torch.empty(part, dtype=full.dtype, device=full.device)
        if self.world_size > 1:
            if full.is_cuda:
                dist.reduce_scatter_tensor(out, full, op=dist.ReduceOp.AVG, group=self.ds_process_group)
            else:
                dist.reduce_scatter_tensor(out, full, group=self.ds_process_group)
                out.div_(self.world_size)
        else:
            out.copy_(full)
        return out

    def _partition_gradients(self, param_list, partition_buffers=None, accumulate=False):
        if partition_buffers is None:
            partition_buffers = [None] * len(param_list)
        for param, buf in zip(param_list, partition_buffers):
            part = self._partitioned_size(param)
            start = part * self.rank
            flat = param.grad.contiguous().view(-1)
            if buf is None:
                buf = torch.zeros(part, dtype=flat.dtype, device=flat.device)
            end = min(start + part, param.ds_numel)
            if start < param.ds_numel:
                if accumulate:
                    buf.narrow(0, 0, end - start).add_(flat.narrow(0, start, end - start))
                else:
                    buf.narrow(0, 0, end - start).copy_(flat.narrow(0, start, end - start))
            param.grad.data = buf.data


class GatheredParameters:
    def __init__(self, params, modifier_rank=None, fwd_module=None, enabled=True):
        """Temporarily gather partitioned parameters. With `modifier_rank` set, modifications made by that rank
        inside the context are broadcast to everybody before re-partitioning; `fwd_module` registers the params as
        external parameters of that module instead."""
        self.enabled = enabled
        if not enabled:
            return
        if isinstance(params, torch.nn.Module):
            params = list(params.parameters())
        elif isinstance(params, torch.nn.Parameter) or torch.is_tensor(params):
            params = [params]
        else:
            params = list(params)
        self.params = [p for p in params if is_zero_param(p)]
        self.src_rank = None
        if modifier_rank is not None and self.params:
            group = self.params[0].ds_process_group
            if group is None or group == dist.group.WORLD:
                self.src_rank = modifier_rank
            else:
                self.src_rank = dist.get_global_rank(group, modifier_rank)
        self.fwd_module = fwd_module
        if self.fwd_module is not None:
            for p in self.params:
                register_external_parameter(self.fwd_module, p)

    def __enter__(self):
        if not self.enabled:
            return
        for p in self.params:
            if p.ds_status == ZeroParamStatus.NOT_AVAILABLE:
                p.all_gather()

    def __exit__(self, *exc):
        if not self.enabled:
            return
        if self.src_rank is None:
            for p in self.params:
                if not getattr(p, "ds_persist", False):
                    p.partition(has_been_updated=False)
            return
        for p in self.params:
            if dist.get_world_size(group=p.ds_process_group) > 1:
                dist.broadcast(p.data, self.src_rank, group=p.ds_process_group)
            p.partition(has_been_updated=True)
            if getattr(p, "ds_persist", False):
                p.all_gather()
        # fused all-gather + GEMM layers read the peers' shards directly from the symmetric heap: publish the edit there
        # (the optimizer re-homes ds_tensor into its flat buffer, so the heap copy is a separate tensor) and make sure
        # every rank has done so before anyone's next forward
        edited = [p for p in self.params if getattr(p, "ds_fused_ag", False)]
        for p in edited:
            if p.ds_heap_shard.data_ptr() != p.ds_tensor.data_ptr():
                p.ds_heap_shard.copy_(p.ds_tensor.view(-1)[:p.ds_heap_shard.numel()])
        if edited and dist.get_world_size(group=edited[0].ds_process_group) > 1:
            dist.barrier(group=edited[0].ds_process_group)
