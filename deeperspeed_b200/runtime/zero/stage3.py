"""ZeRO stage 3 (+ ZeRO-Offload / ZeRO-Infinity): parameters, gradients and optimizer state are all sharded.

Reference: deepspeed/runtime/zero/stage3.py:120-3257. Behavioural contract kept: every parameter is a ds param
(partition_parameters.py) gathered right before the sub-module that uses it runs and released right after (forward
and backward), with look-ahead prefetch along the recorded execution trace, `max_live_parameters`,
`max_reuse_distance` and `param_persistence_threshold`; gradients are reduce-scattered to the owning rank as soon as
they are produced; the optimizer steps sub-group by sub-group so fp32 state can live on CPU / NVMe.

Mechanism (torch 2.x): module forward pre/post hooks; for the backward side, gradient hooks on the tensors each
forward call received and produced (`_BackwardOfOneCall` - no autograd-function wrapping of outputs as in
stage3.py:1364-1559 and none of torch's module backward hooks, so in-place updates of a sub-module's output work and a
returned parameter stays a parameter); post-accumulate-grad hooks for the gradient reduce-scatter. Each sub-group's
shards live in one flat low-precision arena + one flat gradient-shard arena, so the step is the same fused device
sequence as stages 1/2 (fused norm/overflow kernel, fused Adam with low-precision write-back)."""
import os
from collections import OrderedDict

import torch
import torch.distributed as dist

from ...ops import kernels as K
from ...utils.logging import logger
from ..fp16.loss_scaler import LossScaler, DynamicLossScaler
from .partition_parameters import (Init, ZeroParamStatus, ZeROOrderedDict, FWD_MODULE_STACK, inject_parameters,
                                   is_zero_param, _init_external_params, register_external_parameter,
                                   unregister_external_parameter)


def _dist_ready():
    return dist.is_available() and dist.is_initialized()


def _tensors_in(value, depth=2):
    """Tensors a module received or handed back: the value itself, the members of (nested) lists / tuples / dicts, or
    the tensor attributes of a plain result object."""
    if torch.is_tensor(value):
        return [value]
    if depth == 0:
        return []
    if isinstance(value, (list, tuple)):
        members = value
    elif isinstance(value, dict):
        members = value.values()
    elif hasattr(value, "__dict__") and not isinstance(value, (torch.nn.Module, type)):
        members = [v for k, v in vars(value).items() if not k.startswith("__")]
    else:
        return []
    found = []
    for m in members:
        found.extend(_tensors_in(m, depth - 1))
    return found


def _zero_params_in(output):
    """Partitioned parameters that a module handed back in its output."""
    return [t for t in _tensors_in(output) if is_zero_param(t)]


class _BackwardOfOneCall:
    """Backward-side bookkeeping of ONE forward call of a sub-module, driven by gradient hooks on the tensors that
    went in and came out (nothing is wrapped: outputs stay the tensors the module produced, so callers may modify them
    in place and a returned parameter is still that parameter). `before` runs when the first gradient arrives at an
    output - ahead of every autograd node recorded inside the module - and gathers the module's parameters; `after`
    runs when the gradient of an input is complete, i.e. the module's own nodes are done."""

    __slots__ = ("owner", "module", "started", "finished")

    def __init__(self, owner, module):
        self.owner, self.module = owner, module
        self.started = self.finished = False

    def before(self, grad):
        if not self.started:
            self.started = True
            self.owner._pre_backward(self.module)

    def after(self, grad):
        if self.started and not self.finished:
            self.finished = True
            self.owner._post_backward(self.module)


class ExecutionTrace:
    """The order in which sub-modules fetch their parameters during ONE whole iteration - forward and backward -,
    recorded while the first iteration runs and replayed afterwards: it tells the coordinator what to gather ahead
    (also in backward, where the reference looks ahead as well, stage3.py:365-440) and how soon a module runs again
    (`max_reuse_distance`: the last layers of the forward pass are the first of the backward pass and are not worth
    releasing in between). A step that leaves the recorded order (data-dependent control flow, an evaluation pass of a
    different shape) truncates the trace at that point and records the rest again."""

    def __init__(self):
        self.order = []              # module ids, one entry per fetch
        self.recording = True
        self.position = -1           # index in `order` of the latest fetch
        self.fetched_at = {}         # module id -> position of its latest fetch
        self._next_use = []          # per position: position of the next fetch of the same module (None: never again)
        self._before = []            # per position: elements fetched by all earlier positions (prefix sum)

    @property
    def ready(self):
        return not self.recording and bool(self.order)

    def begin_pass(self):
        """The root module starts a forward pass."""
        self.position = -1

    def advance(self, module_id):
        self.position += 1
        if not self.recording:
            if self.position < len(self.order) and self.order[self.position] == module_id:
                self.fetched_at[module_id] = self.position
                return
            del self.order[self.position:]          # off the recorded path: record the rest of this iteration anew
            self.recording = True
        self.order.append(module_id)
        self.fetched_at[module_id] = self.position

    def end_iteration(self, numel_of):
        """The backward pass is over. `numel_of(module_id)` sizes the entries of a freshly recorded trace."""
        if self.recording and self.order:
            self.recording = False
            sizes = [numel_of(mid) for mid in self.order]
            self._before = [0] * (len(sizes) + 1)
            for i, n in enumerate(sizes):
                self._before[i + 1] = self._before[i] + n
            self._next_use, later = [None] * len(self.order), {}
            for i in range(len(self.order) - 1, -1, -1):
                self._next_use[i] = later.get(self.order[i])
                later[self.order[i]] = i
        self.position = -1

    def upcoming(self, count):
        return self.order[self.position + 1:self.position + 1 + count] if self.ready else []

    def reuse_distance(self, module_id):
        """Elements other modules fetch before `module_id` fetches again (inf: not again in this iteration)."""
        at = self.fetched_at.get(module_id)
        if not self.ready or at is None or at >= len(self._next_use) or self._next_use[at] is None:
            return float("inf")
        return self._before[self._next_use[at]] - self._before[at + 1]


class PartitionedParameterCoordinator:
    def __init__(self, comm_stream=None, max_reuse_distance_in_numel=int(5e8), max_available_parameters_in_numel=int(7e8)):
        self.in_flight_handles = []
        self.params_in_flight = []
        self.comm_stream = comm_stream
        self.trace = ExecutionTrace()
        # weights served by the fused all-gather + GEMM kernels are not worth gathering ahead; the optimizer clears
        # this in backward when the input-gradient kernel is not in use (then backward gathers them like any other)
        self.fused_weights_stay_partitioned = True
        self.hierarchy = 0
        self.total_available_parameter_numel = 0
        self.max_available_parameters_in_numel = max_available_parameters_in_numel
        self.max_reuse_distance_in_numel = max_reuse_distance_in_numel
        self.module_by_id = {}

    def reset_step(self):
        """End of an iteration (after backward, or after a forward pass that will have none)."""
        self.trace.end_iteration(lambda mid: sum(p.ds_numel for p in self._params_of(self.module_by_id[mid])))

    def _params_of(self, sub_module):
        params = [p for p in sub_module.parameters(recurse=False) if is_zero_param(p)]
        if hasattr(sub_module, "ds_external_parameters"):
            params += [p for _, p in sub_module.ds_external_parameters() if is_zero_param(p)]
        return params

    @staticmethod
    def _fused_forward_weight(sub_module, p):
        """True when `p` is the weight of a Linear whose forward gathers it inside the GEMM kernel."""
        return (getattr(p, "ds_fused_ag", False) and type(sub_module) is torch.nn.Linear
                and p is getattr(sub_module, "weight", None) and "forward" in sub_module.__dict__)

    def fetch_sub_module(self, sub_module, forward=False):
        """Gather the parameters of `sub_module` (waiting on any prefetch already in flight). In forward, weights
        served by the fused all-gather + GEMM stay partitioned."""
        params = self._params_of(sub_module)
        to_fetch = [p for p in params if p.ds_status == ZeroParamStatus.NOT_AVAILABLE
                    and not (forward and self._fused_forward_weight(sub_module, p))]
        for p in params:
            p.ds_active_sub_modules += 1
        if to_fetch:
            for p in to_fetch:
                p.all_gather()
                self.total_available_parameter_numel += p.ds_numel
        for p in params:
            if p.ds_status == ZeroParamStatus.INFLIGHT:
                self._wait(p)
        self.trace.advance(sub_module.id)

    def _wait(self, p):
        still_in_flight = []
        for entry in self.in_flight_handles:
            params, handles = entry
            if not any(q is p for q in params):
                still_in_flight.append(entry)
                continue
            for h in handles:
                if h is not None:
                    h.wait()
            for q in params:
                q.ds_status = ZeroParamStatus.AVAILABLE
        self.in_flight_handles = still_in_flight      # (list.remove would compare the tensors element-wise)
        if p.ds_status == ZeroParamStatus.INFLIGHT:
            p.ds_status = ZeroParamStatus.AVAILABLE

    def prefetch_next_sub_modules(self, sub_module, numel=5000000):
        """Issue asynchronous all-gathers for the next sub-modules of the recorded trace, up to `numel` elements."""
        budget = numel
        for mid in self.trace.upcoming(8):
            mod = self.module_by_id.get(mid)
            if mod is None:
                continue
            params = [p for p in self._params_of(mod) if p.ds_status == ZeroParamStatus.NOT_AVAILABLE
                      and not (self.fused_weights_stay_partitioned and getattr(p, "ds_fused_ag", False))]
            cost = sum(p.ds_numel for p in params)
            if not params:
                continue
            # shards on NVMe: start the file read now and let it run under the current module's compute; the
            # gather is issued by a later call (or by the module's own fetch), when the arena is in memory
            if any([swapper.prefetch(arena) for swapper, arena in {getattr(p, "ds_swap", None) for p in params} - {None}]):
                break
            if cost > budget or self.total_available_parameter_numel + cost > self.max_available_parameters_in_numel:
                break
            handles = []
            for p in params:
                handles.extend(p.all_gather(async_op=True) or [])
                self.total_available_parameter_numel += p.ds_numel
            self.in_flight_handles.append((params, handles))
            budget -= cost

    def release_sub_module(self, sub_module):
        # a module that fetches again soon (the tail of the forward pass at the turn into backward) keeps its
        # parameters, as long as the live-parameter budget holds
        soon = (self.trace.reuse_distance(sub_module.id) < self.max_reuse_distance_in_numel
                and self.total_available_parameter_numel <= self.max_available_parameters_in_numel)
        for p in self._params_of(sub_module):
            p.ds_active_sub_modules = max(p.ds_active_sub_modules - 1, 0)
            if (p.ds_active_sub_modules == 0 and not p.ds_persist and p.ds_status == ZeroParamStatus.AVAILABLE
                    and not soon):
                p.partition()
                self.total_available_parameter_numel -= p.ds_numel

    def release_and_reset_parameter(self, p):
        p.ds_active_sub_modules = 0
        if p.ds_status == ZeroParamStatus.INFLIGHT:
            self._wait(p)
        if not p.ds_persist and p.ds_status == ZeroParamStatus.AVAILABLE:
            p.partition()
            self.total_available_parameter_numel -= p.ds_numel


class FP16_DeepSpeedZeroOptimizer_Stage3:
    def __init__(self, module, init_optimizer, timers=None, static_loss_scale=1.0, dynamic_loss_scale=False,
                 dynamic_loss_args=None, verbose=True, contiguous_gradients=True, reduce_bucket_size=500000000,
                 prefetch_bucket_size=50000000, max_reuse_distance=1000000000, max_live_parameters=1000000000,
                 param_persistence_threshold=100000, dp_process_group=None, reduce_scatter=True, overlap_comm=False,
                 offload_optimizer_config=None, offload_param_config=None, sub_group_size=1000000000000, mpu=None,
                 clip_grad=0.0, allreduce_always_fp32=False, postscale_gradients=True, gradient_predivide_factor=1.0,
                 gradient_accumulation_steps=1, elastic_checkpoint=False, aio_config=None, precision=torch.float16,
                 fused_allgather_gemm=None):
        self.module = module
        self.optimizer = init_optimizer
        self.timers = timers
        self.mpu = mpu
        self.clip_grad = clip_grad
        self.verbose = verbose
        self.elastic_checkpoint = elastic_checkpoint
        self.overflow = False
        self.precision = precision
        self.dp_process_group = dp_process_group
        self.partition_count = dist.get_world_size(group=dp_process_group) if _dist_ready() else 1
        self.dp_rank = dist.get_rank(group=dp_process_group) if _dist_ready() else 0
        self.prefetch_elements = int(prefetch_bucket_size)
        self.sub_group_size = int(sub_group_size)
        self.micro_step_id = 0
        self.is_gradient_accumulation_boundary = True
        self._global_grad_norm = 0.0
        self.gradient_accumulation_steps = gradient_accumulation_steps

        if dynamic_loss_scale and precision == torch.float16:
            self.loss_scaler = DynamicLossScaler() if dynamic_loss_args is None else DynamicLossScaler(**dynamic_loss_args)
            self.dynamic_loss_scale = True
        else:
            self.loss_scaler = LossScaler(scale=static_loss_scale if precision == torch.float16 else 1.0)
            self.dynamic_loss_scale = False

        # ---- offload configuration
        self.offload_optimizer = offload_optimizer_config is not None
        self.offload_optimizer_device = offload_optimizer_config["device"] if self.offload_optimizer else None
        self.offload_optimizer_pin = bool(offload_optimizer_config.get("pin_memory", False)) if self.offload_optimizer else False
        self.offload_param = offload_param_config is not None
        self.swap_optimizer = self.offload_optimizer and self.offload_optimizer_device == "nvme"
        self.optimizer_swapper = None
        self.swap_params = self.offload_param and offload_param_config.get("device") == "nvme"
        self.param_swapper = None

        # ---- make sure every parameter is a (partitioned) ds param
        if any(not is_zero_param(p) for p in module.parameters()):
            remote = offload_param_config["device"] if self.offload_param else None
            Init(module=module, data_parallel_group=dp_process_group, mem_efficient_linear=False,
                 fused_allgather_gemm=fused_allgather_gemm,
                 remote_device=remote, pin_memory=bool(offload_param_config.get("pin_memory", False)) if self.offload_param else False)
        self.device = next((p.ds_tensor.device for p in module.parameters() if torch.cuda.is_available() is False), None)
        self.device = torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")
        for p in module.parameters():
            p.ds_persist = p.ds_numel <= param_persistence_threshold
        inject_parameters(module, ZeROOrderedDict)

        self.param_coordinator = PartitionedParameterCoordinator(
            comm_stream=torch.cuda.Stream() if (overlap_comm and torch.cuda.is_available()) else None,
            max_reuse_distance_in_numel=int(max_reuse_distance), max_available_parameters_in_numel=int(max_live_parameters))
        self._setup_module_hooks()

        if self.swap_params:
            # ZeRO-Infinity: the low-precision shard arenas live in files, a pool of pinned buffers caches the ones in
            # use (swap_tensor/arena_swapper.py); a sub-group must fit one buffer
            from ..swap_tensor.arena_swapper import ArenaSwapper
            self.param_swapper = ArenaSwapper(offload_param_config["nvme_path"], next(module.parameters()).dtype,
                                              offload_param_config.get("buffer_count", 5),
                                              offload_param_config.get("buffer_size", int(1e8)), aio_config or {})
            self.sub_group_size = min(self.sub_group_size, self.param_swapper.largest_arena())

        # ---- sub-groups: flat low-precision shard arena + flat grad-shard arena + fp32 master per sub-group
        self.fp16_groups = []
        self.sub_groups = []          # list of lists of params
        self.fp16_partitioned_groups_flat = []
        self.grad_partitions_flat = []
        self.fp32_partitioned_groups_flat = []
        self.sub_group_to_group_id = {}
        self._param_slot = {}         # param -> (sub_group index, offset in flat shard)
        for gid, group in enumerate(self.optimizer.param_groups):
            params = [p for p in group["params"] if p.requires_grad]
            self.fp16_groups.append(params)
            cur, cur_n = [], 0
            for p in params:
                n = p.ds_tensor.numel()
                if cur and cur_n + n > self.sub_group_size:
                    self._add_sub_group(cur, gid)
                    cur, cur_n = [], 0
                cur.append(p)
                cur_n += n
            if cur:
                self._add_sub_group(cur, gid)
        self._rebuild_inner_param_groups()
        self._persist_small_params()
        self._peer_shard_fence()
        if self.swap_optimizer:
            from ..swap_tensor.partitioned_optimizer_swapper import PartitionedOptimizerSwapper
            from ..swap_tensor.pipelined_optimizer_swapper import PipelinedOptimizerSwapper
            pipelined = offload_optimizer_config.get("pipeline_read", False) or offload_optimizer_config.get("pipeline_write", False)
            cls = PipelinedOptimizerSwapper if pipelined else PartitionedOptimizerSwapper
            self.optimizer_swapper = cls(swap_config=offload_optimizer_config, aio_config=aio_config or {},
                                         base_folder=offload_optimizer_config["nvme_path"], optimizer=self.optimizer,
                                         largest_numel=max(t.numel() for t in self.fp32_partitioned_groups_flat),
                                         device=self.device, dtype=torch.float32, timers=timers)
        self._install_grad_hooks()
        for params in self.sub_groups:
            for p in params:
                p.ds_shard_updated = self._master_follows_shard
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._sumsq = torch.zeros(2, dtype=torch.float32, device=self.device)
        self.fused_inner = (bool(getattr(self.optimizer, "supports_flat_fused", False)) and not self.offload_optimizer
                            and not self.offload_param)      # the fused kernel writes the low-precision arena on the device

    # ------------------------------------------------------------------ construction helpers
    def _add_sub_group(self, params, gid):
        idx = len(self.sub_groups)
        total = sum(p.ds_tensor.numel() for p in params)
        lp_dev = params[0].ds_tensor.device
        if self.swap_params:
            flat = self._register_swapped_arena(idx, params, total)
        else:
            flat = torch.zeros(total, dtype=params[0].dtype, device=lp_dev)
            if lp_dev.type == "cpu" and torch.cuda.is_available() and self.offload_param:
                flat = flat.pin_memory()
        gflat = torch.zeros(total, dtype=params[0].dtype, device=self.device)
        off = 0
        for p in params:
            n = p.ds_tensor.numel()
            flat.narrow(0, off, n).copy_(p.ds_tensor)
            p.ds_tensor.data = flat.narrow(0, off, n)
            self._param_slot[p] = (idx, off)
            off += n
        master_dev = torch.device("cpu") if self.offload_optimizer else self.device
        master = flat.detach().to(master_dev, dtype=torch.float32).clone()
        if master_dev.type == "cpu" and self.offload_optimizer_pin and torch.cuda.is_available():
            master = master.pin_memory()
        master.requires_grad = True
        if self.swap_params:
            self.param_swapper.write_back(idx)
            self.param_swapper.release(idx)
            flat = None                 # lives in its file; `_lp_arena(idx)` brings it in
        self.sub_groups.append(params)
        self.fp16_partitioned_groups_flat.append(flat)
        self.grad_partitions_flat.append(gflat)
        self.fp32_partitioned_groups_flat.append(master)
        self.sub_group_to_group_id[idx] = gid

    def _register_swapped_arena(self, idx, params, total):
        """Sub-group `idx` gets a file instead of a permanent arena. Returns the (held) buffer to fill."""
        sizes = [p.ds_tensor.numel() for p in params]
        placeholder = torch.zeros(0, dtype=params[0].dtype)

        def attach(view):
            off = 0
            for p, n in zip(params, sizes):
                p.ds_tensor.data = placeholder if view is None else view.narrow(0, off, n)
                off += n

        for p in params:
            p.ds_swap = (self.param_swapper, idx)
        # the attach callback must not fire before the shards have been copied: register with a no-op first
        self.param_swapper.register(idx, total, lambda view: None)
        flat = self.param_swapper.acquire(idx, load=False)
        self.param_swapper.attach[idx] = attach
        return flat

    class _HeldArena:
        def __init__(self, owner, idx, load, dirty, async_write=False):
            self.owner, self.idx, self.load, self.dirty, self.async_write = owner, idx, load, dirty, async_write

        def __enter__(self):
            o = self.owner
            if o.param_swapper is None:
                return o.fp16_partitioned_groups_flat[self.idx]
            return o.param_swapper.acquire(self.idx, load=self.load)

        def __exit__(self, *exc):
            o = self.owner
            if o.param_swapper is not None:
                if self.dirty and exc[0] is None:
                    o.param_swapper.write_back(self.idx, async_op=self.async_write)
                o.param_swapper.release(self.idx)

    def _lp_arena(self, idx, load=False, dirty=True, async_write=False):
        """`with self._lp_arena(idx) as lp:` - the low-precision shard arena of sub-group `idx`, brought in from its
        file when parameters are swapped to NVMe (and written back on exit when `dirty`)."""
        return self._HeldArena(self, idx, load, dirty, async_write)

    def _master_follows_shard(self, param):
        """A shard was rewritten from outside the optimizer (`GatheredParameters(modifier_rank=...)`, weights loaded
        after `initialize`): its slice of the fp32 master takes the new values, otherwise the next step would undo the
        edit (the reference leaves the master stale, partition_parameters.py:700-745)."""
        idx, off = self._param_slot[param]
        master = self.fp32_partitioned_groups_flat[idx]
        swapped = self.optimizer_swapper is not None and master.data.numel() == 0
        if swapped:
            self.optimizer_swapper.swap_in_optimizer_state(parameter=master)
        n = param.ds_tensor.ds_numel
        master.data.narrow(0, off, n).copy_(param.ds_tensor.data.view(-1)[:n])
        if swapped:
            self.optimizer_swapper.swap_out_optimizer_state(parameter=master)

    def refresh_fp32_params(self):
        """fp32 masters <- current low-precision shards, every sub-group (reference stage3.py:3106)."""
        for idx, master in enumerate(self.fp32_partitioned_groups_flat):
            swapped = self.optimizer_swapper is not None and master.data.numel() == 0
            if swapped:
                self.optimizer_swapper.swap_in_optimizer_state(parameter=master)
            with self._lp_arena(idx, load=True, dirty=False) as lp:
                master.data.copy_(lp)
            if swapped:
                self.optimizer_swapper.swap_out_optimizer_state(parameter=master)

    def _rebuild_inner_param_groups(self):
        """The inner optimizer sees one flat fp32 parameter per sub-group (hyper-parameters copied from its group)."""
        old = self.optimizer.param_groups
        new_groups = []
        for idx, master in enumerate(self.fp32_partitioned_groups_flat):
            g = {k: v for k, v in old[self.sub_group_to_group_id[idx]].items() if k != "params"}
            g["params"] = [master]
            new_groups.append(g)
        self.optimizer.param_groups = new_groups

    def _persist_small_params(self):
        for p in self.module.parameters():
            if p.ds_persist and p.ds_status == ZeroParamStatus.NOT_AVAILABLE:
                p.all_gather()

    # ------------------------------------------------------------------ module hooks
    def _setup_module_hooks(self):
        self._module_id = 0
        self.hook_handles = []

        def register(module):
            module.id = self._module_id
            self.param_coordinator.module_by_id[module.id] = module
            self._module_id += 1
            _init_external_params(module)
            w = getattr(module, "weight", None)
            if type(module) is torch.nn.Linear and getattr(w, "ds_fused_ag", False):
                # route this Linear through the ZeRO-3 linear function, which runs the fused all-gather + GEMM while
                # the weight is partitioned (F.linear would see the 1-element placeholder)
                import types
                from .linear import FusedAGLinearFunction, WeightHolder
                module._ds_weight_holder = WeightHolder(w)

                def _fused_linear_forward(mod, inp):
                    wt = dict.__getitem__(mod._parameters, "weight")   # no auto-gather
                    if wt.numel() != wt.ds_numel and inp.is_cuda:
                        h = mod._ds_weight_holder
                        return FusedAGLinearFunction.apply(inp, mod.bias, h.dummy, h)
                    return torch.nn.functional.linear(inp, wt, mod.bias)
                module.forward = types.MethodType(_fused_linear_forward, module)
            self.hook_handles.append(module.register_forward_pre_hook(self._pre_forward))
            self.hook_handles.append(module.register_forward_hook(self._post_forward))
            for child in module.children():
                register(child)

        register(self.module)

    def _pre_forward(self, module, *args):
        FWD_MODULE_STACK.append(module)
        pc = self.param_coordinator
        if module is self.module:
            pc.trace.begin_pass()
        pc.fused_weights_stay_partitioned = True
        pc.fetch_sub_module(module, forward=True)
        pc.prefetch_next_sub_modules(module, numel=self.prefetch_elements)

    def _post_forward(self, module, inputs, output):
        if FWD_MODULE_STACK and FWD_MODULE_STACK[-1] is module:
            FWD_MODULE_STACK.pop()
        self._hand_returned_params_to_caller(module, output)
        self._arm_backward(module, inputs, output)
        self.param_coordinator.release_sub_module(module)

    def _arm_backward(self, module, inputs, output):
        """Gradient hooks for this forward call (see _BackwardOfOneCall). torch's module backward hooks are not used:
        they re-wrap every output in an autograd function, which forbids in-place updates of a sub-module's output
        (`h += x`, ReLU(inplace=True)), strips the ds attributes from a returned parameter and does not see outputs
        that are dicts or result objects. Reference: stage3.py:1448-1497 wraps outputs in Pre/PostBackwardFunction."""
        if not torch.is_grad_enabled():
            return
        outs = [t for t in _tensors_in(output) if t.requires_grad and t.grad_fn is not None]
        if not outs:
            return
        call = _BackwardOfOneCall(self, module)
        for t in outs:
            t.register_hook(call.before)
        for t in _tensors_in(inputs):
            if t.requires_grad and not is_zero_param(t):
                t.register_hook(call.after)

    def _hand_returned_params_to_caller(self, module, output):
        """A module that returns one of its parameters (`return out, self.bias`, the skip-bias-add idiom) leaves the
        caller holding a tensor this module is about to release. The parameter becomes an external parameter of the
        calling module instead - registered there once, so from the next iteration on the caller gathers it itself -
        and takes a hold on it that the caller's own release drops; a hold the returning module had on it as ITS
        external parameter moves up with it (reference stage3.py:1413-1446, same outcome)."""
        if not FWD_MODULE_STACK:
            return                      # returned by the root module: nobody above needs it kept
        caller = FWD_MODULE_STACK[-1]
        for p in _zero_params_in(output):
            if any(id(p) in getattr(m, "_external_params", ()) for m in FWD_MODULE_STACK):
                continue                # an enclosing module already looks after it
            register_external_parameter(caller, p)
            p.ds_active_sub_modules += 1
            if id(p) in getattr(module, "_external_params", ()):
                unregister_external_parameter(module, p)
                p.ds_active_sub_modules = max(p.ds_active_sub_modules - 1, 0)
            if p.ds_status == ZeroParamStatus.NOT_AVAILABLE:
                p.all_gather()
                self.param_coordinator.total_available_parameter_numel += p.ds_numel

    def _pre_backward(self, module):
        # weights on the fused path stay partitioned in backward as well (K-sharded all-gather + GEMM for the input
        # gradient, GEMM + reduce-scatter for the weight gradient); DSB_ZERO3_FUSED_DGRAD=0 gathers them with NCCL
        self.param_coordinator.fused_weights_stay_partitioned = self._fused_dgrad
        self.param_coordinator.fetch_sub_module(module, forward=self._fused_dgrad)
        self.param_coordinator.prefetch_next_sub_modules(module, numel=self.prefetch_elements)    # backward looks ahead too

    def _post_backward(self, module):
        # trainable parameters are released from their own post-accumulate-grad hook (AccumulateGrad needs them at
        # full shape and may run after this); modules without a differentiable input never get here - whatever they
        # leave gathered goes back at the end of backward()
        for p in self.param_coordinator._params_of(module):
            p.ds_active_sub_modules = max(p.ds_active_sub_modules - 1, 0)
            if not p.requires_grad:
                self.param_coordinator.release_and_reset_parameter(p)

    # ------------------------------------------------------------------ gradients
    def _install_grad_hooks(self):
        for params in self.sub_groups:
            for p in params:
                hook = self._make_grad_hook(p)
                p.ds_grad_hook = hook   # also called directly by FusedAGLinearFunction.backward
                p.register_post_accumulate_grad_hook(hook)
        self._setup_fused_wgrad_reduce_scatter()

    @property
    def _fused_dgrad(self):
        # measured (profiles/ag_gemm_bench_*): the K-sharded kernel beats all-gather + GEMM between two GPUs and
        # trails it (0.86 - 0.98 x) from four on, where every k-panel waits on a different peer; an explicit
        # DSB_ZERO3_FUSED_DGRAD=1 / 0 overrides the choice
        choice = os.environ.get("DSB_ZERO3_FUSED_DGRAD")
        return self.partition_count <= 2 if choice is None else choice == "1"

    def _setup_fused_wgrad_reduce_scatter(self):
        """Weights on the fused all-gather + GEMM path also get their gradient reduce-scattered INSIDE the wgrad GEMM
        (parallel/ag_gemm.py:GemmReduceScatter): the kernel hands back this rank's averaged shard, which is added
        to the gradient partition without any NCCL call. DSB_ZERO3_FUSED_RS=0 keeps the NCCL reduce-scatter."""
        self._gemm_rs = None
        fused = [p for params in self.sub_groups for p in params if getattr(p, "ds_fused_ag", False)]
        if not fused or self.partition_count < 2 or os.environ.get("DSB_ZERO3_FUSED_RS", "1") != "1":
            return
        from ...parallel.ag_gemm import GemmReduceScatter
        try:
            self._gemm_rs = GemmReduceScatter(max(p.ds_numel for p in fused), dtype=fused[0].dtype,
                                              group=self.dp_process_group)
        except Exception as e:   # no peer mapping: stay on NCCL, loudly
            logger.warning(f"[zero3] fused wgrad reduce-scatter unavailable ({e!r}); using NCCL reduce_scatter")
            return
        for p in fused:
            if p.dtype != fused[0].dtype:
                continue
            p.ds_gemm_rs = self._gemm_rs
            p.ds_reduced_grad_hook = self._make_reduced_grad_hook(p)

    def _make_reduced_grad_hook(self, param):
        def hook(shard):
            """`shard`: this rank's rows of the weight gradient, already summed over the ranks and averaged."""
            idx, off = self._param_slot[param]
            dst = self.grad_partitions_flat[idx].narrow(0, off, shard.numel())
            dst.add_(shard.view(-1).to(dst.dtype))     # accumulates over micro-batches
            param.grad = None
            self.param_coordinator.release_and_reset_parameter(param)
        return hook

    def _make_grad_hook(self, param):
        init = Init.__new__(Init)   # size helpers only

        def hook(p):
            if p.grad is None:
                return
            world = self.partition_count
            idx, off = self._param_slot[param]
            part = param.ds_tensor.ds_numel       # (numel() is 0 while the shard is swapped out)
            dst = self.grad_partitions_flat[idx].narrow(0, off, part)
            g = p.grad
            if g.numel() != param.ds_numel:
                p.grad = None
                return
            full = torch.zeros(part * world, dtype=g.dtype, device=g.device)
            full.narrow(0, 0, param.ds_numel).copy_(g.contiguous().view(-1))
            if world > 1:
                out = torch.empty(part, dtype=g.dtype, device=g.device)
                if g.is_cuda:
                    dist.reduce_scatter_tensor(out, full, op=dist.ReduceOp.AVG, group=self.dp_process_group)
                else:
                    dist.reduce_scatter_tensor(out, full, group=self.dp_process_group)
                    out.div_(world)
            else:
                out = full
            dst.add_(out.to(dst.dtype))     # accumulates over micro-batches
            p.grad = None
            self.param_coordinator.release_and_reset_parameter(param)
        return hook

    def zero_grad(self, set_grads_to_None=True):
        for g in self.grad_partitions_flat:
            g.zero_()
        for params in self.sub_groups:
            for p in params:
                p.grad = None

    def overlapping_partition_gradients_reduce_epilogue(self):
        pass   # reduce-scatters are issued from the per-parameter hooks as gradients appear

    def reduce_gradients(self):
        pass

    @property
    def loss_scale(self):
        return self.loss_scaler.loss_scale

    @property
    def cur_scale(self):
        return self.loss_scaler.cur_scale

    @cur_scale.setter
    def cur_scale(self, v):
        self.loss_scaler.cur_scale = v

    def backward(self, loss, retain_graph=False):
        self.micro_step_id += 1
        scaled = loss.float() * self.loss_scale
        scaled.backward(retain_graph=retain_graph)
        self.end_of_pass()

    def end_of_pass(self):
        """After backward (or after a forward pass that will have no backward): whatever is still gathered - unused
        in this backward, or kept for a backward that is not coming - goes back to shards; the trace starts over."""
        for p in self.module.parameters():
            if not p.ds_persist and p.ds_status != ZeroParamStatus.NOT_AVAILABLE:
                self.param_coordinator.release_and_reset_parameter(p)
        self.param_coordinator.reset_step()

    # ------------------------------------------------------------------ step
    def _norm_and_overflow(self):
        self._sumsq.zero_()
        self._flag.zero_()
        acc = self._sumsq.narrow(0, 0, 1)
        for g in self.grad_partitions_flat:
            K.sumsq(g, out=acc, flag=self._flag, accumulate=True)
        self._sumsq[1] = self._flag[0].float()
        if _dist_ready():
            if self.partition_count > 1:
                dist.all_reduce(self._sumsq, group=self.dp_process_group)
            if self.mpu is not None and self.mpu.get_model_parallel_world_size() > 1:
                dist.all_reduce(self._sumsq, group=self.mpu.get_model_parallel_group())
        vals = self._sumsq.tolist()
        norm = vals[0] ** 0.5 if vals[0] == vals[0] and vals[0] >= 0 else float("nan")
        overflow = vals[1] > 0 or norm != norm or norm == float("inf")
        return norm, overflow

    def _combined_scale(self, norm):
        scale = float(self.loss_scale)
        combined = scale
        if self.clip_grad > 0.0:
            clip = (norm / scale + 1e-6) / self.clip_grad
            if clip > 1:
                combined = clip * scale
        return combined

    def _optimizer_step(self, idx, inv):
        # every element is rewritten from the master: no read needed; the file write of this sub-group runs under the
        # next sub-group's step (the swapper waits for it before the buffer is reused or the file is read)
        with self._lp_arena(idx, async_write=True) as lp:
            self._optimizer_step_into(idx, inv, lp)

    def _optimizer_step_into(self, idx, inv, lp):
        master = self.fp32_partitioned_groups_flat[idx]
        group = self.optimizer.param_groups[idx]
        g = self.grad_partitions_flat[idx]
        if self.fused_inner:
            st = self.optimizer.state[master]
            if len(st) == 0:
                st["exp_avg"] = torch.zeros_like(master.data)
                st["exp_avg_sq"] = torch.zeros_like(master.data)
            self.optimizer.step_flat(group, master.data, st["exp_avg"], st["exp_avg_sq"], g, lp=lp, inv_scale=inv)
            return
        g32 = g.float().mul_(inv)
        master.grad = g32.to(master.device, non_blocking=True)
        if master.device.type == "cpu" and torch.cuda.is_available():
            torch.cuda.current_stream().synchronize()
        # step only this sub-group's parameter (state of the others may be swapped out)
        saved = self.optimizer.param_groups
        self.optimizer.param_groups = [group]
        from ...ops.adam import DeepSpeedCPUAdam
        if isinstance(self.optimizer, DeepSpeedCPUAdam) and lp.is_cuda and lp.dtype in (torch.float16, torch.bfloat16):
            self.optimizer.step(fp16_param_groups=[[lp]])
        else:
            self.optimizer.step()
            lp.copy_(master.data, non_blocking=True)
        self.optimizer.param_groups = saved
        master.grad = None

    def step(self, closure=None):
        norm, overflow = self._norm_and_overflow()
        self.overflow = overflow
        prev = self.loss_scale
        if overflow:
            self.loss_scaler.update_scale(True)
            self._global_grad_norm = -1
            logger.info("[deepspeed] OVERFLOW! Rank {} Skipping step. Attempted loss scale: {}, reducing to {}".format(
                dist.get_rank() if _dist_ready() else 0, prev, self.loss_scale))
            self.zero_grad()
            return self.overflow
        self._global_grad_norm = norm / float(prev)
        inv = 1.0 / self._combined_scale(norm)
        for idx in range(len(self.sub_groups)):
            if self.optimizer_swapper is not None:
                self.optimizer_swapper.swap_in_optimizer_state(parameter=self.fp32_partitioned_groups_flat[idx])
            self._optimizer_step(idx, inv)
            if self.optimizer_swapper is not None:
                self.optimizer_swapper.swap_out_optimizer_state(parameter=self.fp32_partitioned_groups_flat[idx])
        if self.param_swapper is not None:
            self.param_swapper.flush()          # the last sub-group's file write
        self.loss_scaler.update_scale(False)
        self.zero_grad()
        # persistent (always-gathered) params must see their updated values
        for p in self.module.parameters():
            if p.ds_persist:
                if p.ds_status == ZeroParamStatus.AVAILABLE:
                    p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                p.all_gather()
        self._peer_shard_fence()
        return self.overflow

    def _peer_shard_fence(self):
        """Fused all-gather + GEMM forwards read the peers' shards directly: every rank's shard update must have
        executed before anyone's next forward (NCCL all-gathers gave that ordering for free)."""
        fused = [p for p in self.module.parameters() if getattr(p, "ds_fused_ag", False)]
        if not fused:
            return
        for p in fused:   # the optimizer keeps the shard in its flat buffer: refresh the peer-visible copy
            if p.ds_heap_shard.data_ptr() != p.ds_tensor.data_ptr():
                p.ds_heap_shard.copy_(p.ds_tensor)
        if _dist_ready():
            dist.barrier(group=self.dp_process_group)

    def get_global_grad_norm(self):
        return self._global_grad_norm

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @property
    def state(self):
        return self.optimizer.state

    # ------------------------------------------------------------------ checkpoints (another DP degree re-shards on load)
    def save_checkpoint_prologue(self):
        for p in self.module.parameters():
            self.param_coordinator.release_and_reset_parameter(p)
            if p.ds_persist and p.ds_status == ZeroParamStatus.AVAILABLE:
                p.ds_persist = False
                p.partition()
                p.ds_persist = True

    def save_checkpoint_epilogue(self):
        self._persist_small_params()

    def state_dict(self):
        if self.optimizer_swapper is not None:
            masters, inner = self._swapped_state_for_checkpoint()
        else:
            masters, inner = [m.data.clone() for m in self.fp32_partitioned_groups_flat], self.optimizer.state_dict()
        return {
            "zero_stage": 3,
            "loss_scaler": self.loss_scaler,
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "overflow": self.overflow,
            "partition_count": self.partition_count,
            "optimizer_state_dict": inner,
            "fp32_flat_groups": masters,
            # what a loader with a different data-parallel degree needs to cut the arenas apart again
            "sub_group_numels": [[p.ds_numel for p in params] for params in self.sub_groups],
        }

    def _swapped_state_for_checkpoint(self):
        """Optimizer state on NVMe (the reference refuses to checkpoint then, stage3.py:3030): one sub-group at a time
        is read in, copied into ordinary host memory for the checkpoint and dropped again; its files stay as they
        are."""
        position, n = {}, 0
        for group in self.optimizer.param_groups:
            for p in group["params"]:
                position[id(p)] = n
                n += 1
        masters, states = [], {}
        for master in self.fp32_partitioned_groups_flat:
            self.optimizer_swapper.swap_in_optimizer_state(parameter=master)
            masters.append(master.data.clone())
            kept = {k: (v.detach().clone() if torch.is_tensor(v) else v)
                    for k, v in self.optimizer.state.get(master, {}).items()}
            if kept:
                states[position[id(master)]] = kept
            self.optimizer_swapper.release_optimizer_state(parameter=master)
        inner = self.optimizer.state_dict()
        inner["state"] = states
        return masters, inner

    def load_state_dict(self, state_dict_list, load_optimizer_states=True, load_from_fp32_weights=False):
        if isinstance(state_dict_list, dict):
            state_dict_list = [state_dict_list]
        if state_dict_list[0]["partition_count"] != self.partition_count:
            sd = self._reshard_for_this_rank(state_dict_list)
        else:
            sd = state_dict_list[min(self.dp_rank, len(state_dict_list) - 1)]
        self.loss_scaler = sd["loss_scaler"]
        self.dynamic_loss_scale = sd["dynamic_loss_scale"]
        self.overflow = sd["overflow"]
        swapper = self.optimizer_swapper
        if swapper is not None:
            for master in self.fp32_partitioned_groups_flat:
                if load_optimizer_states:
                    swapper.forget(master)          # the loaded state tensors replace the tracked ones
                else:
                    swapper.swap_in_optimizer_state(parameter=master)    # keep the moments, replace the master
        if load_optimizer_states:
            self.optimizer.load_state_dict(sd["optimizer_state_dict"])
        for idx, (cur, saved) in enumerate(zip(self.fp32_partitioned_groups_flat, sd["fp32_flat_groups"])):
            if cur.data.numel() == saved.numel():
                cur.data.copy_(saved.data)
            else:                                   # swapped out: no storage behind it at the moment
                cur.data = saved.data.to(torch.float32).clone()
            with self._lp_arena(idx) as lp:
                lp.copy_(cur.data)
            if swapper is not None:
                swapper.swap_out_optimizer_state(parameter=cur)
        for p in self.module.parameters():
            if p.ds_persist:
                p.ds_status = ZeroParamStatus.NOT_AVAILABLE
                p.all_gather()
        self._peer_shard_fence()

    def _reshard_for_this_rank(self, shards):
        """A checkpoint written with another data-parallel degree (the reference's ZeRO-3 checkpoints are rigid,
        stage3.py:3072): every saved arena is cut back into its parameters' shards, the shards of all saved ranks are
        joined into the full fp32 vectors (master and every per-element optimizer state), and those are cut again
        for this rank's arenas. Returns a state dict in the layout `load_state_dict` expects from a same-degree save."""
        saved_world = shards[0]["partition_count"]
        assert len(shards) == saved_world, f"need all {saved_world} ZeRO shards of the checkpoint, got {len(shards)}"
        layout = shards[0].get("sub_group_numels")
        assert layout is not None, "this ZeRO-3 checkpoint predates re-shardable saves: load it with the same DP degree"
        saved_numels = [n for group in layout for n in group]
        mine = [p.ds_numel for params in self.sub_groups for p in params]
        assert saved_numels == mine, "the checkpoint belongs to a different model (parameter sizes differ)"

        def tensors_of(shard):
            """Per saved sub-group: {name: flat tensor} of the master and the per-element optimizer states."""
            out = []
            for g, master in enumerate(shard["fp32_flat_groups"]):
                state = shard["optimizer_state_dict"]["state"].get(g, {})
                per = {"": master}
                per.update({k: v for k, v in state.items() if torch.is_tensor(v) and v.numel() == master.numel()})
                out.append(per)
            return out

        per_rank = [tensors_of(shard) for shard in shards]
        names = sorted(set().union(*[set(g) for g in per_rank[0]]))
        full = {name: [] for name in names}          # name -> full vectors, one per parameter, in parameter order
        for g, numels in enumerate(layout):
            offset = 0
            for n in numels:
                part = -(-n // saved_world)
                for name in names:
                    if name in per_rank[0][g]:
                        pieces = [per_rank[r][g][name].narrow(0, offset, part) for r in range(saved_world)]
                        full[name].append(torch.cat(pieces)[:n].float())
                    else:
                        full[name].append(None)
                offset += part
        world, rank = self.partition_count, self.dp_rank
        masters, states, k = [], {}, 0
        scalars = shards[0]["optimizer_state_dict"]["state"]
        for g, params in enumerate(self.sub_groups):
            arenas = {name: [] for name in names}
            for p in params:
                part = -(-p.ds_numel // world)
                for name in names:
                    vec = full[name][k]
                    if vec is None:
                        continue
                    padded = torch.zeros(part * world, dtype=torch.float32)
                    padded[:p.ds_numel] = vec
                    arenas[name].append(padded.narrow(0, rank * part, part))
                k += 1
            masters.append(torch.cat(arenas[""]))
            state = {name: torch.cat(parts) for name, parts in arenas.items() if name and parts}
            if state:
                # scalar entries (the step count) are the same in every saved sub-group
                state.update({key: v for key, v in next(iter(scalars.values())).items() if not torch.is_tensor(v)
                              or v.numel() == 1})
                states[g] = state
        inner = {"state": states, "param_groups": self.optimizer.state_dict()["param_groups"]}
        saved_groups = shards[0]["optimizer_state_dict"]["param_groups"]
        for group in inner["param_groups"]:          # hyper-parameters (lr after scheduling, betas, ...) as saved
            group.update({key: v for key, v in saved_groups[0].items() if key != "params"})
        out = dict(shards[0])
        out.update(partition_count=world, fp32_flat_groups=masters, optimizer_state_dict=inner)
        return out

    def __repr__(self):
        return repr(self.optimizer)
