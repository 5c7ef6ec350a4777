"""Activation checkpointing with a model-parallel RNG tracker.

API and options per reference deepspeed/runtime/activation_checkpointing/checkpointing.py:122-861:
`checkpoint(fn, *args)`, `configure(...)`, `get_cuda_rng_tracker()`, `model_parallel_cuda_manual_seed(seed)`,
`partition_activations` (each tensor-parallel rank keeps 1/mp of every checkpointed input and the group all-gathers it
for the recompute), `cpu_checkpointing` (the kept copy lives in host memory), `contiguous_memory_optimization` (the
kept slices of all checkpoints of a pass are carved from one arena per argument position),
`synchronize_checkpoint_boundary`, `profile`. Works on CPU tensors too (CI).

Mechanism: every tensor argument of a checkpoint is handed to a *stash* (`_KeepStash`, `_HostStash`, `_SliceStash`)
that decides what is kept between forward and backward and how the full tensor comes back; shapes are plain tuples in
the autograd context, so there are no (tensor, size-tensor) pairs to interleave and re-merge. RNG capture / replay is
one context manager shared by both passes.
"""
import contextlib
import copy

import torch
import torch.distributed as dist

from ...utils.logging import logger
from ...utils.timer import SynchronizedWallClockTimer as Timers
from ..config import DeepSpeedConfig

# ---- module state (names are part of the interface: Megatron-style code reads and sets some of them)
mpu = None
mp_rank = None
mp_size = None
mp_group = None
num_layers = None
contiguous_data_buffers = []     # one arena per argument position (contiguous_memory_optimization)
data_offsets = []
contiguous_size_buffers = []     # unused here (shapes are tuples); kept so that callers clearing them keep working
size_offsets = []
timers = None

PARTITION_ACTIVATIONS = False
PA_TO_CPU = False
CONTIGUOUS_CHECKPOINTING = False
SYNCHRONIZE = False
PROFILE_TIME = False
deepspeed_checkpointing_enabled = False
transport_stream = None
cuda_device = None

_MODEL_PARALLEL_RNG_TRACKER_NAME = "model-parallel-rng"


def _cuda():
    return torch.cuda.is_available()


# ----------------------------------------------------------------------------------------------- RNG
def _set_cuda_rng_state(new_state, device=-1):
    if not _cuda():
        return
    if device == -1:
        index = torch.cuda.current_device()
    else:
        dev = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        index = dev.index if dev.index is not None else torch.cuda.current_device()
    torch.cuda.default_generators[index].set_state(new_state)


def _device_rng_get():
    return torch.cuda.get_rng_state() if _cuda() else torch.get_rng_state()


def _device_rng_set(state):
    if _cuda():
        _set_cuda_rng_state(state)
    else:
        torch.set_rng_state(state)


class CudaRNGStatesTracker:
    """Named generator states. `fork(name)` runs a block under the named state and stores the advanced state back, so
    dropout inside tensor-parallel regions differs between TP ranks while everything outside stays replicated."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def get_states(self):
        return copy.copy(self.states_)

    def set_states(self, states):
        self.states_ = states

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception("seed {} already exists".format(seed))
        if name in self.states_:
            raise Exception("cuda rng state {} already exists".format(name))
        self.seeds_.add(seed)
        outside = _device_rng_get()
        (torch.cuda.manual_seed if _cuda() else torch.manual_seed)(seed)
        self.states_[name] = _device_rng_get()
        _device_rng_set(outside)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception("cuda rng state {} is not added".format(name))
        outside = _device_rng_get()
        _device_rng_set(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = _device_rng_get()
            _device_rng_set(outside)


_CUDA_RNG_STATE_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _CUDA_RNG_STATE_TRACKER


def model_parallel_cuda_manual_seed(seed):
    """Default generator: `seed` on every TP rank; tracker state 'model-parallel-rng': seed + 2718 + tp_rank."""
    tp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
    tp_seed = seed + 2718 + tp_rank
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info("> initializing model parallel cuda seeds on global rank {}, model parallel rank {}: "
                    "model parallel seed: {} and data parallel seed: {}".format(dist.get_rank(), tp_rank, tp_seed, seed))
    _CUDA_RNG_STATE_TRACKER.reset()
    (torch.cuda.manual_seed if _cuda() else torch.manual_seed)(seed)
    _CUDA_RNG_STATE_TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, tp_seed)


def _capture_rng():
    return (torch.get_rng_state(), torch.cuda.get_rng_state() if _cuda() else None,
            get_cuda_rng_tracker().get_states())


def _restore_rng(snapshot):
    cpu_state, cuda_state, tracker = snapshot
    torch.set_rng_state(cpu_state)
    if cuda_state is not None:
        _set_cuda_rng_state(cuda_state)
    get_cuda_rng_tracker().set_states(tracker)


@contextlib.contextmanager
def _replayed_rng(snapshot):
    """Run a block under the generator states captured in forward, then continue from where backward was."""
    current = _capture_rng()
    _restore_rng(snapshot)
    try:
        yield
    finally:
        _restore_rng(current)


# ----------------------------------------------------------------------------------------------- helpers (interface)
def detach_variable(inputs, device=None):
    """Detached copies (optionally moved to `device`) that keep each tensor's requires_grad."""
    if not isinstance(inputs, tuple):
        raise RuntimeError("Only tuple of tensors is supported. Got Unsupported input type: ", type(inputs).__name__)
    out = []
    for item in inputs:
        if isinstance(item, torch.Tensor):
            moved = item if device is None else item.to(device=device)
            item = moved.detach().requires_grad_(item.requires_grad)
        out.append(item)
    return tuple(out)


def move_to_device(item, device):
    if device is None:
        return item
    if torch.is_tensor(item):
        return item.to(device)
    if isinstance(item, (list, tuple)):
        return type(item)(move_to_device(v, device) for v in item)
    return item


def extract_tensors(all_objects):
    """(tensors, non-tensors, is-tensor flags) of a tuple or list, preserving order within each class."""
    flags = [torch.is_tensor(v) for v in all_objects]
    tensors = [v for v, f in zip(all_objects, flags) if f]
    others = [v for v, f in zip(all_objects, flags) if not f]
    wrap = tuple if type(all_objects) is tuple else list
    return wrap(tensors), wrap(others), wrap(flags)


def merge_tensors(tensor_objects, non_tensor_objects, tensor_flags):
    """Inverse of extract_tensors."""
    tensors, others = iter(tensor_objects), iter(non_tensor_objects)
    return tuple(next(tensors) if f else next(others) for f in tensor_flags)


def get_partition_size(item):
    return item.numel() // mp_size


def get_partition_start(item):
    return get_partition_size(item) * mp_rank


def _lazy_topology():
    """Model-parallel coordinates and the side stream are resolved at first use (configure() may precede mpu set-up)."""
    global mp_rank, mp_size, mp_group, cuda_device, transport_stream
    if mp_rank is None:
        if mpu is not None:
            mp_rank, mp_size = mpu.get_model_parallel_rank(), mpu.get_model_parallel_world_size()
            mp_group = mpu.get_model_parallel_group()
        else:
            mp_rank, mp_size, mp_group = 0, 1, None
    if cuda_device is None and _cuda():
        cuda_device = torch.cuda.current_device()
        if PARTITION_ACTIVATIONS or PA_TO_CPU:
            transport_stream = torch.cuda.Stream(device=cuda_device)


def get_full_inputs(tensors, device=None, fp32_comm=False):
    """Interface helper: `tensors` = (slice_0, shape_0, ..., slice_k, shape_k, last, last_shape); every slice is
    all-gathered over the model-parallel group back into its full tensor (in place: slice.data becomes the full data)."""
    out = []
    for i in range(len(tensors) // 2 - 1):
        piece, shape = tensors[2 * i], tensors[2 * i + 1]
        shape = [int(v) for v in (shape.tolist() if torch.is_tensor(shape) else shape)]
        piece.data = _gather_slices(piece, shape, device, fp32_comm).data
        out.append(piece)
    out.append(tensors[-2])
    return tuple(out)


def _gather_slices(piece, shape, device=None, fp32_comm=False):
    """Full tensor of `shape` from this rank's flat slice and the slices of the other model-parallel ranks."""
    dev = device if device is not None else piece.device
    n = piece.numel()
    wire = torch.float32 if fp32_comm else piece.dtype
    flat = torch.empty(n * mp_size, dtype=wire, device=dev)
    mine = flat.narrow(0, n * mp_rank, n)
    mine.copy_(piece.to(dev))
    if mp_group is not None and mp_size > 1:
        dist.all_gather([flat.narrow(0, n * r, n) for r in range(mp_size)], mine.clone(), group=mp_group)
    return flat.to(piece.dtype).view(shape)


# ----------------------------------------------------------------------------------------------- stashes
class _KeepStash:
    """Keep the tensor itself (plain checkpointing)."""

    def pack(self, index, tensor):
        return tensor, None

    def unpack(self, kept, meta, original):
        return kept


class _HostStash:
    """Keep a host copy; bring it back to the compute device for the recompute."""

    def pack(self, index, tensor):
        return tensor.detach().to("cpu", non_blocking=True), None

    def unpack(self, kept, meta, original):
        return move_to_device(kept, cuda_device)


class _SliceStash:
    """Keep 1/mp of the flattened tensor (optionally on the host, optionally inside a per-argument arena that holds
    the slices of all `num_layers` checkpoints of a pass) and all-gather the rest back in backward. The caller's
    tensor is shrunk to the slice while it waits (that is where the memory saving comes from) and restored before the
    recompute."""

    def __init__(self, to_host, contiguous):
        self.to_host, self.contiguous = to_host, contiguous

    def _arena_view(self, index, piece):
        while len(contiguous_data_buffers) <= index:
            contiguous_data_buffers.append(None)
            data_offsets.append(0)
        if contiguous_data_buffers[index] is None:
            where = torch.device("cpu") if self.to_host else piece.device
            contiguous_data_buffers[index] = torch.empty(num_layers * piece.numel(), dtype=piece.dtype, device=where)
            data_offsets[index] = 0
        view = contiguous_data_buffers[index].narrow(0, data_offsets[index], piece.numel())
        data_offsets[index] += piece.numel()
        return view.copy_(piece)

    def pack(self, index, tensor):
        n = get_partition_size(tensor)
        piece = tensor.detach().contiguous().view(-1).narrow(0, get_partition_start(tensor), n).clone()
        if self.contiguous:
            piece = self._arena_view(index, piece)
        elif self.to_host:
            piece = piece.to("cpu")
        return piece, tuple(tensor.shape)

    def unpack(self, kept, meta, original):
        full = _gather_slices(kept, list(meta), device=cuda_device if self.to_host else None)
        if original is not None:
            original.data = full.data       # un-shrink the caller's tensor
            return original
        return full


def _release_arenas():
    for i in range(len(contiguous_data_buffers)):
        contiguous_data_buffers[i] = None
        data_offsets[i] = 0
    for i in range(len(contiguous_size_buffers)):
        contiguous_size_buffers[i] = None
    size_offsets[:] = [0] * len(size_offsets)


@contextlib.contextmanager
def _boundary(phase):
    """synchronize_checkpoint_boundary / profile around one pass of a checkpoint."""
    global timers
    if SYNCHRONIZE and _cuda():
        torch.cuda.synchronize()
    if PROFILE_TIME:
        if timers is None:
            timers = Timers()
        timers(phase).start()
    try:
        yield
    finally:
        if PROFILE_TIME:
            timers(phase).stop()
            timers.log([phase])
        if SYNCHRONIZE and _cuda():
            torch.cuda.synchronize()


class CheckpointFunction(torch.autograd.Function):
    """Run `run_function` without building a graph, keep (some form of) its inputs, and rebuild the graph in backward
    under the generator states of the forward pass."""

    @staticmethod
    def forward(ctx, run_function, all_outputs, *args):
        with _boundary("forward"):
            _lazy_topology()
            ctx.run_function = run_function
            if PARTITION_ACTIVATIONS:
                stash = _SliceStash(PA_TO_CPU, CONTIGUOUS_CHECKPOINTING)
            else:
                stash = _HostStash() if PA_TO_CPU else _KeepStash()
            keep = _KeepStash()
            plan, kept, shrink = [], [], []
            for i, a in enumerate(args):
                if not torch.is_tensor(a):
                    plan.append(("object", a, None, None))
                    continue
                # with partitioning the LAST argument (Megatron: the attention mask) is always kept whole
                chosen = keep if (PARTITION_ACTIVATIONS and i == len(args) - 1) else stash
                token, meta = chosen.pack(i, a)
                plan.append(("tensor", len(kept), chosen, meta))
                kept.append(token)
                if isinstance(chosen, _SliceStash):
                    shrink.append((a, token))
            ctx.rng = _capture_rng()          # the recompute must draw the same dropout masks
            with torch.no_grad():
                outputs = run_function(*(move_to_device(args, cuda_device) if cuda_device is not None else args))
            for original, piece in shrink:
                original.data = piece.data
            ctx.originals = [a for a, _ in shrink]
            ctx.plan = plan
            ctx.save_for_backward(*kept)

        outs = [outputs] if torch.is_tensor(outputs) else list(outputs)
        ctx.mark_non_differentiable(*[o for o in outs if torch.is_tensor(o) and not o.is_floating_point()])
        all_outputs += outs
        if torch.is_tensor(outputs):
            return outputs
        return tuple(o for o in outs if torch.is_tensor(o))

    @staticmethod
    def backward(ctx, *grads):
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), please use .backward() if possible")
        with _boundary("backward"):
            if CONTIGUOUS_CHECKPOINTING:
                _release_arenas()         # every slice of this pass has been (or is about to be) read back
            kept = ctx.saved_tensors
            originals = iter(ctx.originals)
            inputs = []
            for kind, ref, stash, meta in ctx.plan:
                if kind == "object":
                    inputs.append(ref)
                    continue
                original = next(originals) if isinstance(stash, _SliceStash) else None
                full = stash.unpack(kept[ref], meta, original)
                inputs.append(full.detach().requires_grad_(full.requires_grad))
            with _replayed_rng(ctx.rng), torch.enable_grad():
                outputs = ctx.run_function(*inputs)
            outputs = (outputs,) if isinstance(outputs, torch.Tensor) else outputs
            pairs = [(o, g) for o, g in zip((o for o in outputs if torch.is_tensor(o)), grads) if o.requires_grad]
            torch.autograd.backward([o for o, _ in pairs], [g for _, g in pairs])
        return (None, None) + tuple(x.grad if torch.is_tensor(x) else None for x in inputs)


def checkpoint(function, *args):
    """Checkpoint a model or part of the model: activations inside `function` are recomputed in backward."""
    all_outputs = []
    CheckpointFunction.apply(function, all_outputs, *args)
    return all_outputs[0] if len(all_outputs) == 1 else tuple(all_outputs)


# ----------------------------------------------------------------------------------------------- configuration
def partition_activations_in_checkpoint(partition_activation):
    global PARTITION_ACTIVATIONS
    PARTITION_ACTIVATIONS = partition_activation
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(f"**************Partition Activations {PARTITION_ACTIVATIONS}************")


def set_num_layers(nlayers):
    global num_layers
    num_layers = nlayers


def reset():
    """Give the arenas back between passes that do not run a backward (e.g. between training and evaluation)."""
    if CONTIGUOUS_CHECKPOINTING:
        _release_arenas()


_OPTION_GLOBALS = {"partition_activations": "PARTITION_ACTIVATIONS", "contiguous_checkpointing": "CONTIGUOUS_CHECKPOINTING",
                   "num_checkpoints": "num_layers", "checkpoint_in_cpu": "PA_TO_CPU", "synchronize": "SYNCHRONIZE",
                   "profile": "PROFILE_TIME"}


def _configure_using_config_file(deepspeed_config, mpu=None):
    config = DeepSpeedConfig(deepspeed_config, mpu=mpu).activation_checkpointing_config
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(config.repr())
    globals().update(PARTITION_ACTIVATIONS=config.partition_activations,
                     CONTIGUOUS_CHECKPOINTING=config.contiguous_memory_optimization,
                     num_layers=config.number_checkpoints, PA_TO_CPU=config.cpu_checkpointing,
                     SYNCHRONIZE=config.synchronize_checkpoint_boundary, PROFILE_TIME=config.profile)


def _configure_defaults():
    globals().update(PARTITION_ACTIVATIONS=False, CONTIGUOUS_CHECKPOINTING=False, num_layers=False, PA_TO_CPU=False,
                     SYNCHRONIZE=False, PROFILE_TIME=False, deepspeed_checkpointing_enabled=True)


def configure(mpu_, deepspeed_config=None, partition_activations=None, contiguous_checkpointing=None,
              num_checkpoints=None, checkpoint_in_cpu=None, synchronize=None, profile=None):
    """Defaults, then the `activation_checkpointing` block of a config file, then explicit arguments."""
    global mpu
    _configure_defaults()
    if mpu_ is not None:
        mpu = mpu_
    if deepspeed_config is not None:
        _configure_using_config_file(deepspeed_config, mpu=mpu)
    given = dict(partition_activations=partition_activations, contiguous_checkpointing=contiguous_checkpointing,
                 num_checkpoints=num_checkpoints, checkpoint_in_cpu=checkpoint_in_cpu, synchronize=synchronize,
                 profile=profile)
    globals().update({_OPTION_GLOBALS[k]: v for k, v in given.items() if v is not None})
    if PA_TO_CPU or CONTIGUOUS_CHECKPOINTING:
        assert PARTITION_ACTIVATIONS, "CPU Checkpointing/Contiguous Checkpointing is only available with partitioned " \
            "activations. Set partitioned activations to true in deepspeed config"
    if CONTIGUOUS_CHECKPOINTING:
        assert num_layers is not None, "Must specify the number of layers with contiguous memory checkpointing"


def is_configured():
    return deepspeed_checkpointing_enabled
