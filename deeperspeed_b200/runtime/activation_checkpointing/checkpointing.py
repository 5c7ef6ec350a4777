"""Activation checkpointing with a model-parallel RNG tracker.

API and options per reference deepspeed/runtime/activation_checkpointing/checkpointing.py:122-861:
`checkpoint(fn, *args)`, `configure(...)`, `get_cuda_rng_tracker()`, `model_parallel_cuda_manual_seed(seed)`,
partition_activations (each TP rank keeps 1/mp of every saved activation, all-gathered in backward),
cpu_checkpointing (saved activations live in pinned host memory, copied on a side stream),
contiguous_memory_optimization, synchronize_checkpoint_boundary, profile.  Works on CPU tensors too (CI).
"""
import contextlib
import copy

import torch
import torch.distributed as dist

from ...utils.logging import logger
from ...utils.timer import SynchronizedWallClockTimer as Timers
from ..config import DeepSpeedConfig

mpu = None
mp_rank = None
mp_size = None
mp_group = None
num_layers = None
contiguous_data_buffers = []
data_offsets = []
contiguous_size_buffers = []
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


def detach_variable(inputs, device=None):
    if isinstance(inputs, tuple):
        out = []
        for inp in inputs:
            if not isinstance(inp, torch.Tensor):
                out.append(inp)
                continue
            requires_grad = inp.requires_grad
            x = inp.to(device=device) if device is not None else inp
            x = x.detach()
            x.requires_grad = requires_grad
            out.append(x)
        return tuple(out)
    raise RuntimeError("Only tuple of tensors is supported. Got Unsupported input type: ", type(inputs).__name__)


def _set_cuda_rng_state(new_state, device=-1):
    if not _cuda():
        return
    if device == -1:
        device = torch.device("cuda")
    elif isinstance(device, str):
        device = torch.device(device)
    elif isinstance(device, int):
        device = torch.device("cuda", device)
    idx = device.index if device.index is not None else torch.cuda.current_device()
    torch.cuda.default_generators[idx].set_state(new_state)


class CudaRNGStatesTracker:
    """Named RNG states so dropout inside tensor-parallel regions differs across TP ranks while everything else is
    replicated; `fork(name)` runs a block under that state and stores the advanced state back."""

    def __init__(self):
        self.states_ = {}
        self.seeds_ = set()

    def reset(self):
        self.states_ = {}
        self.seeds_ = set()

    def get_states(self):
        return copy.copy(self.states_)

    def set_states(self, states):
        self.states_ = states

    def _get(self):
        return torch.cuda.get_rng_state() if _cuda() else torch.get_rng_state()

    def _set(self, s):
        if _cuda():
            _set_cuda_rng_state(s)
        else:
            torch.set_rng_state(s)

    def add(self, name, seed):
        if seed in self.seeds_:
            raise Exception("seed {} already exists".format(seed))
        self.seeds_.add(seed)
        if name in self.states_:
            raise Exception("cuda rng state {} already exists".format(name))
        orig = self._get()
        if _cuda():
            torch.cuda.manual_seed(seed)
        else:
            torch.manual_seed(seed)
        self.states_[name] = self._get()
        self._set(orig)

    @contextlib.contextmanager
    def fork(self, name=_MODEL_PARALLEL_RNG_TRACKER_NAME):
        if name not in self.states_:
            raise Exception("cuda rng state {} is not added".format(name))
        orig = self._get()
        self._set(self.states_[name])
        try:
            yield
        finally:
            self.states_[name] = self._get()
            self._set(orig)


_CUDA_RNG_STATE_TRACKER = CudaRNGStatesTracker()


def get_cuda_rng_tracker():
    return _CUDA_RNG_STATE_TRACKER


def model_parallel_cuda_manual_seed(seed):
    """Default state: same seed on all TP ranks (data-parallel replicas of one TP group share dropout outside TP
    regions); tracker state: seed + 2718 + tp_rank."""
    global mpu
    tp_rank = mpu.get_model_parallel_rank() if mpu is not None else 0
    offset = seed + 2718
    model_parallel_seed = offset + tp_rank
    data_parallel_seed = seed
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info("> initializing model parallel cuda seeds on global rank {}, model parallel rank {}: "
                    "model parallel seed: {} and data parallel seed: {}".format(dist.get_rank(), tp_rank,
                                                                                model_parallel_seed, data_parallel_seed))
    _CUDA_RNG_STATE_TRACKER.reset()
    if _cuda():
        torch.cuda.manual_seed(data_parallel_seed)
    else:
        torch.manual_seed(data_parallel_seed)
    _CUDA_RNG_STATE_TRACKER.add(_MODEL_PARALLEL_RNG_TRACKER_NAME, model_parallel_seed)


def get_partition_start(item):
    global mp_rank, mp_size
    size = item.numel()
    return int(size / mp_size) * mp_rank


def get_partition_size(item):
    global mp_size
    return int(item.numel() / mp_size)


def get_full_inputs(tensors, device=None, fp32_comm=False):
    """Rebuild full activations from (partition, size) pairs by all-gather over the TP group."""
    inputs = []
    num_args = int(len(tensors) / 2)
    for i in range(num_args - 1):
        item = tensors[2 * i]
        size = tensors[2 * i + 1]
        partition_size = item.numel()
        tensor_size = partition_size * mp_size
        dev = device if device is not None else item.device
        wire_dtype = torch.float32 if fp32_comm else item.dtype
        flat = torch.zeros([tensor_size], dtype=wire_dtype, device=dev)
        parts = [flat.narrow(0, partition_size * r, partition_size) for r in range(mp_size)]
        parts[mp_rank].copy_(item.to(dev))
        if mp_group is not None and mp_size > 1:
            dist.all_gather(parts, parts[mp_rank].clone(), group=mp_group)
        full = flat.to(item.dtype).view(list(size.numpy() if hasattr(size, "numpy") else size))
        item.data = full.data
        inputs.append(item)
    inputs.append(tensors[-2])
    return tuple(inputs)


def extract_tensors(all_objects):
    """Split a tuple into (tensor objects, non-tensor objects, tensor flags)."""
    tensor_objects = [v for v in all_objects if torch.is_tensor(v)]
    non_tensor_objects = [v for v in all_objects if not torch.is_tensor(v)]
    tensor_flags = [torch.is_tensor(v) for v in all_objects]
    if type(all_objects) is tuple:
        return tuple(tensor_objects), tuple(non_tensor_objects), tuple(tensor_flags)
    return tensor_objects, non_tensor_objects, tensor_flags


def merge_tensors(tensor_objects, non_tensor_objects, tensor_flags):
    """Inverse of extract_tensors."""
    merged = []
    t_idx = n_idx = 0
    real_tensor_flags = None
    if PARTITION_ACTIVATIONS:
        # after partitioning every tensor became (partition, size): drop the flag of the size tensor
        real_tensor_flags = []
        previous_flag = False
        for flag in tensor_flags:
            if previous_flag:
                previous_flag = False
                continue
            previous_flag = flag
            real_tensor_flags.append(flag)
    else:
        real_tensor_flags = tensor_flags
    for is_tensor in real_tensor_flags:
        if is_tensor:
            merged.append(tensor_objects[t_idx])
            t_idx += 1
        else:
            merged.append(non_tensor_objects[n_idx])
            n_idx += 1
    return tuple(merged)


class CheckpointFunction(torch.autograd.Function):
    """Recompute-in-backward with RNG state restore (CPU + CUDA + tracker states)."""

    @staticmethod
    def forward(ctx, run_function, all_outputs, *args):
        global mpu, timers, SYNCHRONIZE, PROFILE_TIME, cuda_device, transport_stream
        global mp_rank, mp_size, mp_group, contiguous_data_buffers, contiguous_size_buffers, data_offsets, size_offsets

        def save_args_for_backward(*all_args):
            tensor_args, non_tensor_args, tensor_flags = extract_tensors(all_objects=all_args)
            ctx.save_for_backward(*tensor_args)
            ctx.non_tensor_args = non_tensor_args
            ctx.tensor_flags = tensor_flags

        if SYNCHRONIZE and _cuda():
            torch.cuda.synchronize()
        if timers is None and PROFILE_TIME:
            timers = Timers()
        if PROFILE_TIME:
            timers("forward").start()

        ctx.run_function = run_function
        global num_layers
        if mp_rank is None:
            if mpu is not None:
                mp_rank = mpu.get_model_parallel_rank()
                mp_size = mpu.get_model_parallel_world_size()
                mp_group = mpu.get_model_parallel_group()
            else:
                mp_rank, mp_size, mp_group = 0, 1, None
        if cuda_device is None and _cuda():
            cuda_device = torch.cuda.current_device()
            if PARTITION_ACTIVATIONS or PA_TO_CPU:
                transport_stream = torch.cuda.Stream(device=cuda_device)

        inputs = []
        if PARTITION_ACTIVATIONS:
            for i, item in enumerate(args[:-1]):
                if not torch.is_tensor(item):
                    inputs.append(item)
                    continue
                partition_size = get_partition_size(item)
                partition = item.detach().contiguous().view(-1).narrow(0, get_partition_start(item), partition_size).clone()
                if CONTIGUOUS_CHECKPOINTING:
                    buffer_device = torch.device("cpu") if PA_TO_CPU else partition.device
                    if i >= len(contiguous_data_buffers):
                        contiguous_data_buffers.append(torch.tensor(()).new_empty([num_layers * partition.numel()],
                                                                                  dtype=partition.dtype,
                                                                                  device=buffer_device))
                        data_offsets.append(0)
                    elif contiguous_data_buffers[i] is None:
                        contiguous_data_buffers[i] = torch.tensor(()).new_empty([num_layers * partition.numel()],
                                                                                dtype=partition.dtype,
                                                                                device=buffer_device)
                        data_offsets[i] = 0
                    contiguous_partition = contiguous_data_buffers[i].narrow(0, data_offsets[i],
                                                                             partition.numel()).data.copy_(partition.data)
                    data_offsets[i] = data_offsets[i] + partition.numel()
                    inputs.append(contiguous_partition)
                else:
                    inputs.append(partition.to("cpu") if PA_TO_CPU else partition)
            inputs.append(args[-1])
        elif PA_TO_CPU:
            for item in args:
                inputs.append(item.detach().to("cpu", non_blocking=True) if torch.is_tensor(item) else item)

        inputs_cuda = move_to_device(args, cuda_device) if (cuda_device is not None) else args
        # capture RNG states so the recompute sees identical dropout masks
        ctx.fwd_cpu_rng_state = torch.get_rng_state()
        ctx.fwd_cuda_rng_state = torch.cuda.get_rng_state() if _cuda() else None
        ctx.fwd_cuda_rng_state_tracker = get_cuda_rng_tracker().get_states()

        with torch.no_grad():
            outputs = run_function(*inputs_cuda)
        del inputs_cuda

        if PARTITION_ACTIVATIONS:
            new_args = []
            for i, (arg, inp) in enumerate(zip(args, inputs)):
                if not torch.is_tensor(arg):
                    new_args.append(arg)
                    continue
                size = torch.tensor(arg.size())
                arg.data = inp.data
                new_args.append(arg)
                if CONTIGUOUS_CHECKPOINTING:
                    numel = size.numel()
                    if i >= len(contiguous_size_buffers):
                        contiguous_size_buffers.append(torch.tensor(()).new_empty([numel * num_layers],
                                                                                  dtype=size.dtype, device=size.device))
                        size_offsets.append(0)
                    elif contiguous_size_buffers[i] is None:
                        contiguous_size_buffers[i] = torch.tensor(()).new_empty([numel * num_layers], dtype=size.dtype,
                                                                                device=size.device)
                        size_offsets[i] = 0
                    contiguous_size = contiguous_size_buffers[i].narrow(0, size_offsets[i], numel).data.copy_(size.data)
                    contiguous_size = contiguous_size.view_as(size)
                    size_offsets[i] = size_offsets[i] + numel
                    new_args.append(contiguous_size)
                else:
                    new_args.append(size)
            save_args_for_backward(*new_args)
        elif PA_TO_CPU:
            save_args_for_backward(*inputs)
        else:
            save_args_for_backward(*args)

        if PROFILE_TIME:
            timers("forward").stop()
            timers.log(["forward"])
        if SYNCHRONIZE and _cuda():
            torch.cuda.synchronize()

        if torch.is_tensor(outputs):
            non_grad_outputs = [outputs] if not outputs.is_floating_point() else []
        else:
            non_grad_outputs = [o for o in outputs if torch.is_tensor(o) and not o.is_floating_point()]
        ctx.mark_non_differentiable(*non_grad_outputs)

        if torch.is_tensor(outputs):
            all_outputs += [outputs]
            return outputs
        all_outputs += outputs
        outputs, _, _ = extract_tensors(all_objects=outputs)
        return tuple(outputs)

    @staticmethod
    def backward(ctx, *grads):
        global timers, SYNCHRONIZE, PROFILE_TIME, cuda_device, transport_stream
        global contiguous_data_buffers, contiguous_size_buffers, data_offsets, size_offsets
        if SYNCHRONIZE and _cuda():
            torch.cuda.synchronize()
        if PROFILE_TIME:
            timers("backward").start()
        if CONTIGUOUS_CHECKPOINTING:
            for buffers in (contiguous_data_buffers, contiguous_size_buffers):
                for i in range(len(buffers)):
                    buffers[i] = None
            data_offsets[:] = [0] * len(data_offsets)
            size_offsets[:] = [0] * len(size_offsets)
        if not torch.autograd._is_checkpoint_valid():
            raise RuntimeError("Checkpointing is not compatible with .grad(), please use .backward() if possible")

        if PARTITION_ACTIVATIONS:
            inputs = get_full_inputs(ctx.saved_tensors, device=cuda_device if PA_TO_CPU else None)
            detached_inputs = detach_variable(inputs)
        elif PA_TO_CPU:
            inputs = move_to_device(ctx.saved_tensors, cuda_device)
            detached_inputs = detach_variable(inputs)
        else:
            inputs = ctx.saved_tensors
            detached_inputs = detach_variable(inputs)
        detached_inputs = merge_tensors(tensor_objects=detached_inputs, non_tensor_objects=ctx.non_tensor_args,
                                        tensor_flags=ctx.tensor_flags)

        bwd_cpu_rng_state = torch.get_rng_state()
        bwd_cuda_rng_state = torch.cuda.get_rng_state() if _cuda() else None
        bwd_tracker = get_cuda_rng_tracker().get_states()
        torch.set_rng_state(ctx.fwd_cpu_rng_state)
        if ctx.fwd_cuda_rng_state is not None:
            _set_cuda_rng_state(ctx.fwd_cuda_rng_state)
        get_cuda_rng_tracker().set_states(ctx.fwd_cuda_rng_state_tracker)

        with torch.enable_grad():
            outputs = ctx.run_function(*detached_inputs)

        torch.set_rng_state(bwd_cpu_rng_state)
        if bwd_cuda_rng_state is not None:
            _set_cuda_rng_state(bwd_cuda_rng_state)
        get_cuda_rng_tracker().set_states(bwd_tracker)

        if isinstance(outputs, torch.Tensor):
            outputs = (outputs,)
        outputs, _, _ = extract_tensors(all_objects=outputs)
        # only outputs that require grad take part in backward
        output_tensors, grad_tensors = [], []
        for out, grad in zip(outputs, grads):
            if out.requires_grad:
                output_tensors.append(out)
                grad_tensors.append(grad)
        torch.autograd.backward(output_tensors, grad_tensors)

        if PROFILE_TIME:
            timers("backward").stop()
            timers.log(["backward"])
        if SYNCHRONIZE and _cuda():
            torch.cuda.synchronize()
        ret_list = [None, None]
        for inp in detached_inputs:
            ret_list.append(inp.grad if torch.is_tensor(inp) else None)
        return tuple(ret_list)


def move_to_device(item, device):
    if device is None:
        return item
    if torch.is_tensor(item):
        return item.to(device)
    if isinstance(item, (list, tuple)):
        return type(item)(move_to_device(v, device) for v in item)
    return item


def checkpoint(function, *args):
    """Checkpoint a model or part of the model (recompute in backward)."""
    all_outputs = []
    CheckpointFunction.apply(function, all_outputs, *args)
    if len(all_outputs) == 1:
        return all_outputs[0]
    return tuple(all_outputs)


def partition_activations_in_checkpoint(partition_activation):
    global PARTITION_ACTIVATIONS
    PARTITION_ACTIVATIONS = partition_activation
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(f"**************Partition Activations {PARTITION_ACTIVATIONS}************")


def set_num_layers(nlayers):
    global num_layers
    num_layers = nlayers


def reset():
    """Free contiguous buffers between forward passes (e.g. between train and eval)."""
    global contiguous_data_buffers, contiguous_size_buffers, data_offsets, size_offsets
    if CONTIGUOUS_CHECKPOINTING:
        for buffers in (contiguous_data_buffers, contiguous_size_buffers):
            for i in range(len(buffers)):
                buffers[i] = None       # re-allocated by the next forward (the reference stores an empty tensor here,
                                        # which its own forward then fails to narrow)
        data_offsets[:] = [0] * len(data_offsets)
        size_offsets[:] = [0] * len(size_offsets)


def _configure_using_config_file(deepspeed_config, mpu=None):
    global num_layers, PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, PA_TO_CPU, SYNCHRONIZE, PROFILE_TIME
    config = DeepSpeedConfig(deepspeed_config, mpu=mpu).activation_checkpointing_config
    if dist.is_initialized() and dist.get_rank() == 0:
        logger.info(config.repr())
    PARTITION_ACTIVATIONS = config.partition_activations
    CONTIGUOUS_CHECKPOINTING = config.contiguous_memory_optimization
    num_layers = config.number_checkpoints
    PA_TO_CPU = config.cpu_checkpointing
    SYNCHRONIZE = config.synchronize_checkpoint_boundary
    PROFILE_TIME = config.profile


def _configure_defaults():
    global mpu, num_layers, deepspeed_checkpointing_enabled
    global PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, PA_TO_CPU, SYNCHRONIZE, PROFILE_TIME
    PARTITION_ACTIVATIONS = False
    CONTIGUOUS_CHECKPOINTING = False
    num_layers = False
    PA_TO_CPU = False
    SYNCHRONIZE = False
    PROFILE_TIME = False
    deepspeed_checkpointing_enabled = True


def configure(mpu_, deepspeed_config=None, partition_activations=None, contiguous_checkpointing=None,
              num_checkpoints=None, checkpoint_in_cpu=None, synchronize=None, profile=None):
    """Configure DeepSpeed activation checkpointing (config file first, explicit arguments override)."""
    global mpu, num_layers, deepspeed_checkpointing_enabled
    global PARTITION_ACTIVATIONS, CONTIGUOUS_CHECKPOINTING, PA_TO_CPU, SYNCHRONIZE, PROFILE_TIME
    _configure_defaults()
    if mpu_ is not None:
        mpu = mpu_
    if deepspeed_config is not None:
        _configure_using_config_file(deepspeed_config, mpu=mpu)
    if partition_activations is not None:
        PARTITION_ACTIVATIONS = partition_activations
    if contiguous_checkpointing is not None:
        CONTIGUOUS_CHECKPOINTING = contiguous_checkpointing
    if num_checkpoints is not None:
        num_layers = num_checkpoints
    if checkpoint_in_cpu is not None:
        PA_TO_CPU = checkpoint_in_cpu
    if synchronize is not None:
        SYNCHRONIZE = synchronize
    if profile is not None:
        PROFILE_TIME = profile
    if PA_TO_CPU or CONTIGUOUS_CHECKPOINTING:
        assert PARTITION_ACTIVATIONS, "CPU Checkpointing/Contiguous Checkpointing is only available with partitioned activations. Set partitioned activations to true in deepspeed config"
    if CONTIGUOUS_CHECKPOINTING:
        assert num_layers is not None, "Must specify the number of layers with contiguous memory checkpointing"


def is_configured():
    return deepspeed_checkpointing_enabled
