"""Optimizer-state swapping base (ZeRO-Infinity): the fp32 master parameter of a sub-group and all of its optimizer
state tensors live in files under <nvme_path>/zero_stage_3/optimizer/rank<R>/ and are resident only while that
sub-group is being stepped. Reference: deepspeed/runtime/swap_tensor/optimizer_utils.py:118-526 (layout in SURVEY C.13)."""
import os

import torch
import torch.distributed as dist

from ...ops.aio import aio_handle
from ...utils.logging import logger
from .constants import (AIO_BLOCK_SIZE, AIO_QUEUE_DEPTH, AIO_THREAD_COUNT, AIO_SINGLE_SUBMIT, AIO_OVERLAP_EVENTS,
                        AIO_ALIGNED_BYTES, MIN_AIO_BYTES, AIO_BLOCK_SIZE_DEFAULT, AIO_QUEUE_DEPTH_DEFAULT,
                        AIO_THREAD_COUNT_DEFAULT, AIO_SINGLE_SUBMIT_DEFAULT, AIO_OVERLAP_EVENTS_DEFAULT)


def make_aio_handle(aio_config):
    c = aio_config or {}
    return aio_handle(c.get(AIO_BLOCK_SIZE, AIO_BLOCK_SIZE_DEFAULT), c.get(AIO_QUEUE_DEPTH, AIO_QUEUE_DEPTH_DEFAULT),
                      c.get(AIO_SINGLE_SUBMIT, AIO_SINGLE_SUBMIT_DEFAULT),
                      c.get(AIO_OVERLAP_EVENTS, AIO_OVERLAP_EVENTS_DEFAULT),
                      c.get(AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT))


class FlattenedTensorSwapInfo:
    def __init__(self, path, length, offset):
        self.path = path
        self.offset = offset
        self.length = length


class OptimizerStateSwapInfo:
    """Book-keeping for one swappable parameter: which tensors (master + state) exist and where their files are."""

    def __init__(self, parameter, numel, base_folder):
        self.tensors = []
        self.param_id = id(parameter)
        self.swap_folder = base_folder
        self.swap_paths = []
        self.swapped_gradients = {}
        self.unswapped_gradients = {}
        self.tensor_numel = numel
        self.tensor_dtype = parameter.dtype
        self.tensor_device = parameter.device
        self.has_state_tensors = False
        self._add_tensors([parameter])

    def numel(self):
        return self.tensor_numel

    def has_gradients(self):
        return self.swapped_gradients or self.unswapped_gradients

    def _add_tensors(self, tensor_list):
        for t in tensor_list:
            self.tensors.append(t)
            self.swap_paths.append(os.path.join(self.swap_folder, f"{self.param_id}_{len(self.tensors) - 1}.tensor.swp"))

    def add_state_tensors(self, tensor_list):
        self.has_state_tensors = True
        self._add_tensors(tensor_list)

    def device(self):
        return self.tensor_device

    def dtype(self):
        return self.tensor_dtype

    def release_memory(self):
        for tensor in self.tensors:
            tensor.data = torch.Tensor()

    def get_or_create_gradient_paths(self, offsets, lengths):
        gradient_paths = []
        for offset, length in zip(offsets, lengths):
            if offset not in self.swapped_gradients:
                path = os.path.join(self.swap_folder, f"{self.param_id}_gradient_{offset}_{length}.tensor.swp")
                self.swapped_gradients[offset] = FlattenedTensorSwapInfo(path, length, offset)
            gradient_paths.append(self.swapped_gradients[offset].path)
        return gradient_paths

    def set_swap_buffers(self, buffers):
        compute_lengths = [self.numel()] * len(self.tensors)
        for t, b, n in zip(self.tensors, buffers, compute_lengths):
            t.data = b.narrow(0, 0, n).data

    def get_swap_gradient_buffers(self, swap_buffer):
        assert self.numel() <= swap_buffer.numel()
        return [swap_buffer.narrow(0, g.offset, g.length) for g in self.swapped_gradients.values()]

    def get_swap_gradient_paths(self):
        return [g.path for g in self.swapped_gradients.values()]

    def get_unpinned_state_tensors(self):
        return [t for t in self.tensors if not t.is_pinned()]

    def read_unswapped_gradients(self, dest_buffer):
        num_elem_count = 0
        for offset, grad_partition in self.unswapped_gradients.items():
            dst = dest_buffer.narrow(0, offset, grad_partition.numel())
            dst.data.copy_(grad_partition.data)
            num_elem_count += grad_partition.numel()
        return num_elem_count

    def release_unswapped_gradients(self):
        self.unswapped_gradients = {}


SWAPPER_DEBUG_MODE = False
SWAP_OUT_GRADIENT_TIMER = "swap_out_gradient"


class OptimizerSwapper:
    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers):
        self.swap_config = swap_config
        self.aio_config = aio_config
        self.swap_params_info = {}
        self.swap_element_size = torch.tensor([], dtype=dtype).element_size()
        rank = dist.get_rank() if dist.is_initialized() else 0
        self.swap_folder = os.path.join(base_folder, "zero_stage_3", "optimizer", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.optimizer = optimizer
        block = (aio_config or {}).get(AIO_BLOCK_SIZE, AIO_BLOCK_SIZE_DEFAULT)
        threads = (aio_config or {}).get(AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT)
        self.min_aio_bytes = max(MIN_AIO_BYTES, block)
        self.aligned_bytes = AIO_ALIGNED_BYTES * threads
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.largest_numel = self._io_aligned_numel(largest_numel)
        self.dtype = dtype
        self.device = device
        self.timers = timers
        self.timer_names = set()

    def swappable_tensor(self, param=None, numel=None):
        assert param is not None or numel is not None, "Either param or numel must be provided"
        if param is not None:
            return self.min_aio_bytes <= (param.numel() * self.swap_element_size)
        return self.min_aio_bytes <= (numel * self.swap_element_size)

    def init_timers(self):
        self.timer_names = set()

    def log_timers(self):
        if self.timer_names and self.timers:
            self.timers.log(list(self.timer_names))

    def pre_backward(self):
        self.init_timers()

    def post_backward(self):
        pass

    def _io_aligned_numel(self, numel):
        remainder = numel % self.numel_alignment
        return numel if remainder == 0 else (numel + self.numel_alignment - remainder)

    def _get_param_swap_info(self, parameter):
        """Create (lazily) the swap record of a parameter; pick up optimizer state tensors once they exist."""
        pid = id(parameter)
        info = self.swap_params_info.get(pid)
        if info is None:
            info = OptimizerStateSwapInfo(parameter=parameter, numel=parameter.numel(), base_folder=self.swap_folder)
            self.swap_params_info[pid] = info
        if not info.has_state_tensors:
            state = self.optimizer.state.get(parameter, {})
            tensors = [v for v in state.values() if torch.is_tensor(v) and v.numel() == parameter.numel()]
            if tensors:
                info.add_state_tensors(tensors)
        return info

    def _start_timer(self, name):
        if self.timers:
            self.timers(name).start()

    def _stop_timer(self, name):
        if self.timers:
            self.timers(name).stop()
            self.timer_names.add(name)
