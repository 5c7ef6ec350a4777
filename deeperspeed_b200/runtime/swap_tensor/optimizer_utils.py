"""Optimizer-state swapping base (ZeRO-Infinity): the fp32 master parameter of a sub-group and all of its optimizer
state tensors live in files under <nvme_path>/zero_stage_3/optimizer/rank<R>/ and are resident only while that
sub-group is being stepped. File naming and class names per reference deepspeed/runtime/swap_tensor/optimizer_utils.py
(layout in SURVEY C.13); the book-keeping is one small record per parameter: a list of (tensor, file) pairs plus the
gradient pieces that were written out, keyed by their offset inside the flat partition.
"""
import os
from collections import namedtuple

import torch
import torch.distributed as dist

from ...ops.aio import aio_handle
from . import constants as SC
from .utils import SwapBufferManager, get_sized_buffer, get_sized_buffers  # noqa: F401

# a piece of a flat gradient partition kept in its own file
FlattenedTensorSwapInfo = namedtuple("FlattenedTensorSwapInfo", ["path", "length", "offset"])

_AIO_KEYS = ((SC.AIO_BLOCK_SIZE, SC.AIO_BLOCK_SIZE_DEFAULT), (SC.AIO_QUEUE_DEPTH, SC.AIO_QUEUE_DEPTH_DEFAULT),
             (SC.AIO_SINGLE_SUBMIT, SC.AIO_SINGLE_SUBMIT_DEFAULT), (SC.AIO_OVERLAP_EVENTS, SC.AIO_OVERLAP_EVENTS_DEFAULT),
             (SC.AIO_THREAD_COUNT, SC.AIO_THREAD_COUNT_DEFAULT))


def make_aio_handle(aio_config):
    """aio_handle(block_size, queue_depth, single_submit, overlap_events, thread_count) from an `aio` config block."""
    given = aio_config or {}
    return aio_handle(*[given.get(key, default) for key, default in _AIO_KEYS])


class OptimizerStateSwapInfo:
    """Swap record of one parameter: tensors[0] is the fp32 master, the rest its optimizer state; file i holds tensor i."""

    def __init__(self, parameter, numel, base_folder):
        self.param_id = id(parameter)
        self.swap_folder = base_folder
        self.tensor_numel, self.tensor_dtype, self.tensor_device = numel, parameter.dtype, parameter.device
        self.tensors, self.swap_paths = [], []
        self.swapped_gradients = {}        # offset -> FlattenedTensorSwapInfo
        self.unswapped_gradients = {}      # offset -> tensor too small for a file of its own
        self.has_state_tensors = False
        self._track([parameter])

    def _file(self, suffix):
        return os.path.join(self.swap_folder, f"{self.param_id}_{suffix}.tensor.swp")

    def _track(self, tensors):
        for t in tensors:
            self.swap_paths.append(self._file(len(self.tensors)))
            self.tensors.append(t)

    def add_state_tensors(self, tensor_list):
        self.has_state_tensors = True
        self._track(tensor_list)

    # ---- queries
    def numel(self):
        return self.tensor_numel

    def dtype(self):
        return self.tensor_dtype

    def device(self):
        return self.tensor_device

    def has_gradients(self):
        return bool(self.swapped_gradients or self.unswapped_gradients)

    def get_unpinned_state_tensors(self):
        return [t for t in self.tensors if not t.is_pinned()]

    # ---- residency
    def set_swap_buffers(self, buffers):
        """Point every tracked tensor at the first `numel` elements of its (io-aligned) buffer."""
        for tensor, buffer in zip(self.tensors, buffers):
            tensor.data = buffer.narrow(0, 0, self.tensor_numel).data

    def release_memory(self):
        for tensor in self.tensors:
            tensor.data = torch.Tensor()

    # ---- gradients
    def get_or_create_gradient_paths(self, offsets, lengths):
        for offset, length in zip(offsets, lengths):
            if offset not in self.swapped_gradients:
                self.swapped_gradients[offset] = FlattenedTensorSwapInfo(self._file(f"gradient_{offset}_{length}"),
                                                                         length, offset)
        return [self.swapped_gradients[offset].path for offset in offsets]

    def get_swap_gradient_paths(self):
        return [piece.path for piece in self.swapped_gradients.values()]

    def get_swap_gradient_buffers(self, swap_buffer):
        assert self.tensor_numel <= swap_buffer.numel()
        return [swap_buffer.narrow(0, piece.offset, piece.length) for piece in self.swapped_gradients.values()]

    def read_unswapped_gradients(self, dest_buffer):
        """Copy the resident gradient pieces into their places; returns the number of elements copied."""
        copied = 0
        for offset, piece in self.unswapped_gradients.items():
            dest_buffer.narrow(0, offset, piece.numel()).data.copy_(piece.data)
            copied += piece.numel()
        return copied

    def release_unswapped_gradients(self):
        self.unswapped_gradients = {}


SWAPPER_DEBUG_MODE = False
SWAP_OUT_GRADIENT_TIMER = "swap_out_gradient"


class OptimizerSwapper:
    """Common state of the synchronous and the pipelined optimizer swapper."""

    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers):
        self.swap_config, self.aio_config = swap_config, aio_config
        self.optimizer, self.device, self.dtype, self.timers = optimizer, device, dtype, timers
        self.swap_params_info = {}
        self.timer_names = set()
        rank = dist.get_rank() if dist.is_initialized() else 0
        self.swap_folder = os.path.join(base_folder, "zero_stage_3", "optimizer", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        given = aio_config or {}
        # O_DIRECT granularity: every thread of the handle moves whole aligned blocks
        self.swap_element_size = torch.tensor([], dtype=dtype).element_size()
        self.aligned_bytes = SC.AIO_ALIGNED_BYTES * given.get(SC.AIO_THREAD_COUNT, SC.AIO_THREAD_COUNT_DEFAULT)
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.min_aio_bytes = max(SC.MIN_AIO_BYTES, given.get(SC.AIO_BLOCK_SIZE, SC.AIO_BLOCK_SIZE_DEFAULT))
        self.largest_numel = self._io_aligned_numel(largest_numel)

    def _io_aligned_numel(self, numel):
        return -(-numel // self.numel_alignment) * self.numel_alignment

    def swappable_tensor(self, param=None, numel=None):
        """Tensors smaller than one AIO block stay resident."""
        assert param is not None or numel is not None, "Either param or numel must be provided"
        count = param.numel() if param is not None else numel
        return count * self.swap_element_size >= self.min_aio_bytes

    def _get_param_swap_info(self, parameter):
        """The record of `parameter`, created on first use; optimizer state tensors are picked up once they exist."""
        info = self.swap_params_info.get(id(parameter))
        if info is None:
            info = self.swap_params_info[id(parameter)] = OptimizerStateSwapInfo(parameter, parameter.numel(),
                                                                                self.swap_folder)
        if not info.has_state_tensors:
            state = [v for v in self.optimizer.state.get(parameter, {}).values()
                     if torch.is_tensor(v) and v.numel() == parameter.numel()]
            if state:
                info.add_state_tensors(state)
        return info

    # ---- checkpoints
    def release_optimizer_state(self, parameter):
        """Drop the resident copy of a state that was swapped in and NOT modified (its files are current)."""
        if id(parameter) in getattr(self, "_on_disk", ()):
            self._get_param_swap_info(parameter).release_memory()
            getattr(self, "_resident", {}).pop(id(parameter), None)

    def forget(self, parameter):
        """Forget the swap record of `parameter` (its tensors were replaced, e.g. by load_state_dict): the next
        swap-out tracks the new tensors and rewrites the files."""
        self.swap_params_info.pop(id(parameter), None)
        getattr(self, "_on_disk", set()).discard(id(parameter))
        getattr(self, "_resident", {}).pop(id(parameter), None)
        getattr(self, "_prefetched", {}).pop(id(parameter), None)

    # ---- hooks of the ZeRO-3 step and timers
    def pre_backward(self):
        self.init_timers()

    def post_backward(self):
        pass

    def init_timers(self):
        self.timer_names = set()

    def log_timers(self):
        if self.timers and self.timer_names:
            self.timers.log(sorted(self.timer_names))

    def _start_timer(self, name):
        if self.timers:
            self.timers(name).start()

    def _stop_timer(self, name):
        if self.timers:
            self.timers(name).stop()
            self.timer_names.add(name)
