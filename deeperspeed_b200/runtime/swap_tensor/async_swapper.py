"""AsyncTensorSwapper: streams a sequence of tensors (e.g. gradient partitions) out to files through a small ring of
pinned staging slabs, so that filling one slab overlaps the file writes of the previous ones.

Same role and public surface as the reference class (deepspeed/runtime/swap_tensor/async_swapper.py:16-173:
add_buffers / swap_out_tensors / release_buffers / has_buffers / get_timer_names); the mechanics here are a three-queue
slab pipeline:

    idle slabs  --fill-->  the slab being packed  --seal-->  in-flight slabs (writes submitted)  --drain-->  idle

A slab is sealed when the next tensor does not fit; sealing submits one asynchronous write per packed tensor. A drain
(`aio_handle.wait()`) happens only when no idle slab is left, or at `release_buffers()`."""
from collections import deque

import torch

from ...utils.logging import logger
from .utils import SwapBuffer

ASYNC_SWAPPER_WAIT_TIMER = "async_swap_gradient_wait"


class AsyncTensorSwapper:
    def __init__(self, aio_handle, numel_alignment, timers):
        self.aio_handle = aio_handle
        self.numel_alignment = int(numel_alignment)
        self.timers = timers
        self._slabs = []            # every SwapBuffer handed to us
        self._idle = deque()        # slabs with nothing in them
        self._packing = None        # the slab currently receiving tensors
        self._in_flight = []        # sealed slabs whose writes have been submitted
        self._writes_pending = 0
        self._timer_names = set()
        self._elements_out = 0
        self._dtype = None
        self._slab_capacity = 0

    # ------------------------------------------------------------------ buffers
    def has_buffers(self):
        return bool(self._slabs)

    def add_buffers(self, buffer_list):
        assert not self._slabs, "buffers were already handed over; release_buffers() first"
        assert buffer_list and all(not b.is_cuda for b in buffer_list), "staging buffers live in (pinned) host memory"
        dtypes = {b.dtype for b in buffer_list}
        assert len(dtypes) == 1, f"staging buffers must share one dtype, got {dtypes}"
        self._dtype = dtypes.pop()
        self._slabs = [SwapBuffer(b) for b in buffer_list]
        self._idle = deque(self._slabs)
        self._slab_capacity = max(b.numel() for b in buffer_list)
        self._timer_names = set()

    def release_buffers(self):
        """Flush everything that is still staged, then hand the raw pinned buffers back to the caller."""
        self._log_volume("swapped out (before final flush)")
        self._seal()
        self._drain()
        assert len(self._idle) == len(self._slabs)
        self._log_volume("swapped out (after final flush)")
        raw = [slab.buffer for slab in self._slabs]
        self._slabs, self._idle, self._packing = [], deque(), None
        self._elements_out, self._dtype = 0, None
        return raw

    def get_timer_names(self):
        return sorted(self._timer_names)

    # ------------------------------------------------------------------ swapping
    def swap_out_tensors(self, tensor_list, path_list):
        assert self._slabs, "add_buffers() first"
        for tensor, path in zip(tensor_list, path_list):
            need = self._round_up(tensor.numel())
            assert need <= self._slab_capacity, f"tensor of {tensor.numel()} elements exceeds the staging slab size"
            if self._packing is not None and not self._packing.has_space(need):
                self._seal()
            if self._packing is None:
                if not self._idle:
                    self._drain()
                self._packing = self._idle.popleft()
            self._packing.insert_tensor(tensor, path, need)

    def _round_up(self, numel):
        a = self.numel_alignment
        return (numel + a - 1) // a * a

    def _seal(self):
        """Submit the writes of the slab being packed (if any) and move it to the in-flight list."""
        slab, self._packing = self._packing, None
        if slab is None:
            return
        tensors, paths = slab.get_swap_tensors(), slab.get_swap_paths()
        for t, p in zip(tensors, paths):
            assert self.aio_handle.async_pwrite(t, p) == 0
        self._writes_pending += len(tensors)
        self._in_flight.append(slab)

    def _drain(self):
        """Wait for every submitted write, recycle the slabs."""
        if not self._in_flight:
            return
        if self.timers:
            self.timers(ASYNC_SWAPPER_WAIT_TIMER).start()
        done = self.aio_handle.wait()
        if self.timers:
            self.timers(ASYNC_SWAPPER_WAIT_TIMER).stop()
            self._timer_names.add(ASYNC_SWAPPER_WAIT_TIMER)
        assert done == self._writes_pending, f"expected {self._writes_pending} completed writes, got {done}"
        self._writes_pending = 0
        for slab in self._in_flight:
            self._elements_out += slab.get_num_elem()
            slab.reset()
            self._idle.append(slab)
        self._in_flight = []

    def _log_volume(self, what):
        if self._dtype is None or not torch.distributed.is_initialized() or torch.distributed.get_rank() != 0:
            return
        gib = self._elements_out * torch.empty((), dtype=self._dtype).element_size() / 2 ** 30
        logger.info(f"{what}: {self._elements_out} elements, {gib:5.2f} GiB")

    def _log_timers(self, name_list, force=False):
        if self.timers and force:
            self.timers.log(name_list)
