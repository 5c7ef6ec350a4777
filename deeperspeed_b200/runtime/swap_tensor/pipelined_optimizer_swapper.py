"""PipelinedOptimizerSwapper: overlaps the read of the NEXT sub-group's state and the write of the PREVIOUS one with
the current sub-group's optimizer step, using separate read / write aio handles
(reference deepspeed/runtime/swap_tensor/pipelined_optimizer_swapper.py:60-284)."""
import torch

from .optimizer_utils import OptimizerSwapper, make_aio_handle
from .partitioned_optimizer_swapper import _pinned


class PipelinedOptimizerSwapper(OptimizerSwapper):
    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers):
        super().__init__(swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers)
        self.write_aio_handle = make_aio_handle(aio_config)
        self.read_aio_handle = make_aio_handle(aio_config)
        self.async_swap_in = bool((swap_config or {}).get("pipeline_read", False))
        self.async_swap_out = bool((swap_config or {}).get("pipeline_write", False))
        self._on_disk = set()
        self._prefetched = {}      # param id -> buffers being read
        self._pending_write = None  # (info, staged buffers)

    def initialize_parameters(self, parameters, src_tensors):
        for p, src in zip(parameters, src_tensors):
            p.data = src.detach().float().clone()
            self.swap_out_optimizer_state(parameter=p)
        self._flush_write()

    def _issue_read(self, parameter):
        info = self._get_param_swap_info(parameter)
        if self._pending_write is not None and self._pending_write[0] is info:
            # this very state is still on its way to the file (asynchronous swap-out of the previous step; with a
            # single sub-group nothing else flushes it in between): reading now would return the old contents
            self._flush_write()
        n = info.numel()
        bufs = []
        for path in info.swap_paths:
            buf = _pinned(self._io_aligned_numel(n), info.dtype())
            self.read_aio_handle.async_pread(buf, path)
            bufs.append(buf)
        self._prefetched[id(parameter)] = bufs

    def swap_in_optimizer_state(self, parameter, async_parameter=None):
        if id(parameter) in self._on_disk:
            if id(parameter) not in self._prefetched:
                self._issue_read(parameter)
            self.read_aio_handle.wait()
            info = self._get_param_swap_info(parameter)
            info.set_swap_buffers(self._prefetched.pop(id(parameter)))
        # start fetching the next sub-group while this one is being stepped
        if self.async_swap_in and async_parameter is not None and id(async_parameter) in self._on_disk \
                and id(async_parameter) not in self._prefetched:
            self._issue_read(async_parameter)

    def _flush_write(self):
        if self._pending_write is not None:
            self.write_aio_handle.wait()
            info, _ = self._pending_write
            self._pending_write = None

    def swap_out_optimizer_state(self, parameter, async_swap=False):
        self._flush_write()
        info = self._get_param_swap_info(parameter)
        n = info.numel()
        staged = []
        for t, path in zip(info.tensors, info.swap_paths):
            buf = _pinned(self._io_aligned_numel(n), info.dtype())
            buf.narrow(0, 0, n).copy_(t.data.view(-1))
            self.write_aio_handle.async_pwrite(buf, path)
            staged.append(buf)
        info.release_memory()
        self._on_disk.add(id(parameter))
        self._pending_write = (info, staged)
        if not (self.async_swap_out or async_swap):
            self._flush_write()
