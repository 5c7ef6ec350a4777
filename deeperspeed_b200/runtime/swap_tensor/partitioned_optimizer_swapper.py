"""PartitionedOptimizerSwapper: synchronous swap-in before / swap-out after each sub-group's optimizer step
(reference deepspeed/runtime/swap_tensor/partitioned_optimizer_swapper.py:27-260)."""
import torch

from ...utils.logging import logger
from .optimizer_utils import OptimizerSwapper, make_aio_handle

SWAP_IN_PARAM_TIMER = "swap_in_param"
SWAP_OUT_PARAM_TIMER = "swap_out_param"


def _pinned(numel, dtype):
    t = torch.zeros(numel, dtype=dtype, device="cpu")
    return t.pin_memory() if torch.cuda.is_available() else t


class PartitionedOptimizerSwapper(OptimizerSwapper):
    def __init__(self, swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers):
        super().__init__(swap_config, aio_config, base_folder, optimizer, largest_numel, device, dtype, timers)
        self.aio_handle = make_aio_handle(aio_config)
        self._resident = {}     # param id -> list of pinned buffers currently backing its tensors
        self._on_disk = set()

    def initialize_parameters(self, parameters, src_tensors):
        """Write the initial fp32 values straight to their swap files."""
        for p, src in zip(parameters, src_tensors):
            p.data = src.detach().float().clone()
            self.swap_out_optimizer_state(parameter=p)

    def swap_in_optimizer_state(self, parameter, async_parameter=None):
        info = self._get_param_swap_info(parameter)
        if id(parameter) not in self._on_disk:
            return      # never swapped out yet: everything is resident
        self._start_timer(SWAP_IN_PARAM_TIMER)
        buffers = []
        n = info.numel()
        for t, path in zip(info.tensors, info.swap_paths):
            buf = _pinned(self._io_aligned_numel(n), info.dtype())
            self.aio_handle.async_pread(buf, path)
            buffers.append(buf)
        self.aio_handle.wait()
        info.set_swap_buffers(buffers)
        self._resident[id(parameter)] = buffers
        self._stop_timer(SWAP_IN_PARAM_TIMER)

    def swap_out_optimizer_state(self, parameter, async_swap=False):
        info = self._get_param_swap_info(parameter)
        self._start_timer(SWAP_OUT_PARAM_TIMER)
        n = info.numel()
        staged = []
        for t, path in zip(info.tensors, info.swap_paths):
            buf = _pinned(self._io_aligned_numel(n), info.dtype())
            buf.narrow(0, 0, n).copy_(t.data.view(-1))
            self.aio_handle.async_pwrite(buf, path)
            staged.append(buf)
        self.aio_handle.wait()
        info.release_memory()
        self._resident.pop(id(parameter), None)
        self._on_disk.add(id(parameter))
        self._stop_timer(SWAP_OUT_PARAM_TIMER)

    def swap_out_gradients(self, parameter, gradient_offsets, gradient_tensors):
        info = self._get_param_swap_info(parameter)
        paths = info.get_or_create_gradient_paths(gradient_offsets, [g.numel() for g in gradient_tensors])
        for g, path in zip(gradient_tensors, paths):
            buf = _pinned(self._io_aligned_numel(g.numel()), g.dtype)
            buf.narrow(0, 0, g.numel()).copy_(g.view(-1))
            self.aio_handle.async_pwrite(buf, path)
        self.aio_handle.wait()
