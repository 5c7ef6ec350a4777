"""Host staging memory for tensor swapping: a slab (`SwapBuffer`) packs tensors back to back at I/O-aligned offsets,
a `SwapBufferPool` spills from one slab to the next, a `SwapBufferManager` owns a fixed set of pinned slabs and lends
them out. Public names follow the reference (deepspeed/runtime/swap_tensor/utils.py:37-241) because the swappers and
user code address them by name; the bookkeeping is a list of `_Slot` records per slab."""
from dataclasses import dataclass
from typing import List, Optional

import torch

from ...utils.logging import logger
from .constants import AIO_ALIGNED_BYTES, MIN_AIO_BYTES  # noqa: F401


def swap_in_tensors(swap_handle, tensor_buffers, swap_paths):
    """Submit one asynchronous read per (buffer, file); the caller waits on the handle."""
    for buf, path in zip(tensor_buffers, swap_paths):
        rc = swap_handle.async_pread(buf, path)
        assert rc == 0, f"async read of {path} failed with {rc}"


def swap_out_tensors(swap_handle, tensor_buffers, swap_paths):
    """Submit one asynchronous write per (buffer, file); the caller waits on the handle."""
    for buf, path in zip(tensor_buffers, swap_paths):
        rc = swap_handle.async_pwrite(buf, path)
        assert rc == 0, f"async write of {path} failed with {rc}"


def print_object(obj, name, exclude_list=()):
    logger.info(f"{name}:")
    for key in sorted(vars(obj)):
        if key not in exclude_list:
            logger.info(f"  {key} {'.' * max(1, 29 - len(key))} {getattr(obj, key)}")


def _host_slab(numel, dtype):
    t = torch.zeros(numel, dtype=dtype, device="cpu")
    return t.pin_memory() if torch.cuda.is_available() else t


@dataclass
class _Slot:
    offset: int                 # element offset inside the slab
    io_view: torch.Tensor       # aligned region that is read / written as one file
    data_view: torch.Tensor     # the leading `numel` elements of io_view: what compute sees
    path: Optional[str]


class SwapBuffer:
    """One slab of staging memory with tensors packed in arrival order."""

    def __init__(self, buffer):
        self.buffer = buffer
        self._slots: List[_Slot] = []
        self._cursor = 0

    def reset(self):
        self._slots = []
        self._cursor = 0

    # capacity
    def has_space(self, numel):
        return self._cursor + numel <= self.buffer.numel()

    def get_num_elem(self):
        return sum(s.data_view.numel() for s in self._slots)

    # packing
    def allocate_tensor(self, swap_path, numel, aligned_numel):
        assert self.has_space(aligned_numel), "slab is full"
        io_view = self.buffer.narrow(0, self._cursor, aligned_numel)
        slot = _Slot(self._cursor, io_view, io_view.narrow(0, 0, numel), swap_path)
        self._slots.append(slot)
        self._cursor += aligned_numel
        return slot.io_view, slot.data_view

    def insert_tensor(self, tensor, swap_path, aligned_numel):
        io_view, data_view = self.allocate_tensor(swap_path, tensor.numel(), aligned_numel)
        data_view.copy_(tensor.detach().reshape(-1))
        return io_view, data_view

    # views
    def get_swap_tensors(self):
        return [s.io_view for s in self._slots]

    def get_compute_tensors(self):
        return [s.data_view for s in self._slots]

    def get_swap_paths(self):
        return [s.path for s in self._slots]

    def _at(self, offset):
        return next((s for s in self._slots if s.offset == offset), None)

    def get_swap_tensor(self, offset):
        s = self._at(offset)
        return None if s is None else s.io_view

    def get_compute_tensor(self, offset):
        s = self._at(offset)
        return None if s is None else s.data_view

    def get_swap_path(self, offset):
        s = self._at(offset)
        return None if s is None else s.path


class SwapBufferPool:
    """Several slabs filled one after the other."""

    def __init__(self, buffers):
        self.buffers = [SwapBuffer(b) for b in buffers]
        self._active = 0

    def reset(self):
        self._active = 0
        for slab in self.buffers:
            slab.reset()

    def has_space(self, numel):
        """True if the request fits the active slab or the next one (which then becomes active)."""
        if self.buffers[self._active].has_space(numel):
            return True
        if self._active + 1 < len(self.buffers) and self.buffers[self._active + 1].has_space(numel):
            self._active += 1
            return True
        return False

    def allocate_tensor(self, numel, swap_path, aligned_numel):
        if not self.has_space(aligned_numel):
            return None, None
        return self.buffers[self._active].allocate_tensor(swap_path, numel, aligned_numel)

    def insert_tensor(self, tensor, swap_path, aligned_numel):
        if not self.has_space(aligned_numel):
            return None, None
        return self.buffers[self._active].insert_tensor(tensor, swap_path, aligned_numel)

    def _collect(self, getter):
        out = []
        for slab in self.buffers[:self._active + 1]:
            out.extend(getter(slab))
        return out

    def get_swap_tensors(self):
        return self._collect(SwapBuffer.get_swap_tensors)

    def get_swap_paths(self):
        return self._collect(SwapBuffer.get_swap_paths)

    def get_compute_tensors(self):
        return self._collect(SwapBuffer.get_compute_tensors)

    def _transfer(self, submit, aio_handle, async_op):
        views, paths = self.get_swap_tensors(), self.get_swap_paths()
        assert all(p is not None for p in paths), "every staged tensor needs a swap path"
        submit(aio_handle, views, paths)
        if not async_op:
            done = aio_handle.wait()
            assert done == len(views), f"{done} of {len(views)} transfers completed"

    def swap_out(self, aio_handle, async_op=False):
        self._transfer(swap_out_tensors, aio_handle, async_op)

    def swap_in(self, aio_handle, async_op=False):
        self._transfer(swap_in_tensors, aio_handle, async_op)


class SwapBufferManager:
    """Owns `count` pinned slabs of `num_elems` elements and lends (prefixes of) them out."""

    def __init__(self, num_elems, count, dtype):
        self.num_elems, self.count, self.dtype = num_elems, count, dtype
        self.all_buffers = [_host_slab(num_elems, dtype) for _ in range(count)]
        self._available = list(range(count))
        self._lent = {}             # id(view handed out) -> slab index
        self.gigabytes = self.all_buffers[0].element_size() * num_elems * count / 2 ** 30

    def allocate(self, num_elems, count, dtype):
        assert dtype == self.dtype and num_elems <= self.num_elems
        if count > len(self._available):
            return None
        views = []
        for _ in range(count):
            idx = self._available.pop()
            view = self.all_buffers[idx].narrow(0, 0, num_elems)
            self._lent[id(view)] = idx
            views.append(view)
        return views

    def allocate_all(self, num_elems, dtype):
        return self.allocate(num_elems, len(self._available), dtype)

    def free(self, buffers):
        keys = [id(b) for b in buffers]
        unknown = [k for k in keys if k not in self._lent]
        assert not unknown, "free() of a buffer this manager did not lend"
        for k in keys:
            self._available.append(self._lent.pop(k))


def get_sized_buffer(buffer, num_elems):
    assert num_elems <= buffer.numel(), f"{num_elems} elements requested from a buffer of {buffer.numel()}"
    return buffer if num_elems == buffer.numel() else buffer.narrow(0, 0, num_elems)


def get_sized_buffers(buffer_list, num_elems_list):
    return [get_sized_buffer(b, n) for b, n in zip(buffer_list, num_elems_list)]
