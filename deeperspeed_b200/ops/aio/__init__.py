"""aio_handle: asynchronous tensor <-> file I/O (kernel AIO, O_DIRECT, multi-threaded) for the NVMe tier.
API per the reference pybind module `async_io` (csrc/aio/py_lib/py_ds_aio.cpp:12-41):
  aio_handle(block_size, queue_depth, single_submit, overlap_events, num_threads)
    .sync_pread/.sync_pwrite/.async_pread/.async_pwrite(tensor, filename) , .wait(), getters
  aio_read / aio_write (one-shot), deepspeed_memcpy. CUDA tensors bounce through a pinned host buffer."""
import ctypes as C

import torch

from .. import native as N

_p, _i, _l = C.c_void_p, C.c_int, C.c_long
_bound = False


def _lib():
    global _bound
    lib = N.host_lib()
    if not _bound:
        lib.dsb_aio_handle_create.restype = _p
        lib.dsb_aio_handle_create.argtypes = [_l, _i, _i, _i, _i]
        lib.dsb_aio_handle_destroy.argtypes = [_p]
        lib.dsb_aio_submit.argtypes = [_p, _i, _p, _l, C.c_char_p]
        lib.dsb_aio_wait.restype = _l
        lib.dsb_aio_wait.argtypes = [_p]
        lib.dsb_aio_pending.restype = _l
        lib.dsb_aio_pending.argtypes = [_p]
        lib.dsb_aio_sync_rw.restype = _l
        lib.dsb_aio_sync_rw.argtypes = [_i, _p, _l, C.c_char_p, _l, _i, _i, _i]
        lib.dsb_aio_alloc_aligned.restype = _p
        lib.dsb_aio_alloc_aligned.argtypes = [_l, _i]
        lib.dsb_aio_free_aligned.argtypes = [_p, _l, _i]
        lib.dsb_parallel_memcpy.argtypes = [_p, _p, _l]
        _bound = True
    return lib


class aio_handle:
    def __init__(self, block_size=1048576, queue_depth=8, single_submit=False, overlap_events=True, num_threads=1):
        self._lib = _lib()
        self._block_size, self._queue_depth = block_size, queue_depth
        self._single_submit, self._overlap_events, self._num_threads = single_submit, overlap_events, num_threads
        self._h = self._lib.dsb_aio_handle_create(block_size, queue_depth, int(single_submit), int(overlap_events),
                                                  num_threads)
        self._keep = []      # (tensor, bounce buffer) kept alive until wait()

    def __del__(self):
        try:
            if self._h:
                self._lib.dsb_aio_wait(self._h)
                self._lib.dsb_aio_handle_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def get_block_size(self):
        return self._block_size

    def get_queue_depth(self):
        return self._queue_depth

    def get_single_submit(self):
        return self._single_submit

    def get_overlap_events(self):
        return self._overlap_events

    def get_thread_count(self):
        return self._num_threads

    def _submit(self, is_read, tensor, filename, validate=False):
        assert tensor.is_contiguous(), "aio needs contiguous tensors"
        nbytes = tensor.numel() * tensor.element_size()
        if tensor.is_cuda:
            bounce = torch.empty(tensor.shape, dtype=tensor.dtype, device="cpu", pin_memory=True)
            if not is_read:
                bounce.copy_(tensor)
            host = bounce
        else:
            bounce = None
            host = tensor
        rc = self._lib.dsb_aio_submit(self._h, int(is_read), host.data_ptr(), nbytes, str(filename).encode())
        if rc < 0:
            raise OSError(-rc, f"aio submit failed for {filename}")
        self._keep.append((tensor, bounce, is_read))
        return 0

    def wait(self):
        r = self._lib.dsb_aio_wait(self._h)
        if r < 0:
            self._keep = []
            raise OSError(-r, "aio operation failed")
        for tensor, bounce, is_read in self._keep:
            if bounce is not None and is_read:
                tensor.copy_(bounce, non_blocking=False)
        self._keep = []
        return int(r)

    def read(self, tensor, filename, validate=False):
        self._submit(True, tensor, filename, validate)
        return self.wait()

    def write(self, tensor, filename, validate=False):
        self._submit(False, tensor, filename, validate)
        return self.wait()

    def pread(self, tensor, filename, validate, async_op):
        self._submit(True, tensor, filename, validate)
        return 0 if async_op else self.wait()

    def pwrite(self, tensor, filename, validate, async_op):
        self._submit(False, tensor, filename, validate)
        return 0 if async_op else self.wait()

    def sync_pread(self, tensor, filename):
        return self.pread(tensor, filename, False, False)

    def sync_pwrite(self, tensor, filename):
        return self.pwrite(tensor, filename, False, False)

    def async_pread(self, tensor, filename):
        return self.pread(tensor, filename, False, True)

    def async_pwrite(self, tensor, filename):
        return self.pwrite(tensor, filename, False, True)


def aio_read(tensor, filename, block_size=1048576, queue_depth=8, single_submit=False, overlap_events=True, validate=False):
    r = _lib().dsb_aio_sync_rw(1, tensor.data_ptr(), tensor.numel() * tensor.element_size(), str(filename).encode(),
                               block_size, queue_depth, int(single_submit), int(overlap_events))
    if r < 0:
        raise OSError(-r, f"aio_read failed for {filename}")
    return int(r)


def aio_write(tensor, filename, block_size=1048576, queue_depth=8, single_submit=False, overlap_events=True, validate=False):
    r = _lib().dsb_aio_sync_rw(0, tensor.data_ptr(), tensor.numel() * tensor.element_size(), str(filename).encode(),
                               block_size, queue_depth, int(single_submit), int(overlap_events))
    if r < 0:
        raise OSError(-r, f"aio_write failed for {filename}")
    return int(r)


def deepspeed_memcpy(dest, src):
    assert dest.is_contiguous() and src.is_contiguous() and not dest.is_cuda and not src.is_cuda
    n = min(dest.numel() * dest.element_size(), src.numel() * src.element_size())
    return _lib().dsb_parallel_memcpy(dest.data_ptr(), src.data_ptr(), n)


class AsyncIOBuilder:
    """`AsyncIOBuilder().load()` returns this module-like object (reference op_builder/async_io.py)."""
    NAME = "async_io"

    def is_compatible(self):
        return N.available() or True

    def load(self, verbose=True):
        import sys
        return sys.modules[__name__]
