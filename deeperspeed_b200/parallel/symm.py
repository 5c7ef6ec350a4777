"""Symmetric heap over CUDA IPC: every rank of a process group allocates the same number of bytes, exchanges IPC
handles once and maps all peers, so kernels can load from / store to any peer's memory over NVLink / NVSwitch.
Layout of each rank's heap: [flag block (4 KiB) | user region]. Native side: csrc/comm/fused_zero.cu (dsb_symm_*,
dsb_ipc_*). This is the data plane SURVEY 5.8 asks for; torch.distributed stays the control plane."""
import ctypes as C

import torch
import torch.distributed as dist

from ..ops import native as N

FLAG_BYTES = 4096
_p, _i, _l = C.c_void_p, C.c_int, C.c_long

N.register_sigs({
    "dsb_symm_alloc": [_l, C.POINTER(C.c_void_p)],
    "dsb_symm_free": [_p],
    "dsb_ipc_get_handle": [_p, _p],
    "dsb_ipc_open": [_p, C.POINTER(C.c_void_p)],
    "dsb_ipc_close": [_p],
})


class _RawCudaBuffer:
    """Minimal __cuda_array_interface__ holder so torch can view memory we allocated with cudaMalloc."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
                                         "strides": None}


class SymmetricHeap:
    def __init__(self, nbytes, group=None):
        assert torch.cuda.is_available()
        self.group = group
        self.world = dist.get_world_size(group=group)
        self.rank = dist.get_rank(group=group)
        self.nbytes = (int(nbytes) + 255) // 256 * 256 + FLAG_BYTES
        lib = N.cuda_lib()
        base = C.c_void_p()
        N.check(lib.dsb_symm_alloc(self.nbytes, C.byref(base)), "dsb_symm_alloc")
        self.base = base.value
        handle = (C.c_ubyte * 64)()
        N.check(lib.dsb_ipc_get_handle(self.base, handle), "dsb_ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=group)
        self.peer_base = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.peer_base.append(self.base)
            else:
                buf = (C.c_ubyte * 64).from_buffer_copy(h)
                out = C.c_void_p()
                N.check(lib.dsb_ipc_open(buf, C.byref(out)), "dsb_ipc_open")
                self.peer_base.append(out.value)
        self._bytes = torch.as_tensor(_RawCudaBuffer(self.base, self.nbytes), device=torch.device("cuda", torch.cuda.current_device()))
        self._cursor = FLAG_BYTES
        self.epoch = 0
        dist.barrier(group=group)

    def flag_ptrs(self):
        return list(self.peer_base)

    def alloc(self, numel, dtype):
        """Carve a tensor out of the local heap; the same call sequence on every rank yields symmetric offsets."""
        esize = torch.empty((), dtype=dtype).element_size()
        nbytes = (numel * esize + 255) // 256 * 256
        assert self._cursor + nbytes <= self.nbytes, "symmetric heap exhausted"
        off = self._cursor
        self._cursor += nbytes
        t = self._bytes.narrow(0, off, numel * esize).view(dtype)
        t.zero_()
        return t, off

    def peer_ptrs(self, offset):
        return [b + offset for b in self.peer_base]

    def next_epoch(self):
        self.epoch += 1
        return self.epoch
