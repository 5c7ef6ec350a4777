"""Symmetric heap: every rank of a process group allocates the same number of bytes and maps all peers, so kernels can
load from / store to any peer's memory over NVLink / NVSwitch - and, on the VMM backend, reduce in the switch through
the heap's multicast address (NVLS).
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


def _group_name(group):
    return (group if group is not None else dist.group.WORLD).group_name


class SymmetricHeap:
    """`backend`: "vmm" = torch symmetric memory (cuMem VMM allocation bound to an NVSwitch multicast object: besides
    the per-peer addresses there is ONE multicast address, `mc_base`, on which multimem.ld_reduce / multimem.st work -
    NVLS), "ipc" = cudaMalloc + CUDA IPC handles (peer addresses only, mc_base == 0), "auto" = vmm if it can be set up.
    DSB_SYMM_BACKEND overrides."""

    def __init__(self, nbytes, group=None, backend="auto"):
        assert torch.cuda.is_available()
        self.group = group
        self.world = dist.get_world_size(group=group)
        self.rank = dist.get_rank(group=group)
        self.nbytes = (int(nbytes) + 255) // 256 * 256 + FLAG_BYTES
        self.mc_base = 0
        import os
        backend = os.environ.get("DSB_SYMM_BACKEND", backend)
        ok = backend in ("auto", "vmm") and self._init_vmm()
        if backend == "vmm" and not ok:
            raise RuntimeError("symmetric heap: torch symmetric memory / multicast is unavailable")
        # every rank must take the same branch: kernels address peers through one scheme
        agree = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=group)
        if ok and not int(agree.item()):
            ok, self.mc_base = False, 0
        self.backend = "vmm" if ok else "ipc"
        if not ok:
            self._init_ipc()
        self._bytes.narrow(0, 0, FLAG_BYTES).zero_()
        self._cursor = FLAG_BYTES
        self.epoch = 0
        torch.cuda.synchronize()
        dist.barrier(group=group)

    def _init_vmm(self):
        try:
            import torch.distributed._symmetric_memory as symm_mem
            name = _group_name(self.group)
            enable = getattr(symm_mem, "enable_symm_mem_for_group", None)
            if enable is not None:
                try:
                    import warnings
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore")
                        enable(name)
                except Exception:
                    pass
            buf = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=torch.device("cuda", torch.cuda.current_device()))
            hdl = symm_mem.rendezvous(buf, name)
            self._bytes, self._handle = buf, hdl
            self.base = buf.data_ptr()
            self.peer_base = [int(p) for p in hdl.buffer_ptrs]
            assert self.peer_base[self.rank] == self.base
            self.mc_base = int(hdl.multicast_ptr or 0)
            return True
        except Exception as e:  # older torch, no fabric / multicast support, ...
            from ..utils.logging import logger
            logger.warning(f"[symm] torch symmetric memory unavailable ({e!r}); using the CUDA IPC heap")
            return False

    def _init_ipc(self):
        lib = N.cuda_lib()
        base = C.c_void_p()
        N.check(lib.dsb_symm_alloc(self.nbytes, C.byref(base)), "dsb_symm_alloc")
        self.base = base.value
        handle = (C.c_ubyte * 64)()
        N.check(lib.dsb_ipc_get_handle(self.base, handle), "dsb_ipc_get_handle")
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle), group=self.group)
        self.peer_base = []
        for r, h in enumerate(handles):
            if r == self.rank:
                self.peer_base.append(self.base)
            else:
                buf = (C.c_ubyte * 64).from_buffer_copy(h)
                out = C.c_void_p()
                N.check(lib.dsb_ipc_open(buf, C.byref(out)), "dsb_ipc_open")
                self.peer_base.append(out.value)
        self._bytes = torch.as_tensor(_RawCudaBuffer(self.base, self.nbytes),
                                      device=torch.device("cuda", torch.cuda.current_device()))

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

    def mc_ptr(self, offset):
        """Multicast address of heap offset `offset` (0 when the heap has no multicast mapping)."""
        return self.mc_base + offset if self.mc_base else 0

    def next_epoch(self):
        self.epoch += 1
        return self.epoch
