"""Fused all-gather + GEMM for ZeRO-3 linears: y = x @ W^T (+ bias) where W [N, K] is row-sharded over the ranks of a
symmetric heap. The tcgen05 GEMM's TMA producer reads each weight tile directly from the owning rank's shard over
NVLink (one tensor map per rank), so the gathered weight never exists in memory
(csrc/gemm/gemm_tcgen05.cu: dsb_gemm_gather_b). Replaces `_allgather_param` + F.linear
(reference deepspeed/runtime/zero/partition_parameters.py:746-855, zero/linear.py:29-99)."""
import ctypes as C

import torch
import torch.distributed as dist

from ..ops import kernels as K
from ..ops import native as N
from .symm import SymmetricHeap

_p, _i, _l, _f = C.c_void_p, C.c_int, C.c_long, C.c_float
N.register_sigs({
    "dsb_gemm_gather_b": [_p, _p, _i, _p, _i, _i, _i, _l, _l, _l, _i, _i, _f, _p, _p, _l, _p],
    "dsb_gemm_ag_overlap": [_p, _p, _i, _i, _p, _p, _p, _i, _i, _i, _l, _l, _i, _i, _f, _p, _p, _l, _p],
})


class ShardedWeight:
    """A weight [N, K] whose rows are split evenly over the ranks of `group`; every shard lives at the same offset of a
    symmetric heap, so each rank knows the device address of every other shard."""

    def __init__(self, full_weight_or_shape, dtype=torch.float16, group=None, heap=None, init_from_rank0=True):
        self.group = group
        self.world = dist.get_world_size(group=group)
        self.rank = dist.get_rank(group=group)
        if torch.is_tensor(full_weight_or_shape):
            Nn, Kk = full_weight_or_shape.shape
            dtype = full_weight_or_shape.dtype
        else:
            Nn, Kk = full_weight_or_shape
        assert Nn % self.world == 0, "rows must split evenly over the ranks"
        self.N, self.K = Nn, Kk
        self.rows = Nn // self.world
        self.heap = heap or SymmetricHeap(self.rows * Kk * 2 + 4096, group=group)
        self.shard, off = self.heap.alloc(self.rows * Kk, dtype)
        self.shard = self.shard.view(self.rows, Kk)
        self.peer_ptrs = self.heap.peer_ptrs(off)
        if torch.is_tensor(full_weight_or_shape):
            self.shard.copy_(full_weight_or_shape[self.rank * self.rows:(self.rank + 1) * self.rows])
        torch.cuda.synchronize()
        dist.barrier(group=group)

    def supported(self):
        return self.rows % 256 == 0 and self.K % 8 == 0 and self.world <= 8


def ag_linear(x, sw, bias=None, direct=False):
    """x [.., K] @ gather(sw)^T (+ bias) with the all-gather inside the GEMM kernel.

    Default: two warps of every CTA stream the remote shards into a local scratch panel by panel while the tensor
    pipe works on the panels already present (own shard first); every weight byte crosses NVLink once and the scratch
    is shared by all layers (`dsb_gemm_ag_overlap`). direct=True keeps the first design (TMA loads every remote tile
    straight from the peer; only sensible for a handful of rows since tiles are re-fetched per M block)."""
    x2 = x.reshape(-1, x.shape[-1]).contiguous()
    M = x2.shape[0]
    out = torch.empty(M, sw.N, dtype=x.dtype, device=x.device)
    ptrs = (C.c_void_p * sw.world)(*sw.peer_ptrs)
    if not direct:
        gathered = N.scratch(sw.N * sw.K * 2, x.device, "ag_gather")
        flags = N.scratch(4 * (sw.N // 256), x.device, "ag_flags")
        rc = N.cuda_lib().dsb_gemm_ag_overlap(N.ptr(x2), ptrs, sw.world, sw.rank, N.ptr(gathered), N.ptr(flags),
                                              N.ptr(out), M, sw.N, sw.K, x2.stride(0), out.stride(0),
                                              N.dtype_code(x.dtype), N.EPI_BIAS if bias is not None else 0, 1.0,
                                              N.ptr(bias), None, 0, N.stream_ptr())
        N.check(rc, "dsb_gemm_ag_overlap")
        K._count()
        return out.view(*x.shape[:-1], sw.N)
    rc = N.cuda_lib().dsb_gemm_gather_b(N.ptr(x2), ptrs, sw.world, N.ptr(out), M, sw.N, sw.K, x2.stride(0), sw.K,
                                        out.stride(0), N.dtype_code(x.dtype), N.EPI_BIAS if bias is not None else 0,
                                        1.0, N.ptr(bias), None, 0, N.stream_ptr())
    N.check(rc, "dsb_gemm_gather_b")
    K._count()
    return out.view(*x.shape[:-1], sw.N)
