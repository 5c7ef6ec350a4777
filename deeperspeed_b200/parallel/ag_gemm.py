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
    "dsb_gemm_ag_overlap_k": [_p, _p, _i, _i, _p, _p, _p, _i, _i, _i, _l, _l, _i, _f, _p],
    "dsb_gemm_reduce_scatter": [_p, _p, _p, _p, _p, _i, _i, _p, _i, _i, _i, _l, _l, _i, _i, _i, _f, _f, C.c_uint, _p],
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


def ag_linear_dgrad(grad_out, sw):
    """grad_out [.., N_out] @ gather(sw) -> [.., K_in]: the input gradient of a linear whose weight [N_out, K_in] is
    row-sharded, with the all-gather inside the GEMM kernel (`dsb_gemm_ag_overlap_k`: the shards split the reduction
    dimension, every tile starts on the own shard's range while copier warps bring in the others)."""
    g2 = grad_out.reshape(-1, grad_out.shape[-1])
    if not g2.is_contiguous():
        g2 = g2.contiguous()
    M = g2.shape[0]
    out = torch.empty(M, sw.K, dtype=g2.dtype, device=g2.device)
    ptrs = (C.c_void_p * sw.world)(*sw.peer_ptrs)
    gathered = N.scratch(sw.N * sw.K * 2, g2.device, "ag_gather")
    flags = N.scratch(4 * (sw.N // 256), g2.device, "ag_flags")
    rc = N.cuda_lib().dsb_gemm_ag_overlap_k(N.ptr(g2), ptrs, sw.world, sw.rank, N.ptr(gathered), N.ptr(flags),
                                            N.ptr(out), M, sw.K, sw.N, g2.stride(0), out.stride(0),
                                            N.dtype_code(g2.dtype), 1.0, N.stream_ptr())
    N.check(rc, "dsb_gemm_ag_overlap_k")
    K._count()
    return out.view(*grad_out.shape[:-1], sw.K)


class GemmReduceScatter:
    """GEMM followed by a reduce-scatter of its result, in ONE kernel (csrc/gemm/gemm_tcgen05.cu:
    dsb_gemm_reduce_scatter): every rank computes its local product D [M, N]; each output tile is TMA-stored into
    the row owner's staging buffer over NVLink while the tensor pipe works on the next tiles, then every rank sums
    the slots that arrived in its own staging. Used for the weight gradient of ZeRO-3 linears (reference:
    `_reduce_scatter_gradient`, deepspeed/runtime/zero/partition_parameters.py:888-926, after the wgrad GEMM).

    Two staging buffers of `max_elems` elements alternate from call to call, so a peer that is still summing the
    previous call's slots is never overwritten."""

    def __init__(self, max_elems, dtype=torch.float16, group=None):
        self.group = group
        self.world = dist.get_world_size(group=group)
        self.rank = dist.get_rank(group=group)
        self.max_elems = int(max_elems)
        self.dtype = dtype
        self.heap = SymmetricHeap(2 * (self.max_elems * 2 + 512) + 4096, group=group)
        self.stage_ptrs = []
        for _ in range(2):
            _, off = self.heap.alloc(self.max_elems, dtype)
            self.stage_ptrs.append((C.c_void_p * self.world)(*self.heap.peer_ptrs(off)))
        self.flag_ptrs = (C.c_void_p * self.world)(*self.heap.flag_ptrs())
        self.sync = torch.zeros(2, dtype=torch.int32, device="cuda")
        self.calls = 0
        torch.cuda.synchronize()
        dist.barrier(group=group)

    def supported(self, M, Nn, K):
        return (M % self.world == 0 and (M // self.world) % 128 == 0 and Nn % 8 == 0 and M * Nn <= self.max_elems
                and self.world <= 8)

    def __call__(self, a, b, trans_a=False, trans_b=True, average=True, out=None):
        """out [M / world, N] = (sum over ranks of op(a) @ op(b)) rows owned by this rank (/ world if average).
        Layouts as ops.kernels.gemm: trans_a=True: a is [K, M]; trans_b=False: b is [K, N]."""
        if trans_a:
            Kk, M = a.shape
        else:
            M, Kk = a.shape
        Nn = b.shape[0] if trans_b else b.shape[1]
        assert self.supported(M, Nn, Kk), (M, Nn, Kk, self.max_elems)
        assert a.stride(-1) == 1 and b.stride(-1) == 1 and a.dtype == self.dtype and b.dtype == self.dtype
        rows = M // self.world
        if out is None:
            out = torch.empty(rows, Nn, dtype=a.dtype, device=a.device)
        stage = self.stage_ptrs[self.calls & 1]
        self.calls += 1
        rc = N.cuda_lib().dsb_gemm_reduce_scatter(
            N.ptr(a), N.ptr(b), N.ptr(out), stage, self.flag_ptrs, self.world, self.rank, N.ptr(self.sync), M, Nn, Kk,
            a.stride(0), b.stride(0), int(trans_a), int(not trans_b), N.dtype_code(a.dtype), 1.0,
            1.0 / self.world if average else 1.0, self.heap.next_epoch(), N.stream_ptr())
        N.check(rc, "dsb_gemm_reduce_scatter")
        K._count()
        return out
