"""NcclBackend.compressed_allreduce: error-compensated 1-bit all-reduce (math in SURVEY C.7).

On CUDA the whole algorithm is ONE cooperative kernel that packs signs with warp ballots and exchanges words and
scales through NVLink peer stores into a symmetric heap (csrc/comm/onebit.cu). The torch.distributed restatement
below (`_compressed_allreduce_torch`) is the measurement baseline and the CPU/gloo path.
API per reference deepspeed/runtime/comm/nccl.py:14-186 (which needs cupy; this one does not)."""
import ctypes as C

import torch
import torch.distributed as dist

from ...ops import native as N
from ..compression.cupy import packbits, unpackbits

_p, _i, _l, _u = C.c_void_p, C.c_int, C.c_long, C.c_uint
N.register_sigs({"dsb_onebit_allreduce": [_p, _p, _p, _l, _i, _i, _p, _l, _p, _p, _u, _p]})
N.register_sigs({"dsb_onebit_heap_bytes": [_l, _i]}, restype=_l)
_f = C.c_float
N.register_sigs({"dsb_onebit_adam_step": [_p, _p, _p, _p, _p, _p, _l, _l, _i, _i, _p, _l, _p, _p, _u, _f, _f, _f, _f, _p]})


class NcclBackend:
    def __init__(self, mpu=None, use_fused=True):
        if mpu is None:
            self.world_group = dist.new_group(ranks=range(dist.get_world_size()))
        else:
            self.mpu = mpu
            self.world_group = self.mpu.get_data_parallel_group()
        self.rank = dist.get_rank(group=self.world_group)
        self.size = dist.get_world_size(group=self.world_group)
        self.compression_backend = None
        self.use_fused = use_fused
        self._heaps = {}
        self._scratch = None

    def my_igather(self, rank, size, group, sendbuf, recvbuf, root):
        req = []
        if rank == root:
            for idx in range(size):
                if idx != rank:
                    req.append(dist.irecv(recvbuf[idx], src=idx, group=group))
                else:
                    recvbuf[rank] = sendbuf
        else:
            req.append(dist.isend(sendbuf, group=group, dst=root))
        return req

    def my_gather(self, rank, size, group, sendbuf, recvbuf, root):
        if rank == root:
            for idx in range(size):
                if idx != rank:
                    dist.recv(recvbuf[idx], src=idx, group=group)
                else:
                    recvbuf[rank] = sendbuf
        else:
            dist.send(sendbuf, group=group, dst=root)

    # ------------------------------------------------------------------ fused sm_100a path
    def _fused_ok(self, t):
        return (self.use_fused and t.is_cuda and t.dtype == torch.float32 and self.size > 1 and self.size <= 16
                and t.numel() % (32 * self.size) == 0)

    def _heap_for(self, n):
        if n not in self._heaps:
            from ...parallel.symm import SymmetricHeap
            lib = N.cuda_lib()
            heap = SymmetricHeap(lib.dsb_onebit_heap_bytes(n, self.size), group=self.world_group)
            self._heaps[n] = heap
        return self._heaps[n]

    def _compressed_allreduce_fused(self, buffer_m, worker_error, server_error):
        from ...parallel.symm import FLAG_BYTES
        n = buffer_m.numel()
        heap = self._heap_for(n)
        lib = N.cuda_lib()
        if self._scratch is None:
            self._scratch = torch.zeros(4096, dtype=torch.float32, device=buffer_m.device)
        bases = (C.c_void_p * self.size)(*heap.peer_base)
        rc = lib.dsb_onebit_allreduce(N.ptr(buffer_m), N.ptr(worker_error), N.ptr(server_error), n, self.size,
                                      self.rank, bases, FLAG_BYTES, bases, N.ptr(self._scratch), heap.next_epoch(),
                                      N.stream_ptr())
        N.check(rc, "dsb_onebit_allreduce")
        from ...ops import kernels as K
        K._count()
        return buffer_m

    def phase_times_ms(self):
        """Phase durations of the last fused launch from its in-kernel timestamps: A (compensate + norm), B (quantise +
        scatter), wait for the peers, C + D (server average, requantise, broadcast), wait, E (decompress / update)."""
        torch.cuda.synchronize()
        t = self._scratch[1024:1024 + 14].view(torch.int64).tolist()
        names = ["A_norm", "B_pack_scatter", "barrier_1", "CD_server", "barrier_2", "E_unpack"]
        return {n: round((t[i + 1] - t[i]) / 1e6, 4) for i, n in enumerate(names)}

    def fused_adam_supported(self, momentum, worker_error):
        """Can `compressed_adam_step` run for this momentum tensor (fp32 on a GPU, length a multiple of 4, error buffers
        padded to whole sign words per server chunk)?"""
        n = worker_error.numel()
        return (self.use_fused and momentum.is_cuda and momentum.dtype == torch.float32 and 1 < self.size <= 16
                and momentum.is_contiguous() and momentum.numel() % 4 == 0 and momentum.numel() <= n
                and n % (32 * self.size) == 0)

    def compressed_adam_step(self, momentum, grad, param, variance, worker_error, server_error, beta1, lr, eps,
                             weight_decay):
        """1-bit Adam's compression-stage step as ONE kernel (csrc/comm/onebit.cu): momentum <- beta1 * momentum +
        (1 - beta1) * grad, error-compensated 1-bit average over the group, param <- param - lr * (momentum /
        (sqrt(variance) + eps) + weight_decay * param). All tensors fp32 and flat."""
        from ...parallel.symm import FLAG_BYTES
        n = worker_error.numel()
        heap = self._heap_for(n)
        lib = N.cuda_lib()
        if self._scratch is None:
            self._scratch = torch.zeros(4096, dtype=torch.float32, device=momentum.device)
        bases = (C.c_void_p * self.size)(*heap.peer_base)
        rc = lib.dsb_onebit_adam_step(N.ptr(momentum), N.ptr(grad), N.ptr(param), N.ptr(variance), N.ptr(worker_error),
                                      N.ptr(server_error), n, momentum.numel(), self.size, self.rank, bases, FLAG_BYTES,
                                      bases, N.ptr(self._scratch), heap.next_epoch(), float(beta1), float(lr),
                                      float(eps), float(weight_decay), N.stream_ptr())
        N.check(rc, "dsb_onebit_adam_step")
        from ...ops import kernels as K
        K._count()

    # ------------------------------------------------------------------ torch.distributed restatement
    def _compressed_allreduce_torch(self, buffer_m, worker_error, server_error):
        W = self.size
        n = buffer_m.numel()
        c = n // W
        buffer_m.add_(worker_error)
        worker_scale = torch.norm(buffer_m) / (n ** 0.5)
        pos = buffer_m >= 0
        sgn = pos.to(buffer_m.dtype) * 2 - 1
        worker_error.set_(buffer_m - worker_scale * sgn)
        packed = packbits(pos)                               # n / 8 bytes
        recv = torch.empty_like(packed)
        if W > 1:
            try:
                dist.all_to_all_single(recv, packed, group=self.world_group)
            except Exception:
                gathered = [torch.empty_like(packed) for _ in range(W)]
                dist.all_gather(gathered, packed, group=self.world_group)
                cb = packed.numel() // W
                recv = torch.cat([g[self.rank * cb:(self.rank + 1) * cb] for g in gathered])
            scales = [torch.empty_like(worker_scale) for _ in range(W)]
            dist.all_gather(scales, worker_scale, group=self.world_group)
        else:
            recv = packed
            scales = [worker_scale]
        bits = unpackbits(recv).view(W, c).to(buffer_m.dtype) * 2 - 1
        y = (bits * torch.stack(scales).view(W, 1)).sum(0) / W
        y.add_(server_error)
        server_scale = torch.norm(y) / (c ** 0.5)
        spos = y >= 0
        server_error.set_(y - server_scale * (spos.to(y.dtype) * 2 - 1))
        spacked = packbits(spos)
        if W > 1:
            allp = [torch.empty_like(spacked) for _ in range(W)]
            dist.all_gather(allp, spacked, group=self.world_group)
            alls = [torch.empty_like(server_scale) for _ in range(W)]
            dist.all_gather(alls, server_scale, group=self.world_group)
        else:
            allp, alls = [spacked], [server_scale]
        out = (unpackbits(torch.cat(allp)).view(W, c).to(buffer_m.dtype) * 2 - 1) * torch.stack(alls).view(W, 1)
        buffer_m.data.copy_(out.view(-1))
        return buffer_m

    def compressed_allreduce(self, buffer_m: torch.tensor, worker_error, server_error, local_rank):
        original_shape = buffer_m.size()
        if len(original_shape) > 1:
            buffer_m = torch.flatten(buffer_m)
        original_size = buffer_m.numel()
        worker_error_size = worker_error.numel()
        if original_size != worker_error_size:
            empty_tensor = torch.zeros(worker_error_size - original_size, device=buffer_m.device)
            buffer_m = torch.cat([buffer_m, empty_tensor])
        if self._fused_ok(buffer_m):
            res = self._compressed_allreduce_fused(buffer_m.contiguous(), worker_error, server_error)
        else:
            res = self._compressed_allreduce_torch(buffer_m, worker_error, server_error)
        if original_size != worker_error_size:
            res = res[0:original_size]
        if len(original_shape) > 1:
            res = res.reshape(original_shape)
        return res
