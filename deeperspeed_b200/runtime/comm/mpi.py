"""MpiBackend: the same error-compensated 1-bit all-reduce over mpi4py (reference deepspeed/runtime/comm/mpi.py:14-290).
mpi4py is optional; without it the constructor raises with a pointer to NcclBackend. The exchange uses Alltoall /
Allgather on host-staged buffers (cuda_aware=True passes device buffers straight to a CUDA-aware MPI)."""
import numpy as np
import torch

from ..compression.cupy import packbits, unpackbits


class MpiBackend:
    def __init__(self, cuda_aware):
        try:
            from mpi4py import MPI
        except ImportError as e:
            raise ImportError("MpiBackend needs mpi4py; use deeperspeed_b200.runtime.comm.nccl.NcclBackend instead") from e
        self.MPI = MPI
        self.comm = MPI.COMM_WORLD
        self.rank = self.comm.Get_rank()
        self.size = self.comm.Get_size()
        self.cuda_aware = cuda_aware

    def _alltoall(self, t):
        send = t.detach().cpu().numpy()
        recv = np.empty_like(send)
        self.comm.Alltoall(send, recv)
        return torch.from_numpy(recv).to(t.device)

    def _allgather(self, t):
        send = t.detach().cpu().numpy().reshape(-1)
        recv = np.empty(send.size * self.size, dtype=send.dtype)
        self.comm.Allgather(send, recv)
        return torch.from_numpy(recv).to(t.device)

    def compressed_allreduce(self, buffer_m: torch.tensor, worker_error, server_error, local_rank):
        original_shape = buffer_m.size()
        if len(original_shape) > 1:
            buffer_m = torch.flatten(buffer_m)
        original_size = buffer_m.numel()
        n = worker_error.numel()
        if original_size != n:
            buffer_m = torch.cat([buffer_m, torch.zeros(n - original_size, device=buffer_m.device)])
        W, c = self.size, n // self.size
        buffer_m.add_(worker_error)
        ws = torch.norm(buffer_m) / (n ** 0.5)
        pos = buffer_m >= 0
        worker_error.set_(buffer_m - ws * (pos.to(buffer_m.dtype) * 2 - 1))
        recv = self._alltoall(packbits(pos))
        scales = self._allgather(ws.reshape(1))
        bits = unpackbits(recv).view(W, c).to(buffer_m.dtype) * 2 - 1
        y = (bits * scales.view(W, 1)).sum(0) / W + server_error
        ss = torch.norm(y) / (c ** 0.5)
        spos = y >= 0
        server_error.set_(y - ss * (spos.to(y.dtype) * 2 - 1))
        allp = self._allgather(packbits(spos))
        alls = self._allgather(ss.reshape(1))
        out = (unpackbits(allp).view(W, c).to(buffer_m.dtype) * 2 - 1) * alls.view(W, 1)
        buffer_m.data.copy_(out.view(-1))
        res = buffer_m[0:original_size] if original_size != n else buffer_m
        return res.reshape(original_shape) if len(original_shape) > 1 else res
