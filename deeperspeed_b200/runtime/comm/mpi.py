"""MpiBackend: the error-compensated 1-bit all-reduce over mpi4py (reference deepspeed/runtime/comm/mpi.py:14-290).

mpi4py is optional; without it the constructor raises with a pointer to NcclBackend. Two transports, chosen by
`cuda_aware` as in the reference (`gather_cuda` / `gather_host`, reference mpi.py:22-123):

  * cuda_aware=True  - the packed sign bytes and the scales stay on the GPU and the device tensors themselves are handed
    to MPI (mpi4py >= 3.1 takes any object with `__cuda_array_interface__` / DLPack, which torch tensors have); the
    producing stream is synchronised first because MPI knows nothing about CUDA streams;
  * cuda_aware=False - the buffers are staged through pinned host memory (one pinned pair per message size, reused).

The exchange is one Alltoall of sign chunks + one Allgather of scales per phase instead of the reference's W rooted
gathers of non-blocking sends: same bytes, W times fewer messages.
"""
import numpy as np
import torch

from ..compression.cupy import packbits, unpackbits


class MpiBackend:
    def __init__(self, cuda_aware, comm=None):
        if comm is None:
            try:
                from mpi4py import MPI
            except ImportError as e:
                raise ImportError(
                    "MpiBackend needs mpi4py; use deeperspeed_b200.runtime.comm.nccl.NcclBackend instead") from e
            comm = MPI.COMM_WORLD
        self.comm = comm
        self.rank = self.comm.Get_rank()
        self.size = self.comm.Get_size()
        self.cuda_aware = bool(cuda_aware)
        self._pinned = {}

    # ---- transports ---------------------------------------------------------------------------------------------
    def _host_pair(self, tag, send_like, recv_numel):
        """Reusable (send, recv) pinned host tensors for one message kind."""
        key = (tag, send_like.numel(), recv_numel, send_like.dtype)
        pair = self._pinned.get(key)
        if pair is None:
            pin = send_like.is_cuda
            pair = (torch.empty(send_like.numel(), dtype=send_like.dtype, pin_memory=pin),
                    torch.empty(recv_numel, dtype=send_like.dtype, pin_memory=pin))
            self._pinned[key] = pair
        return pair

    def _exchange(self, tag, collective, t, recv_numel):
        t = t.detach().contiguous().view(-1)
        if self.cuda_aware:
            recv = torch.empty(recv_numel, dtype=t.dtype, device=t.device)
            if t.is_cuda:
                torch.cuda.current_stream(t.device).synchronize()      # MPI reads the buffer outside stream order
            collective(t, recv)
            return recv
        send_h, recv_h = self._host_pair(tag, t, recv_numel)
        send_h.copy_(t)                                                 # blocking device -> pinned host copy
        collective(send_h.numpy(), recv_h.numpy())
        return recv_h.to(t.device, non_blocking=False)

    def _alltoall(self, t, tag="a2a"):
        return self._exchange(tag, self.comm.Alltoall, t, t.numel())

    def _allgather(self, t, tag="ag"):
        return self._exchange(tag, self.comm.Allgather, t, t.numel() * self.size)

    # ---- the algorithm -----------------------------------------------------------------------------------------
    @staticmethod
    def _signs(bits, rows, cols, dtype):
        return unpackbits(bits).view(rows, cols).to(dtype) * 2 - 1

    def compressed_allreduce(self, buffer_m: torch.tensor, worker_error, server_error, local_rank):
        shape = buffer_m.size()
        flat = torch.flatten(buffer_m) if len(shape) > 1 else buffer_m
        numel, padded = flat.numel(), worker_error.numel()
        if numel != padded:
            flat = torch.cat([flat, torch.zeros(padded - numel, device=flat.device, dtype=flat.dtype)])
        W, chunk = self.size, padded // self.size

        # worker phase: compensate, 1-bit quantise with one scale, remember what the quantisation lost
        flat.add_(worker_error)
        w_scale = torch.norm(flat) / (padded ** 0.5)
        w_pos = flat >= 0
        worker_error.set_(flat - w_scale * (w_pos.to(flat.dtype) * 2 - 1))
        got_bits = self._alltoall(packbits(w_pos), "worker_bits")      # chunk r of every rank lands on rank r
        got_scales = self._allgather(w_scale.reshape(1), "worker_scale")

        # server phase: average my chunk over the workers, compensate, quantise again
        mine = (self._signs(got_bits, W, chunk, flat.dtype) * got_scales.view(W, 1)).sum(0) / W + server_error
        s_scale = torch.norm(mine) / (chunk ** 0.5)
        s_pos = mine >= 0
        server_error.set_(mine - s_scale * (s_pos.to(mine.dtype) * 2 - 1))
        all_bits = self._allgather(packbits(s_pos), "server_bits")
        all_scales = self._allgather(s_scale.reshape(1), "server_scale")

        flat.data.copy_((self._signs(all_bits, W, chunk, flat.dtype) * all_scales.view(W, 1)).view(-1))
        out = flat[:numel] if numel != padded else flat
        return out.reshape(shape) if len(shape) > 1 else out
