"""AsyncPartitionedParameterSwapper: low-precision parameter shards <-> NVMe with a pinned buffer pool
(`buffer_count` x `buffer_size` elements) and separate read / write aio handles. Files live under
<nvme_path>/zero_stage_3/fp16params/rank<R>/ (reference deepspeed/runtime/swap_tensor/partitioned_param_swapper.py:36-308).

This is the reference's per-PARAMETER interface, kept for code that drives a swapper itself. The ZeRO-3 optimizer of
this package does not use it: it swaps whole sub-group arenas through `arena_swapper.ArenaSwapper` (one file and one
sequential transfer per sub-group)."""
import os
from enum import Enum

import torch
import torch.distributed as dist

from ...utils.logging import logger
from .optimizer_utils import make_aio_handle
from .constants import AIO_ALIGNED_BYTES, AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT, AIO_BLOCK_SIZE, AIO_BLOCK_SIZE_DEFAULT, MIN_AIO_BYTES


class PartitionedParamStatus(Enum):
    AVAILABLE = 1       # shard resident in a pool buffer
    NOT_AVAILABLE = 2   # shard only on NVMe
    INFLIGHT = 3        # read issued, not waited


class AsyncPartitionedParameterSwapper:
    def __init__(self, ds_config=None, dtype=torch.float16, nvme_path=None, buffer_count=5, buffer_size=int(1e8),
                 aio_config=None):
        if ds_config is not None:
            off = ds_config.zero_config.offload_param
            nvme_path, buffer_count, buffer_size = off["nvme_path"], off["buffer_count"], off["buffer_size"]
            aio_config = ds_config.aio_config
        self.dtype = dtype
        rank = dist.get_rank() if dist.is_initialized() else 0
        self.swap_folder = os.path.join(nvme_path, "zero_stage_3", f"{'fp16' if dtype != torch.float32 else 'fp32'}params", f"rank{rank}")
        os.makedirs(self.swap_folder, exist_ok=True)
        self.swap_element_size = torch.tensor([], dtype=dtype).element_size()
        threads = (aio_config or {}).get(AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT)
        self.min_aio_bytes = max(MIN_AIO_BYTES, (aio_config or {}).get(AIO_BLOCK_SIZE, AIO_BLOCK_SIZE_DEFAULT))
        self.aligned_bytes = AIO_ALIGNED_BYTES * threads
        self.numel_alignment = self.aligned_bytes // self.swap_element_size
        self.elements_per_buffer = int(buffer_size)
        self.aligned_elements_per_buffer = self._io_aligned_numel(self.elements_per_buffer)
        self.param_buffer_count = int(buffer_count)
        pool = torch.zeros(self.aligned_elements_per_buffer * self.param_buffer_count, dtype=dtype, device="cpu")
        self.buffers = pool.pin_memory() if torch.cuda.is_available() else pool
        self.available_buffer_ids = list(range(self.param_buffer_count))
        self.param_id_to_buffer_id = {}
        self.param_id_to_swap_buffer = {}
        self.param_id_to_numel = {}
        self.id_to_path = {}
        self.inflight_params = []
        self.inflight_numel = 0
        self.available_params = set()
        self.available_numel = 0
        self.aio_read_handle = make_aio_handle(aio_config)
        self.aio_write_handle = make_aio_handle(aio_config)
        self.pending_writes = 0
        self.pending_reads = 0
        self.invalid_buffer = torch.tensor(1).to(dtype)

    def _io_aligned_numel(self, numel):
        remainder = numel % self.numel_alignment
        return numel if remainder == 0 else (numel + self.numel_alignment - remainder)

    def swappable_tensor(self, param=None, numel=None):
        n = param.ds_tensor.ds_numel if param is not None else numel
        return self.min_aio_bytes <= n * self.swap_element_size

    def available_swap_in_buffers(self):
        return len(self.available_buffer_ids)

    def get_path(self, param, must_exist=False):
        pid = param.ds_id
        if pid not in self.id_to_path:
            assert not must_exist, f"Path for param id {pid} does not exist"
            self.id_to_path[pid] = os.path.join(self.swap_folder, f"{pid}_param.tensor.swp")
        return self.id_to_path[pid]

    def _allocate_buffer(self, param):
        pid = param.ds_id
        assert len(self.available_buffer_ids) > 0, "No buffer available for swapping: increase offload_param.buffer_count"
        numel = param.ds_tensor.ds_numel
        assert self._io_aligned_numel(numel) <= self.aligned_elements_per_buffer, "offload_param.buffer_size too small"
        bid = self.available_buffer_ids.pop()
        self.param_id_to_buffer_id[pid] = bid
        self.param_id_to_numel[pid] = numel
        swap = self.buffers.narrow(0, bid * self.aligned_elements_per_buffer, self._io_aligned_numel(numel))
        self.param_id_to_swap_buffer[pid] = swap
        return swap, swap.narrow(0, 0, numel)

    def get_buffer(self, param, numel):
        swap, compute = self._allocate_buffer(param)
        return compute

    def swap_out_and_release(self, params, async_op=False, force_buffer_release=False):
        """Write resident shards to NVMe and give their pool buffers back."""
        for p in params:
            pid = p.ds_id
            if pid not in self.param_id_to_swap_buffer:
                swap, compute = self._allocate_buffer(p)
                compute.copy_(p.ds_tensor.data.view(-1))
            self.aio_write_handle.async_pwrite(self.param_id_to_swap_buffer[pid], self.get_path(p))
            self.pending_writes += 1
        if not async_op:
            self.synchronize_writes()
            self._release(params)

    def _release(self, params):
        for p in params:
            pid = p.ds_id
            if pid in self.param_id_to_buffer_id:
                self.available_buffer_ids.append(self.param_id_to_buffer_id.pop(pid))
                self.param_id_to_swap_buffer.pop(pid, None)
                if pid in self.available_params:
                    self.available_params.remove(pid)
                    self.available_numel -= self.param_id_to_numel[pid]
            p.ds_tensor.data = self.invalid_buffer.data
            p.ds_tensor.status = PartitionedParamStatus.NOT_AVAILABLE

    def remove_partition_and_release_buffers(self, params):
        self._release(params)

    def synchronize_writes(self):
        if self.pending_writes:
            assert self.pending_writes == self.aio_write_handle.wait()
            self.pending_writes = 0

    def swap_in(self, params, async_op=True, swap_in_buffers=None):
        for p in params:
            assert getattr(p.ds_tensor, "status", PartitionedParamStatus.NOT_AVAILABLE) == PartitionedParamStatus.NOT_AVAILABLE, \
                "Some params are already available or in flight"
            swap, compute = self._allocate_buffer(p)
            self.aio_read_handle.async_pread(swap, self.get_path(p, must_exist=True))
            self.pending_reads += 1
            p.ds_tensor.data = compute.data
            p.ds_tensor.status = PartitionedParamStatus.INFLIGHT
            self.inflight_params.append(p.ds_id)
            self.inflight_numel += self.param_id_to_numel[p.ds_id]
        self._inflight_objs = list(params) + getattr(self, "_inflight_objs", [])
        if not async_op:
            self.synchronize_reads()

    def synchronize_reads(self):
        if self.pending_reads:
            assert self.pending_reads == self.aio_read_handle.wait()
            self.pending_reads = 0
        for p in getattr(self, "_inflight_objs", []):
            p.ds_tensor.status = PartitionedParamStatus.AVAILABLE
            self.available_params.add(p.ds_id)
        self._inflight_objs = []
        self.available_numel += self.inflight_numel
        self.inflight_params = []
        self.inflight_numel = 0
