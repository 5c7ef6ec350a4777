"""ArenaSwapper: the low-precision parameter shards of ZeRO-3 on NVMe, one file per SUB-GROUP arena.

Role of the reference's AsyncPartitionedParameterSwapper (deepspeed/runtime/swap_tensor/partitioned_param_swapper.py:
36-308, one file and one pool buffer per parameter). Here the unit is what the rest of this ZeRO-3 already works in: the
flat low-precision arena of a sub-group (`stage3._add_sub_group`), so a read or a write is one large sequential
transfer - what NVMe and the async I/O engine are good at - and the optimizer step of a sub-group writes its file once.

    <nvme_path>/zero_stage_3/<fp16|fp32>params/rank<R>/arena_<idx>.swp

A pool of `buffer_count` pinned host buffers of `buffer_size` elements caches arenas. `acquire(idx)` makes an arena
resident (reading its file unless the caller is about to overwrite all of it), `release(idx)` lets it be evicted again
- least recently used first, never while somebody holds it -, `write_back(idx)` stores a modified arena, `prefetch(idx)`
starts a read without waiting. Whenever an arena moves in or out, the `attach` callback it was registered with re-points
the `ds_tensor` views of its parameters."""
import os
from collections import OrderedDict

import torch
import torch.distributed as dist

from .constants import AIO_ALIGNED_BYTES, AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT
from .optimizer_utils import make_aio_handle


class ArenaSwapper:
    _instances = {}       # folder -> swappers created on it in this process (a second engine gets its own files)

    def __init__(self, nvme_path, dtype, buffer_count, buffer_size, aio_config=None):
        rank = dist.get_rank() if dist.is_initialized() else 0
        kind = "fp32" if dtype == torch.float32 else "fp16"
        base = os.path.join(nvme_path, "zero_stage_3", f"{kind}params", f"rank{rank}")
        nth = ArenaSwapper._instances[base] = ArenaSwapper._instances.get(base, -1) + 1
        self.folder = base if nth == 0 else f"{base}_engine{nth}"
        os.makedirs(self.folder, exist_ok=True)
        self.dtype = dtype
        element = torch.tensor([], dtype=dtype).element_size()
        threads = (aio_config or {}).get(AIO_THREAD_COUNT, AIO_THREAD_COUNT_DEFAULT)
        self.alignment = AIO_ALIGNED_BYTES * threads // element          # O_DIRECT granularity, in elements
        self.slot_numel = self._aligned(int(buffer_size))
        self.slot_count = int(buffer_count)
        pool = torch.zeros(self.slot_numel * self.slot_count, dtype=dtype, device="cpu")
        self.pool = pool.pin_memory() if torch.cuda.is_available() else pool
        self.free_slots = list(range(self.slot_count))
        self.read_handle, self.write_handle = make_aio_handle(aio_config), make_aio_handle(aio_config)
        self.numel, self.attach = {}, {}
        self.resident = OrderedDict()        # arena -> slot, least recently used first
        self.holders = {}                    # arena -> number of acquire() calls not yet released
        self.reading = {}                    # arena -> slot of a read that was issued and not waited for
        self.writes_in_flight = False
        self.being_written = set()           # arenas whose buffer an asynchronous write is still reading
        self.last_use = {}                   # slot -> CUDA event after the last device copy out of / into it
        self.stats = {"reads": 0, "writes": 0, "evictions": 0, "prefetch_hits": 0}

    # ---- geometry
    def _aligned(self, numel):
        return -(-numel // self.alignment) * self.alignment

    def _path(self, idx):
        return os.path.join(self.folder, f"arena_{idx}.swp")

    def _io_view(self, idx, slot):
        return self.pool.narrow(0, slot * self.slot_numel, self._aligned(self.numel[idx]))

    def _view(self, idx, slot):
        return self.pool.narrow(0, slot * self.slot_numel, self.numel[idx])

    def largest_arena(self):
        return self.slot_numel

    # ---- registration
    def register(self, idx, numel, attach):
        """Declare arena `idx` (`numel` elements). `attach(view | None)` is called with the resident buffer view every
        time the arena comes in and with None when it is evicted."""
        if self._aligned(numel) > self.slot_numel:
            raise ValueError(f"offload_param.buffer_size ({self.slot_numel} elements) is smaller than a sub-group's "
                             f"shard arena ({numel}): raise buffer_size or lower sub_group_size")
        self.numel[idx], self.attach[idx], self.holders[idx] = int(numel), attach, 0

    # ---- residency
    def _drain_writes(self):
        if self.writes_in_flight:
            self.write_handle.wait()
            self.writes_in_flight = False
            self.being_written.clear()

    def _wait_device(self, slot):
        event = self.last_use.pop(slot, None)
        if event is not None:
            event.synchronize()

    def _take_slot(self):
        if self.free_slots:
            return self.free_slots.pop()
        for idx, slot in self.resident.items():          # least recently used first
            if self.holders[idx] == 0:
                self._drain_writes()                     # its last write may still be reading the buffer
                self._wait_device(slot)
                del self.resident[idx]
                self.attach[idx](None)
                self.stats["evictions"] += 1
                return slot
        for idx in list(self.reading):                   # an unused prefetch is the next candidate
            self._finish_reads()
            return self._take_slot()
        raise RuntimeError("every parameter swap buffer is in use: raise offload_param.buffer_count "
                           f"(now {self.slot_count})")

    def _finish_reads(self):
        if self.reading:
            self.read_handle.wait()
            for idx, slot in self.reading.items():
                self.resident[idx] = slot
                self.attach[idx](self._view(idx, slot))
            self.reading = {}

    def prefetch(self, idx):
        """Start reading arena `idx` if it is not here and a buffer can be had without waiting for anybody.
        True when a read was issued by this call (the caller then has something to overlap it with)."""
        if idx in self.resident or idx in self.reading or not os.path.exists(self._path(idx)):
            return False
        if not self.free_slots and not any(self.holders[i] == 0 for i in self.resident):
            return False
        slot = self._take_slot()
        if idx in self.being_written:
            self._drain_writes()                         # do not read a file that is still being written
        self.read_handle.async_pread(self._io_view(idx, slot), self._path(idx))
        self.reading[idx] = slot
        self.stats["reads"] += 1
        return True

    def acquire(self, idx, load=True):
        """Resident view of arena `idx`; held until release(). `load=False`: the caller overwrites all of it."""
        if idx in self.reading:
            self._finish_reads()
            self.stats["prefetch_hits"] += 1
        if idx in self.being_written:
            self._drain_writes()                         # the caller may modify the buffer
        if idx in self.resident:
            self.resident.move_to_end(idx)
        else:
            slot = self._take_slot()
            if load:
                self._drain_writes()                     # the file may still be on its way out
                self.read_handle.sync_pread(self._io_view(idx, slot), self._path(idx))
                self.stats["reads"] += 1
            self.resident[idx] = slot
            self.attach[idx](self._view(idx, slot))
        self.holders[idx] += 1
        return self._view(idx, self.resident[idx])

    def release(self, idx):
        assert self.holders[idx] > 0, f"arena {idx} released more often than acquired"
        self.holders[idx] -= 1
        if torch.cuda.is_available():
            event = torch.cuda.Event()
            event.record()                               # device copies queued against this buffer so far
            self.last_use[self.resident[idx]] = event

    def write_back(self, idx, async_op=False):
        """Store the resident arena `idx` in its file."""
        slot = self.resident[idx]
        if torch.cuda.is_available():
            torch.cuda.current_stream().synchronize()    # a device -> host copy into the buffer may be in flight
        self._drain_writes()
        self.write_handle.async_pwrite(self._io_view(idx, slot), self._path(idx))
        self.writes_in_flight = True
        self.being_written.add(idx)
        self.stats["writes"] += 1
        if not async_op:
            self._drain_writes()

    def flush(self):
        self._drain_writes()
        self._finish_reads()

    def resident_arenas(self):
        return list(self.resident)
