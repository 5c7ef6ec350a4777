"""Arena allocator with compaction: hands out views of one preallocated buffer and, when the free space is large
enough but fragmented, slides the live blocks down to the start of the arena and re-points their owners.

Same public surface as the reference allocator (deepspeed/runtime/zero/contiguous_memory_allocator.py:9-283:
allocate_tensor / assign_to_param / release_tensor / release_tensor_with_id / print_allocation / max_allocated).
Bookkeeping here is a single address-ordered list of live blocks; free extents are the gaps between them, so there is
no separate free map to keep consistent and compaction is one ordered pass."""
import bisect
from dataclasses import dataclass, field
from typing import List

import torch


@dataclass(eq=False)
class _Block:
    start: int
    size: int
    view: torch.Tensor                      # the tensor handed to the caller (its .data is re-pointed on compaction)
    params: List[torch.nn.Parameter] = field(default_factory=list)


class ContiguousMemoryAllocator:
    def __init__(self, size, dtype, device):
        self.buffer = torch.zeros(size, dtype=dtype, device=device)
        self.total_size = size
        self.total_free = size
        self._blocks: List[_Block] = []     # sorted by start
        self._by_id = {}                    # id(view) -> block
        self._peak = 0
        self.count = 0                      # allocations served so far

    # ------------------------------------------------------------------ queries
    def _gaps(self):
        """Free extents (start, size) in address order."""
        cursor = 0
        for b in self._blocks:
            if b.start > cursor:
                yield cursor, b.start - cursor
            cursor = b.start + b.size
        if cursor < self.total_size:
            yield cursor, self.total_size - cursor

    @property
    def largest_contiguous(self):
        return max((size for _, size in self._gaps()), default=0)

    def max_allocated(self):
        return self._peak

    def print_allocation(self, resolution=200):
        cells = ["|"] * resolution
        for start, size in self._gaps():
            lo = int(start * resolution / self.total_size)
            hi = int((start + size) * resolution / self.total_size)
            for i in range(lo, min(hi, resolution)):
                cells[i] = "."
        print("".join(cells))

    # ------------------------------------------------------------------ allocate / release
    def allocate_tensor(self, size):
        assert size <= self.total_free, "Not enough memory in buffer. Allocation failed"
        if self.largest_contiguous < size:
            self._compact()
        # best fit: the smallest gap that holds the request
        start = min((g for g in self._gaps() if g[1] >= size), key=lambda g: g[1])[0]
        view = self.buffer.narrow(0, start, size)
        block = _Block(start, size, view)
        bisect.insort(self._blocks, block, key=lambda b: b.start)
        self._by_id[id(view)] = block
        self.total_free -= size
        self._peak = max(self._peak, self.total_size - self.total_free)
        self.count += 1
        return view

    def assign_to_param(self, tensor, param, numel, shape):
        block = self._by_id.get(id(tensor))
        assert block is not None, "No such tensor allocated by the allocator."
        assert tensor.numel() >= numel, "the allocation is smaller than the parameter"
        assert not block.params, "This tensor has already been assigned to a param"
        block.params.append(param)
        param.data = tensor.narrow(0, 0, numel).view(shape).data

    def release_tensor(self, tensor):
        self.release_tensor_with_id(id(tensor))

    def release_tensor_with_id(self, tensor_id):
        block = self._by_id.pop(tensor_id, None)
        assert block is not None, "Invalid tensor id"
        self._blocks.remove(block)
        self.total_free += block.size

    # ------------------------------------------------------------------ compaction
    def _compact(self):
        """Slide every live block down so that all free space ends up in one extent at the top."""
        cursor = 0
        for block in self._blocks:          # address order: a block only ever moves towards lower addresses
            if block.start != cursor:
                src = self.buffer.narrow(0, block.start, block.size)
                dst = self.buffer.narrow(0, cursor, block.size)
                overlap = cursor + block.size > block.start
                dst.copy_(src.clone() if overlap else src)
                block.start = cursor
                block.view.data = dst.data
                for p in block.params:
                    p.data = dst.narrow(0, 0, p.numel()).view(p.data.shape).data
            cursor += block.size
