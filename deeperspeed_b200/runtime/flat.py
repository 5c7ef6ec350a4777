"""Flat parameter / gradient arenas.

The reference flattens parameters at optimizer construction and flattens/unflattens gradients with copies on
every step (reference runtime/engine.py:1367-1395, zero/stage1.py:466-555, csrc/utils/flatten_unflatten.cpp).
Here every parameter of a group is re-homed ONCE into a contiguous arena and its `.grad` is a permanent view of a
matching gradient arena, so autograd accumulates straight into communication-ready memory: collectives, the fused
norm/overflow kernel and the fused optimizer all operate on slices of the arena and no flatten pass exists.
Parameter starts are aligned to 8 elements (16 bytes) so each parameter remains a legal TMA / 128-bit operand.
"""
import torch


def _align_up(x, a):
    return (x + a - 1) // a * a


class Segment:
    __slots__ = ("param", "offset", "numel", "shape")

    def __init__(self, param, offset):
        self.param = param
        self.offset = offset
        self.numel = param.numel()
        self.shape = tuple(param.shape)


class FlatArena:
    def __init__(self, params, shard_count=1, align=8, allocator=None, shard_align=None):
        params = list(params)
        assert len(params) > 0
        self.dtype = params[0].dtype
        self.device = params[0].device
        self.align = align
        self.shard_count = shard_count
        self.segments = []
        off = 0
        for p in params:
            assert p.dtype == self.dtype, "one arena holds one dtype"
            self.segments.append(Segment(p, off))
            off = _align_up(off + p.numel(), align)
        self.numel_used = off
        self.numel_unpadded = sum(s.numel for s in self.segments)
        quantum = shard_count * (shard_align or max(align, 8))
        self.numel = _align_up(max(off, 1), quantum)
        self.shard_size = self.numel // shard_count
        self._alloc = allocator or (lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev))
        self.data = self._alloc(self.numel, self.dtype, self.device)
        for s in self.segments:
            view = self.data.narrow(0, s.offset, s.numel).view(s.shape)
            view.copy_(s.param.data)
            s.param.data = view
        self.grad = None

    # ---- gradients
    def allocate_grads(self, allocator=None):
        if self.grad is None:
            alloc = allocator or self._alloc
            self.grad = alloc(self.numel, self.dtype, self.device)
            self.grad.zero_()
        self.bind_grads()
        return self.grad

    def bind_grads(self):
        """(Re)attach each param.grad as a view into the gradient arena."""
        for s in self.segments:
            g = s.param.grad
            view = self.grad.narrow(0, s.offset, s.numel).view(s.shape)
            if g is None or g.data_ptr() != view.data_ptr():
                if g is not None:
                    view.copy_(g)
                s.param.grad = view

    def absorb_stray_grads(self):
        """If user code replaced a .grad (e.g. set_to_none), fold it back into the arena."""
        if self.grad is None:       # gradients are not kept whole (ZeRO-2 sharded gradient memory)
            return
        for s in self.segments:
            g = s.param.grad
            view = self.grad.narrow(0, s.offset, s.numel).view(s.shape)
            if g is None:
                s.param.grad = view
            elif g.data_ptr() != view.data_ptr():
                view.add_(g.to(view.dtype))
                s.param.grad = view

    def zero_grads(self):
        if self.grad is not None:
            self.grad.zero_()
            self.bind_grads()

    # ---- sharding
    def shard(self, flat, rank):
        return flat.narrow(0, rank * self.shard_size, self.shard_size)

    def rebind_params(self):
        for s in self.segments:
            s.param.data = self.data.narrow(0, s.offset, s.numel).view(s.shape)

    # ---- padding-free (reference-compatible) flat layout for checkpoints
    def to_unpadded(self, flat):
        """arena-layout flat tensor -> plain concatenation of the parameters (reference flat layout)."""
        if self.numel_used == self.numel_unpadded:
            return flat.narrow(0, 0, self.numel_unpadded).clone()
        return torch.cat([flat.narrow(0, s.offset, s.numel) for s in self.segments])

    def from_unpadded(self, src, out=None):
        if out is None:
            out = torch.zeros(self.numel, dtype=src.dtype, device=src.device)
        pos = 0
        for s in self.segments:
            out.narrow(0, s.offset, s.numel).copy_(src.narrow(0, pos, s.numel))
            pos += s.numel
        return out

    def segment_ranges_in(self, start, end):
        """Segments overlapping arena range [start, end) as (segment, lo, hi) with lo/hi relative to the segment."""
        out = []
        for s in self.segments:
            lo, hi = max(start, s.offset), min(end, s.offset + s.numel)
            if lo < hi:
                out.append((s, lo - s.offset, hi - s.offset))
        return out
