"""Data-parallel gradient reduction for engines WITHOUT a flat gradient arena (client optimizers, no fp16 / ZeRO
wrapper): dense gradients go through flattened buckets, embedding-style sparse gradients as CSR (indices, values)
all-gathers. Engines with an arena never come here: their reduction is the in-place bucket collective of
`runtime/flat.py`.

Behaviour per reference deepspeed/runtime/engine.py:1230-1460 (pre / post scaling with `gradient_predivide_factor`,
fp32 all-reduce option, zero-filled missing gradients so that every rank reduces the same tensor list, CSR gradients
only for the module names registered as sparse).
"""
import torch
import torch.distributed as dist

from .csr_tensor import CSRTensor


def flatten_dense(tensors):
    return torch.cat([t.contiguous().view(-1) for t in tensors])


def unflatten_like(flat, tensors):
    out, offset = [], 0
    for t in tensors:
        out.append(flat.narrow(0, offset, t.numel()).view_as(t))
        offset += t.numel()
    return out


def split_half_float_double_csr(tensors):
    """[(type string, tensors of that type)] in order of first appearance."""
    groups = {}
    for t in tensors:
        groups.setdefault(t.type(), []).append(t)
    return list(groups.items())


class GradientReducer:
    """Bound to one engine; reads its data-parallel group and scaling options at call time."""

    def __init__(self, engine):
        self.e = engine

    # ---- dense
    def _scales(self):
        """(factor applied before the all-reduce, factor applied after it)."""
        e = self.e
        if not e.postscale_gradients():
            return 1.0 / e.dp_world_size, 1.0
        pre = e.gradient_predivide_factor()
        post = pre / e.dp_world_size if (e.gradient_average and pre != e.dp_world_size) else 1.0
        return 1.0 / pre, post

    def reduce_flat(self, bucket):
        """All-reduce the tensors of `bucket` as one flat buffer; returns the reduced flat tensor."""
        e = self.e
        flat = flatten_dense(bucket)
        wire = flat.float() if e.allreduce_always_fp32() else flat
        pre, post = self._scales()
        if pre != 1.0:
            wire.mul_(pre)
        dist.all_reduce(wire, group=e.data_parallel_group)
        if post != 1.0:
            wire.mul_(post)
        if wire is not flat:
            flat.copy_(wire)
        return flat

    def reduce_in_place(self, bucket):
        for grad, reduced in zip(bucket, unflatten_like(self.reduce_flat(bucket), bucket)):
            grad.copy_(reduced)

    def reduce_dense(self, tensors, numel_per_bucket):
        """Greedy buckets: a bucket is flushed as soon as it exceeds `numel_per_bucket` elements."""
        pending, count = [], 0
        for t in tensors:
            pending.append(t)
            count += t.numel()
            if count > numel_per_bucket:
                self.reduce_in_place(pending)
                pending, count = [], 0
        if pending:
            self.reduce_in_place(pending)

    # ---- sparse (CSR)
    def gather_rows(self, value):
        """Every rank's `value` (rows may differ per rank): pad to the longest, all-gather, trim."""
        e = self.e
        rows = torch.tensor([value.size(0)], dtype=torch.long, device=e.device)
        all_rows = [torch.zeros_like(rows) for _ in range(e.dp_world_size)]
        dist.all_gather(all_rows, rows, group=e.data_parallel_group)
        longest = int(torch.cat(all_rows).max())
        assert value.dim() in (1, 2)
        if longest > value.size(0):
            pad_shape = (longest - value.size(0),) + tuple(value.shape[1:])
            value = torch.cat([value, value.new_zeros(pad_shape)])
        slots = [torch.zeros_like(value) for _ in range(e.dp_world_size)]
        dist.all_gather(slots, value, group=e.data_parallel_group)
        return [slot[:int(n)] for slot, n in zip(slots, all_rows)]

    def reduce_csr(self, csr):
        """Average = concatenate every rank's (indices, values / world): duplicates add up when densified."""
        csr.values.div_(self.e.dp_world_size)
        csr.indices = torch.cat(self.gather_rows(csr.indices))
        csr.values = torch.cat(self.gather_rows(csr.values))
        return csr

    def reduce_sparse(self, bucket):
        for csr in bucket:
            dense = self.reduce_csr(csr).to_dense()
            if csr.orig_param is not None:
                csr.orig_param.grad = dense
            else:
                csr.orig_dense_tensor.copy_(dense)

    # ---- everything
    def collect(self):
        """The gradient of every parameter in module order: dense tensors, or CSR wrappers for registered sparse
        modules. Missing gradients become zeros (all ranks must reduce identical lists)."""
        e = self.e
        out = []
        for name, p in e.module.named_parameters():
            if p.grad is None:
                p.grad = torch.zeros(p.size(), dtype=p.dtype, device=p.device)
                out.append(p.grad.data)
                continue
            g = p.grad.data
            if e.sparse_gradients_enabled() and name in e.csr_tensor_module_names:
                csr = CSRTensor(g)
                if g.is_sparse:
                    csr.orig_param = p      # the reduced gradient replaces the sparse one as a dense tensor
                out.append(csr)
            elif g.is_sparse:
                p.grad = g.to_dense()
                out.append(p.grad.data)
            else:
                out.append(g)
        return out

    def reduce_all(self, elements_per_buffer):
        for kind, bucket in split_half_float_double_csr(self.collect()):
            if kind == CSRTensor.type():
                self.reduce_sparse(bucket)
            else:
                self.reduce_dense(bucket, elements_per_buffer)
