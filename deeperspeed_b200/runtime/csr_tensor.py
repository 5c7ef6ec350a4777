"""Row-sparse gradient container for embedding tables (`sparse_gradients: true`): the rows of a [rows, cols] gradient
that are non-zero plus their indices. The engine exchanges (indices, values) with padded all-gathers instead of
all-reducing the dense table (engine_train.py: csr_allreduce). API of reference deepspeed/runtime/csr_tensor.py:11-59
(`indices`, `values`, `dense_size`, `to_dense`, `add`, `sparse_size`, `type`)."""
import torch


class CSRTensor:
    def __init__(self, dense_tensor=None):
        self.orig_dense_tensor = dense_tensor
        self.indices = self.values = self.dense_size = None
        self.orig_param = None     # set by the engine when the gradient arrived as a torch sparse tensor
        if dense_tensor is not None and dense_tensor.is_sparse:
            # nn.Embedding(sparse=True) hands over a COO gradient whose first index dimension is the row
            coo = dense_tensor.coalesce()
            self.indices = coo.indices()[0].clone()
            self.values = coo.values().clone()
            self.dense_size = list(dense_tensor.shape)
            self.orig_dense_tensor = None
        elif dense_tensor is not None:
            touched = dense_tensor.ne(0).any(dim=1)          # rows with at least one non-zero entry
            self.indices = touched.nonzero(as_tuple=False).reshape(-1)
            self.values = dense_tensor.index_select(0, self.indices)
            self.dense_size = list(dense_tensor.shape)

    @staticmethod
    def type():
        return "deepspeed.CSRTensor"

    def to_dense(self):
        """Scatter the rows back; duplicate indices (after `add`) accumulate."""
        dense = self.values.new_zeros(self.dense_size)
        return dense.index_add_(0, self.indices, self.values)

    def add(self, b):
        """Concatenate another sparse gradient of the same dense shape (summation happens in `to_dense`)."""
        other = b
        assert self.dense_size == other.dense_size, "both operands must describe the same dense tensor"
        self.indices = torch.cat((self.indices, other.indices))
        self.values = torch.cat((self.values, other.values))

    def sparse_size(self):
        """(elements held: indices + values, elements of the dense tensor)."""
        held = self.indices.numel() + self.values.numel()
        return held, self.dense_size[0] * self.dense_size[1]

    def __repr__(self):
        held, dense = self.sparse_size()
        return (f"DeepSpeed.CSRTensor(indices_size={tuple(self.indices.shape)}, values_size={tuple(self.values.shape)}, "
                f"dense_size={self.dense_size}, device={self.indices.device}, reduction_factor={dense / max(held, 1):.2f})")

    __str__ = __repr__
