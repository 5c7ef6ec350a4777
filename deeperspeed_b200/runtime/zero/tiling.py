"""TiledLinear: a linear layer stored as an out_splits x in_splits grid of small linears, so that ZeRO-3 gathers (and
frees) one tile at a time instead of the whole weight.

Interface per reference deepspeed/runtime/zero/tiling.py:26-294 (constructor keywords, `linears[row][col]`,
`copy_params_from`, the `_reduce_local_output` / `_combine_output_splits` hooks subclasses override). The forward pass
is organised per output slice: the partial products of one row of tiles are reduced as they are produced, so only one
output slice and one gathered tile are live at any time.
"""
import torch

from .partition_parameters import GatheredParameters
from ..utils import partition_uniform as partition


def split_tensor_along_last_dim(tensor, partitions, contiguous_split_chunks=False):
    """Pieces of the last dimension between consecutive boundaries of `partitions` (length = pieces + 1)."""
    widths = [hi - lo for lo, hi in zip(partitions[:-1], partitions[1:])]
    pieces = torch.split(tensor, widths, dim=-1)
    return tuple(p.contiguous() for p in pieces) if contiguous_split_chunks else pieces


class TiledLinear(torch.nn.Module):
    """y = x W^T + b with W cut into tiles. `in_splits` cuts the input features (partial products are summed),
    `out_splits` cuts the output features (slices are concatenated unless `combine_out_splits=False`)."""

    def __init__(self, in_features, out_features, bias=True, in_splits=1, out_splits=1, input_is_already_split=False,
                 combine_out_splits=True, linear_cls=torch.nn.Linear, init_linear=None, **kwargs):
        super().__init__()
        if not 1 <= in_splits <= in_features:
            raise RuntimeError("in splits must be in range [1, in_features].")
        if not 1 <= out_splits <= out_features:
            raise RuntimeError("out splits must be in range [1, out_features].")
        self.in_features, self.out_features = in_features, out_features
        self.use_bias = bias
        self.in_splits, self.out_splits = in_splits, out_splits
        self.input_is_already_split = input_is_already_split
        self.combine_out_splits = combine_out_splits
        self.in_parts = partition(num_items=in_features, num_parts=in_splits)
        self.out_parts = partition(num_items=out_features, num_parts=out_splits)
        assert self.in_parts[0] == 0 and self.in_parts[-1] == in_features and len(self.in_parts) == in_splits + 1
        assert self.out_parts[0] == 0 and self.out_parts[-1] == out_features and len(self.out_parts) == out_splits + 1

        def tile(row, col):
            # the bias of an output slice lives in the LAST tile of its row, so it is added exactly once
            return linear_cls(self.in_parts[col + 1] - self.in_parts[col], self.out_parts[row + 1] - self.out_parts[row],
                              bias=bias and col == in_splits - 1, **kwargs)

        self.linears = torch.nn.ModuleList(torch.nn.ModuleList(tile(r, c) for c in range(in_splits))
                                           for r in range(out_splits))
        if init_linear is not None:
            self.copy_params_from(init_linear)

    def _input_pieces(self, input_):
        if self.in_splits == 1:
            return [input_]
        if self.input_is_already_split:
            assert len(input_) == self.in_splits, \
                f"Col splits {self.in_splits} does not match input splits {len(input_)}"
            return input_
        return self._split_global_input(input_, self.in_parts)

    def forward(self, input_):
        pieces = self._input_pieces(input_)
        slices = []
        for row, tiles in enumerate(self.linears):
            acc = None
            for col, (lin, x) in enumerate(zip(tiles, pieces)):
                acc = self._reduce_local_output(in_id=col, out_id=row, current_out=acc, new_out=lin(x))
            slices.append(acc)
        return self._combine_output_splits(slices) if self.combine_out_splits else slices

    # ---- hooks
    def _split_global_input(self, input, splits):
        return split_tensor_along_last_dim(input, splits)

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        """Fold the product of tile (out_id, in_id) into the running sum of its output slice."""
        return new_out.clone() if current_out is None else current_out + new_out

    def _combine_output_splits(self, outputs):
        assert len(outputs) == self.out_splits
        return torch.cat(outputs, dim=-1)

    @torch.no_grad()
    def copy_params_from(self, other):
        """Fill the tiles from a full linear layer; ZeRO-3 tiles are gathered one at a time while they are written."""
        assert hasattr(other, "weight") and other.weight.size() == (self.out_features, self.in_features)
        if self.use_bias:
            assert getattr(other, "bias", None) is not None and other.bias.size() == (self.out_features,)
        else:
            assert other.bias is None
        for row, tiles in enumerate(self.linears):
            r0, r1 = self.out_parts[row], self.out_parts[row + 1]
            for col, lin in enumerate(tiles):
                with GatheredParameters(lin.weight, modifier_rank=0):
                    lin.weight.copy_(other.weight[r0:r1, self.in_parts[col]:self.in_parts[col + 1]])
                if lin.bias is not None:
                    with GatheredParameters(lin.bias, modifier_rank=0):
                        lin.bias.data.copy_(other.bias[r0:r1].data)


class TiledLinearReturnBias(TiledLinear):
    """For tiles whose forward returns (output, bias) instead of adding the bias (Megatron-style linears)."""

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        if current_out is None:
            return new_out
        (t0, b0), (t1, b1) = current_out, new_out
        both = [b for b in (b0, b1) if b is not None]
        return t0 + t1, (sum(both[1:], both[0]) if both else None)

    def _combine_output_splits(self, outputs):
        tensor = super()._combine_output_splits([o[0] for o in outputs])
        biases = [o[1] for o in outputs if o[1] is not None]
        return tensor, (torch.cat(biases, dim=-1) if biases else None)
