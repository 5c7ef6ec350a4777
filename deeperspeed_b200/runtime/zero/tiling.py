"""TiledLinear: split a huge linear layer into in_splits x out_splits tiles so ZeRO-3 gathers one tile at a time
(reference deepspeed/runtime/zero/tiling.py:26-294)."""
import torch

from .partition_parameters import GatheredParameters, is_zero_param
from ..utils import partition_uniform as partition


def split_tensor_along_last_dim(tensor, partitions, contiguous_split_chunks=False):
    """Split the last dim of `tensor` at the boundaries in `partitions` (a prefix list of length parts + 1)."""
    last_dim = tensor.dim() - 1
    sizes = [partitions[i + 1] - partitions[i] for i in range(len(partitions) - 1)]
    tensor_list = torch.split(tensor, sizes, dim=last_dim)
    if contiguous_split_chunks:
        return tuple(chunk.contiguous() for chunk in tensor_list)
    return tensor_list


class TiledLinear(torch.nn.Module):
    def __init__(self, in_features, out_features, bias=True, in_splits=1, out_splits=1, input_is_already_split=False,
                 combine_out_splits=True, linear_cls=torch.nn.Linear, init_linear=None, **kwargs):
        super().__init__()
        if (in_splits < 1) or (in_splits > in_features):
            raise RuntimeError("in splits must be in range [1, in_features].")
        if (out_splits < 1) or (out_splits > out_features):
            raise RuntimeError("out splits must be in range [1, out_features].")
        self.in_features = in_features
        self.out_features = out_features
        self.use_bias = bias
        self.out_splits = out_splits
        self.in_splits = in_splits
        self.input_is_already_split = input_is_already_split
        self.combine_out_splits = combine_out_splits
        self.in_parts = partition(num_items=in_features, num_parts=in_splits)
        self.out_parts = partition(num_items=out_features, num_parts=out_splits)
        assert len(self.out_parts) == out_splits + 1 and len(self.in_parts) == in_splits + 1
        assert self.out_parts[0] == 0 and self.out_parts[out_splits] == out_features
        assert self.in_parts[in_splits] == in_features
        self.linears = torch.nn.ModuleList()
        for out_id in range(out_splits):
            self.linears.append(torch.nn.ModuleList())
            local_out_dim = self.out_parts[out_id + 1] - self.out_parts[out_id]
            for in_id in range(in_splits):
                # only the last tile of a row carries the bias (it is added once per output slice)
                local_bias = bias if in_id == (in_splits - 1) else False
                local_in_dim = self.in_parts[in_id + 1] - self.in_parts[in_id]
                local = linear_cls(local_in_dim, local_out_dim, bias=local_bias, **kwargs)
                self.linears[out_id].append(local)
        if init_linear is not None:
            self.copy_params_from(init_linear)

    def forward(self, input_):
        if self.in_splits > 1 and not self.input_is_already_split:
            inputs = split_tensor_along_last_dim(input_, self.in_parts)
        elif self.in_splits > 1:
            inputs = input_
            assert len(inputs) == self.in_splits, f"Col splits {self.in_splits} does not match input splits {len(inputs)}"
        else:
            inputs = [input_]
        outputs = [None] * self.out_splits
        for out_id in range(self.out_splits):
            for in_id in range(self.in_splits):
                local_output = self.linears[out_id][in_id](inputs[in_id])
                outputs[out_id] = self._reduce_local_output(in_id=in_id, out_id=out_id, current_out=outputs[out_id],
                                                            new_out=local_output)
        if self.combine_out_splits:
            return self._combine_output_splits(outputs)
        return outputs

    def _split_global_input(self, input, splits):
        return split_tensor_along_last_dim(input, splits)

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        if current_out is None:
            return new_out.clone()
        return current_out + new_out

    def _combine_output_splits(self, outputs):
        assert len(outputs) == self.out_splits
        return torch.cat(outputs, dim=-1)

    @torch.no_grad()
    def copy_params_from(self, other):
        """Copy weight and bias from a full (un-tiled) linear layer, gathering ZeRO-3 params as needed."""
        assert hasattr(other, "weight")
        assert other.weight.size() == (self.out_features, self.in_features)
        if self.use_bias:
            assert hasattr(other, "bias") and other.bias is not None
            assert other.bias.size() == (self.out_features,)
        else:
            assert other.bias is None
        for row in range(self.out_splits):
            rstart, rstop = self.out_parts[row], self.out_parts[row + 1]
            for col in range(self.in_splits):
                cstart, cstop = self.in_parts[col], self.in_parts[col + 1]
                local = self.linears[row][col]
                global_weight = other.weight[rstart:rstop, cstart:cstop]
                with GatheredParameters(local.weight, modifier_rank=0):
                    local.weight.copy_(global_weight)
            if local.bias is not None:
                with GatheredParameters(local.bias, modifier_rank=0):
                    local.bias.data.copy_(other.bias[rstart:rstop].data)


class TiledLinearReturnBias(TiledLinear):
    """TiledLinear for layers whose forward returns (output, bias) (Megatron-style) ."""

    def _reduce_local_output(self, in_id, out_id, current_out, new_out):
        if current_out is None:
            return new_out
        old_tensor, old_bias = current_out
        new_tensor, new_bias = new_out
        tensor = old_tensor + new_tensor
        if old_bias is None:
            bias = new_bias
        elif new_bias is None:
            bias = old_bias
        else:
            bias = old_bias + new_bias
        return tensor, bias

    def _combine_output_splits(self, outputs):
        tensors = [o[0] for o in outputs]
        tensor = super()._combine_output_splits(tensors)
        biases = [o[1] for o in outputs if o[1] is not None]
        bias = super()._combine_output_splits(biases) if len(biases) > 0 else None
        return tensor, bias
