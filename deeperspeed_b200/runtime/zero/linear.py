"""Memory-efficient linear for ZeRO-3: the autograd function keeps a *reference to the Parameter object*, not to
its (gathered) storage, so the weight can be released between forward and backward and re-gathered by the ZeRO-3
hooks (reference deepspeed/runtime/zero/linear.py:29-168). On CUDA the products run on the tcgen05 GEMM kernels."""
import torch
from torch import Tensor

from ...ops import kernels as K

tensor_map = {}     # kept for code that imported it; this implementation does not need a registry


def _tensor_core_ok(x2, out_features):
    return x2.is_cuda and x2.dtype in (torch.float16, torch.bfloat16) and x2.shape[-1] % 8 == 0 and out_features % 8 == 0


class LinearFunctionForZeroStage3(torch.autograd.Function):
    """y = x W^T + b. Only the INPUT goes through save_for_backward; weight and bias are remembered as the Parameter
    objects themselves (plain attributes of ctx), so autograd pins no gathered storage: between forward and backward
    ZeRO-3 may shrink `weight.data` to its placeholder, and by the time backward reads `ctx.weight.data` the
    pre-backward hook has gathered it again."""

    @staticmethod
    def forward(ctx, input, weight, bias=None):
        ctx.save_for_backward(input)
        ctx.weight, ctx.bias = weight, bias
        x2 = input.reshape(-1, input.shape[-1])
        if _tensor_core_ok(x2, weight.shape[0]):
            y = K.gemm(x2, weight, trans_b=True, epilogue=1 if bias is not None else 0, bias=bias)
        elif bias is not None:
            y = torch.addmm(bias, x2, weight.t())
        else:
            y = x2.matmul(weight.t())
        return y.view(*input.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, grad_output):
        (input,) = ctx.saved_tensors
        weight, bias = ctx.weight, ctx.bias
        want_x, want_w, want_b = ctx.needs_input_grad
        g2 = grad_output.reshape(-1, grad_output.shape[-1])
        x2 = input.reshape(-1, input.shape[-1])
        fast = _tensor_core_ok(g2, weight.shape[1]) and g2.shape[0] % 8 == 0 and g2.is_contiguous()
        grad_input = grad_weight = grad_bias = None
        if want_x:
            gx = K.gemm(g2, weight.data, trans_b=False) if fast else g2.matmul(weight)
            grad_input = gx.view(input.shape)
        if want_w:
            grad_weight = K.gemm(g2, x2.contiguous(), trans_a=True, trans_b=False) if fast else g2.t().matmul(x2)
        if bias is not None and want_b:
            grad_bias = g2.sum(0)
        return grad_input, grad_weight, grad_bias

    @staticmethod
    def apply_compat(input, weight, bias=None):
        return LinearFunctionForZeroStage3.apply(input, weight, bias)


class WeightHolder:
    """Keeps a partitioned weight Parameter out of autograd's sight (it is a 1-element placeholder in forward, so
    autograd would reduce its gradient to that shape) plus a dummy leaf that forces the backward to run."""

    def __init__(self, param):
        self.param = param
        self.dummy = torch.zeros((), device=param.device, requires_grad=True)


class FusedAGLinearFunction(torch.autograd.Function):
    """ZeRO-3 linear whose forward gathers the weight INSIDE the GEMM kernel from the peers' shards
    (parallel/ag_gemm.py). The weight is not an autograd input and stays partitioned in backward too: the input
    gradient comes from the K-sharded all-gather + GEMM kernel (`ag_linear_dgrad`), the weight gradient from the
    GEMM + reduce-scatter kernel (`GemmReduceScatter`), whose result goes straight into the gradient partition.
    Where a fused kernel does not apply (shape, or the weight happens to be gathered) the plain GEMMs and the
    parameter's ZeRO gradient hook (NCCL reduce-scatter + release) are used."""

    @staticmethod
    def forward(ctx, input, bias, dummy, holder):
        from ...parallel.ag_gemm import ag_linear
        ctx.holder = holder
        ctx.save_for_backward(input, bias)
        return ag_linear(input, holder.param.ds_sharded, bias)

    @staticmethod
    def backward(ctx, grad_output):
        input, bias = ctx.saved_tensors
        w = ctx.holder.param
        gathered = w.numel() == w.ds_numel
        go2 = grad_output.reshape(-1, grad_output.shape[-1])
        if not go2.is_contiguous():
            go2 = go2.contiguous()
        x2 = input.reshape(-1, input.shape[-1])
        fast = go2.shape[0] % 64 == 0 and go2.shape[1] % 8 == 0 and x2.shape[1] % 8 == 0
        grad_input = grad_bias = None
        if ctx.needs_input_grad[0]:
            if not gathered:
                from ...parallel.ag_gemm import ag_linear_dgrad
                gi = ag_linear_dgrad(go2, w.ds_sharded)
            else:
                gi = K.gemm(go2, w.data, trans_b=False) if fast else go2.matmul(w.data)
            grad_input = gi.view(input.shape)
        if bias is not None and ctx.needs_input_grad[1]:
            grad_bias = go2.float().sum(0).to(go2.dtype)
        if w.requires_grad:
            rs = getattr(w, "ds_gemm_rs", None)
            x2c = x2.contiguous()
            if (rs is not None and fast and w.grad is None and go2.dtype == rs.dtype
                    and rs.supported(go2.shape[1], x2c.shape[1], go2.shape[0])):
                # weight-gradient GEMM with the reduce-scatter inside the kernel: only this rank's averaged rows
                # come back, straight into the gradient partition
                w.ds_reduced_grad_hook(rs(go2, x2c, trans_a=True, trans_b=False, average=True))
                return grad_input, grad_bias, None, None
            gw = K.gemm(go2, x2c, trans_a=True, trans_b=False) if fast else go2.t().matmul(x2)
            w.grad = gw if w.grad is None else w.grad + gw
            hook = getattr(w, "ds_grad_hook", None)
            if hook is not None:
                hook(w)
        return grad_input, grad_bias, None, None


class LinearModuleForZeroStage3(torch.nn.Linear):
    """nn.Linear (same parameters, same initialisation) whose forward is LinearFunctionForZeroStage3."""

    def forward(self, input: Tensor) -> Tensor:
        return LinearFunctionForZeroStage3.apply(input, self.weight, self.bias)
