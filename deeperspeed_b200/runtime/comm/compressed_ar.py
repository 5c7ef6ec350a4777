"""Experimental "24-bit" all-reduce (DeeperSpeed addition, reference deepspeed/runtime/comm/compressed_ar.py:22-54):
split each fp32 value into an fp16 mantissa and an int8 exponent (frexp), all-reduce both, recombine with ldexp.
Not a true sum (exponents add) — kept for API parity, not wired into the engine."""
import torch
import torch.distributed as dist


def tensor_frexp(t):
    mant, exp = torch.frexp(t.float())
    return mant, exp


def tensor_ldexp(mant, exp):
    return torch.ldexp(mant, exp)


def compressed_all_reduce(tensor, group=None, op=dist.ReduceOp.SUM, async_op=False):
    original_dtype = tensor.dtype
    mant, exp = tensor_frexp(tensor)
    mant = mant.to(torch.float16)
    exp = exp.to(torch.int8)
    h1 = dist.all_reduce(mant, op=op, group=group, async_op=async_op)
    h2 = dist.all_reduce(exp, op=op, group=group, async_op=async_op)
    if async_op:
        h1.wait()
        h2.wait()
    out = tensor_ldexp(mant.float(), exp.to(torch.int32)).to(original_dtype)
    tensor.data.copy_(out)
    return tensor
