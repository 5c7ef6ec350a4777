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


# ---- the reference module's own entry points (its cupy variant needs no cupy here: torch.frexp / ldexp run on the GPU)
def decompose(t):
    """fp32 value -> (fp16 mantissa in [0.5, 1), int8 exponent)."""
    mant, exp = tensor_frexp(t)
    return mant.half(), exp.to(torch.int8)


def reconstruct(mantissa, exponent, original_dtype=torch.bfloat16):
    return tensor_ldexp(mantissa.float(), exponent.to(torch.int32)).to(original_dtype)


def compressed_all_reduce_torch(tensor, op=dist.ReduceOp.SUM, group=None, async_op=False):
    """Returns a new tensor (the reference's behaviour); `compressed_all_reduce` above also writes it back in place."""
    mant, exp = decompose(tensor)
    handles = [dist.all_reduce(mant, op=op, group=group, async_op=async_op),
               dist.all_reduce(exp, op=op, group=group, async_op=async_op)]
    if async_op:
        for h in handles:
            h.wait()
    return reconstruct(mant, exp, tensor.dtype)


compressed_all_reduce_cupy = compressed_all_reduce_torch


def decompose_cupy(tensor):
    return decompose(tensor)


def torch2cupy(tensor):
    """No cupy in this framework: tensors stay torch tensors (see runtime/compression/cupy.py)."""
    return tensor


def cupy2torch(cupy_tensor):
    return cupy_tensor

