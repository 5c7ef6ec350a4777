"""Causal self-attention block op for decoder models: fused-QKV in, context out.

forward : rotary (our vectorised out-of-place kernel, reads q/k straight from the fused QKV activation) -> causal
          flash attention -> context [B, S, h]
backward: attention backward -> inverse rotary written DIRECTLY into the dQKV buffer (no slice-backward zero fills,
          no gradient-accumulation adds, no clones).

Two attention back ends, selected with DSB_ATTN (or `set_attention_backend`):
  "native": the tcgen05 flash-attention kernels of csrc/transformer/flash_attn_v2.cu (forward + backward, V read in
            place from the fused activation, dV written in place into the dQKV buffer). Default. Needs head dim 128 and
            seq % 128 == 0; other shapes run the library path.
  "sdpa"  : the library flash kernel behind torch SDPA, kept for comparison (bench/flash_attn_bench.py).
"""
import ctypes as C
import os

import torch
import torch.nn.functional as F

from . import kernels as K
from . import native as N

_p, _i, _l, _f = C.c_void_p, C.c_int, C.c_long, C.c_float
N.register_sigs({
    "dsb_rotary_table": [_p, _p, _i, _i, _f, _p],
    "dsb_rotary_oop": [_p, _p, _p, _p, _i, _i, _i, _i, _i, _l, _l, _l, _l, _l, _l, _p, _p, _i, _i, _p],
    "dsb_rotary_oop_v": [_p, _p, _p, _p, _i, _i, _i, _i, _i, _l, _l, _l, _l, _l, _l, _p, _p, _i, _i, _p, _p, _l, _l,
                         _l, _p],
})

N.register_sigs({
    "dsb_flash_attn_fwd": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _l, _l, _l, _l, _l, _l, _l, _l, _l, _f, _i,
                           _i, _p],
})

N.register_sigs({
    "dsb_flash_attn_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _f, _i, _i, _p],
    "dsb_flash_attn_bwd2": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _f, _i, _i, _p, _p, _i, _p],
})

_tables = {}


def set_flash_bwd_variant(v):
    """2: transposed backward with P^T / dS^T in tensor memory (default); 1: first-generation kernel (returns dq as its
    fp32 accumulator). Returns the previous setting."""
    global _bwd_gen
    assert v in (1, 2)
    prev, _bwd_gen = _bwd_gen, v
    return prev


def set_flash_fwd_variant(keys_per_step):
    """2: two query tiles per CTA, P kept in tensor memory (default; needs S % 256 == 0, else the 64-key kernel runs);
    64: 64-key blocks, two CTAs per SM; 128: 128-key blocks, one CTA per SM. Returns the previous one."""
    lib = N.cuda_lib()
    lib.dsb_flash_attn_fwd_variant.argtypes = [C.c_int]
    lib.dsb_flash_attn_fwd_variant.restype = C.c_int
    return lib.dsb_flash_attn_fwd_variant(int(keys_per_step))


_bwd_gen = 2


def flash_attn_bwd(dout, q, k, v, out, lse, causal=True, scale=None, dk=None, dv=None, dq=None, rotary=None):
    """Backward of `flash_attn_fwd` on the tcgen05 kernels. All tensors [B, S, H, 128] views with unit last stride
    (dq / dk / dv may be slices of a fused dQKV buffer). Returns (dq, dk, dv). With the default (second-generation)
    kernel dq comes back in the input dtype; the first-generation kernel (set_flash_bwd_variant(1)) returns its fp32
    accumulator [B, S, H, 128] and ignores `dq`. `rotary` = (rot_dim, base): q and k are rotated tensors and dq / dk
    are wanted for the unrotated ones; the second-generation kernel applies the inverse rotation itself."""
    B, S, H, D = q.shape
    if scale is None:
        scale = 1.0 / (D ** 0.5)
    if dk is None:
        dk = torch.empty(B, S, H, D, dtype=q.dtype, device=q.device)
    if dv is None:
        dv = torch.empty(B, S, H, D, dtype=q.dtype, device=q.device)
    lib = N.cuda_lib()
    if _bwd_gen == 2:
        if dq is None:
            dq = torch.empty(B, S, H, D, dtype=q.dtype, device=q.device)
        scratch = torch.empty((2 + D) * B * H * S, dtype=torch.float32, device=q.device)
        ts = (dout, q, k, v, out, dq, dk, dv)
        assert all(t.stride(-1) == 1 and t.shape == q.shape for t in ts)
        strides = (C.c_long * 24)(*[t.stride(i) for t in ts for i in range(3)])
        cs = sn = None
        rot_dim = 0
        if rotary is not None and rotary[0] > 0:
            rot_dim = int(rotary[0])
            cs, sn = rotary_table(S, rot_dim, rotary[1], q.device)
        N.check(lib.dsb_flash_attn_bwd2(N.ptr(dout), N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(out), N.ptr(lse),
                                        N.ptr(scratch), N.ptr(dq), N.ptr(dk), N.ptr(dv), B, S, H, D, strides,
                                        float(scale), int(causal), N.dtype_code(q.dtype), N.ptr(cs), N.ptr(sn),
                                        rot_dim, N.stream_ptr()), "dsb_flash_attn_bwd2")
        K._count(3)
        return dq, dk, dv
    assert rotary is None, "the first-generation backward does not apply the inverse rotary"
    dq_acc = torch.zeros(B, S, H, D, dtype=torch.float32, device=q.device)
    delta = torch.empty(B, H, S, dtype=torch.float32, device=q.device)
    ts = (dout, q, k, v, out, dk, dv)
    assert all(t.stride(-1) == 1 and t.shape == q.shape for t in ts)
    strides = (C.c_long * 21)(*[t.stride(i) for t in ts for i in range(3)])
    N.check(lib.dsb_flash_attn_bwd(N.ptr(dout), N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(out), N.ptr(lse), N.ptr(delta),
                                   N.ptr(dq_acc), N.ptr(dk), N.ptr(dv), B, S, H, D, strides, float(scale), int(causal),
                                   N.dtype_code(q.dtype), N.stream_ptr()), "dsb_flash_attn_bwd")
    K._count(2)
    return dq_acc, dk, dv


def flash_attn_fwd(q, k, v, causal=True, scale=None, out=None, return_lse=False):
    """Flash-attention forward on the tcgen05 kernel (csrc/transformer/flash_attn.cu).
    q, k, v: [B, S, H, 128] views (any batch / sequence / head strides that are multiples of 8 elements, unit last
    stride; e.g. slices of the fused QKV activation), fp16 or bf16, S a multiple of 128.
    Returns the context [B, S, H, 128] (contiguous unless `out` is given) and, if asked, the per-row log-sum-exp
    [B, H, S] (fp32, natural log of the sum of exp(scaled scores))."""
    B, S, H, D = q.shape
    assert k.shape == q.shape and v.shape == q.shape and q.stride(-1) == 1 and k.stride(-1) == 1 and v.stride(-1) == 1
    if scale is None:
        scale = 1.0 / (D ** 0.5)
    if out is None:
        out = torch.empty(B, S, H, D, dtype=q.dtype, device=q.device)
    lse = torch.empty(B, H, S, dtype=torch.float32, device=q.device) if return_lse else None
    N.check(N.cuda_lib().dsb_flash_attn_fwd(
        N.ptr(q), N.ptr(k), N.ptr(v), N.ptr(out), N.ptr(lse), B, S, H, D, q.stride(0), q.stride(1), q.stride(2),
        k.stride(0), k.stride(1), k.stride(2), v.stride(0), v.stride(1), v.stride(2), out.stride(0), out.stride(1),
        out.stride(2), float(scale), int(causal), N.dtype_code(q.dtype), N.stream_ptr()), "dsb_flash_attn_fwd")
    K._count()
    return (out, lse) if return_lse else out


def rotary_table(S, rot_dim, base, device):
    key = (S, rot_dim, float(base), device.index)
    t = _tables.get(key)
    if t is None:
        half = rot_dim // 2
        cs = torch.empty(S, half, dtype=torch.float32, device=device)
        sn = torch.empty(S, half, dtype=torch.float32, device=device)
        N.check(N.cuda_lib().dsb_rotary_table(N.ptr(cs), N.ptr(sn), S, rot_dim, float(base), N.stream_ptr()),
                "dsb_rotary_table")
        t = (cs, sn)
        _tables[key] = t
    return t


def rotary_oop(q_in, k_in, q_out, k_out, rot_dim, base, inverse=False, v_in=None, v_out=None):
    """q_in/k_in/q_out/k_out: [B, S, H, D] views with unit last stride (in and out may have different strides).
    v_in/v_out (optional): a third [B, S, H, D] tensor copied verbatim in the same launch; v_out must have q_out's
    strides (it is the V slice of the same fused buffer)."""
    B, S, H, D = q_in.shape
    assert q_in.stride() == k_in.stride() and q_out.stride() == k_out.stride()
    cs, sn = rotary_table(S, rot_dim, base, q_in.device)
    if v_in is not None:
        assert v_out.stride() == q_out.stride() and v_in.stride(-1) == 1 and v_in.shape == q_in.shape
    N.check(N.cuda_lib().dsb_rotary_oop_v(N.ptr(q_in), N.ptr(k_in), N.ptr(q_out), N.ptr(k_out), B, S, H, D, rot_dim,
                                          q_in.stride(0), q_in.stride(1), q_in.stride(2), q_out.stride(0),
                                          q_out.stride(1), q_out.stride(2), N.ptr(cs), N.ptr(sn), int(inverse),
                                          N.dtype_code(q_in.dtype), N.ptr(v_in), N.ptr(v_out),
                                          v_in.stride(0) if v_in is not None else 0,
                                          v_in.stride(1) if v_in is not None else 0,
                                          v_in.stride(2) if v_in is not None else 0, N.stream_ptr()),
            "dsb_rotary_oop_v")
    K._count()


_BACKEND = os.environ.get("DSB_ATTN", "native").lower()


def set_attention_backend(name):
    """"native" (tcgen05 flash-attention kernels) or "sdpa" (library kernel). Returns the previous setting."""
    global _BACKEND
    assert name in ("native", "sdpa")
    prev, _BACKEND = _BACKEND, name
    return prev


def attention_backend():
    return _BACKEND


class NativeCausalSelfAttentionFn(torch.autograd.Function):
    """Same contract as CausalSelfAttentionFn with the attention itself on our tcgen05 kernels."""

    @staticmethod
    def forward(ctx, qkv, H, D, rot_dim, base):
        B, S = qkv.shape[0], qkv.shape[1]
        qkv5 = qkv.view(B, S, 3, H, D)
        if rot_dim > 0:
            q = torch.empty(B, S, H, D, dtype=qkv.dtype, device=qkv.device)
            k = torch.empty_like(q)
            rotary_oop(qkv5[:, :, 0], qkv5[:, :, 1], q, k, rot_dim, base)
        else:
            q, k = qkv5[:, :, 0], qkv5[:, :, 1]
        out, lse = flash_attn_fwd(q, k, qkv5[:, :, 2], causal=True, return_lse=True)
        ctx.save_for_backward(qkv, q if rot_dim > 0 else None, k if rot_dim > 0 else None, out, lse)
        ctx.meta = (B, S, H, D, rot_dim, base)
        return out.view(B, S, H * D)

    @staticmethod
    def backward(ctx, dctx):
        qkv, q, k, out, lse = ctx.saved_tensors
        B, S, H, D, rot_dim, base = ctx.meta
        qkv5 = qkv.view(B, S, 3, H, D)
        if q is None:
            q, k = qkv5[:, :, 0], qkv5[:, :, 1]
        dout = dctx.reshape(B, S, H, D)
        if dout.stride(-1) != 1 or any(st % 8 for st in dout.stride()[:3]):
            dout = dout.contiguous()
        dqkv = torch.empty(B, S, 3, H, D, dtype=qkv.dtype, device=qkv.device)
        if rot_dim > 0 and _bwd_gen == 2:
            # inverse rotary fused into the backward kernels: dq / dk / dv land in the dQKV buffer directly
            flash_attn_bwd(dout, q, k, qkv5[:, :, 2], out, lse, causal=True, dq=dqkv[:, :, 0], dk=dqkv[:, :, 1],
                           dv=dqkv[:, :, 2], rotary=(rot_dim, base))
        elif rot_dim > 0:
            dk_tmp = torch.empty(B, S, H, D, dtype=qkv.dtype, device=qkv.device)
            dq, _, _ = flash_attn_bwd(dout, q, k, qkv5[:, :, 2], out, lse, causal=True, dk=dk_tmp, dv=dqkv[:, :, 2])
            dq16 = torch.empty_like(dk_tmp)
            K.cast(dq, dq16)
            rotary_oop(dq16, dk_tmp, dqkv[:, :, 0], dqkv[:, :, 1], rot_dim, base, inverse=True)
        else:
            dq, _, _ = flash_attn_bwd(dout, q, k, qkv5[:, :, 2], out, lse, causal=True, dk=dqkv[:, :, 1],
                                      dv=dqkv[:, :, 2], dq=dqkv[:, :, 0])
            if dq.dtype != qkv.dtype:
                dqkv[:, :, 0].copy_(dq)
        return dqkv.view(B, S, 3 * H * D), None, None, None, None


class CausalSelfAttentionFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, H, D, rot_dim, base):
        B, S = qkv.shape[0], qkv.shape[1]
        qkv5 = qkv.view(B, S, 3, H, D)
        q = torch.empty(B, S, H, D, dtype=qkv.dtype, device=qkv.device)
        k = torch.empty_like(q)
        if rot_dim > 0:
            rotary_oop(qkv5[:, :, 0], qkv5[:, :, 1], q, k, rot_dim, base)
        else:
            q.copy_(qkv5[:, :, 0])
            k.copy_(qkv5[:, :, 1])
        with torch.enable_grad():
            ql, kl = q.requires_grad_(), k.requires_grad_()
            vl = qkv5[:, :, 2].detach().requires_grad_()
            out = F.scaled_dot_product_attention(ql.transpose(1, 2), kl.transpose(1, 2), vl.transpose(1, 2),
                                                 is_causal=True)
        ctx.graph = (ql, kl, vl, out)
        ctx.meta = (B, S, H, D, rot_dim, base)
        return out.detach().transpose(1, 2).reshape(B, S, H * D)

    @staticmethod
    def backward(ctx, dctx):
        ql, kl, vl, out = ctx.graph
        B, S, H, D, rot_dim, base = ctx.meta
        dout = dctx.view(B, S, H, D).transpose(1, 2)
        dq, dk, dv = torch.autograd.grad(out, (ql, kl, vl), dout)
        ctx.graph = None
        dqkv = torch.empty(B, S, 3, H, D, dtype=dctx.dtype, device=dctx.device)
        if rot_dim > 0:
            if dq.stride(-1) != 1 or dq.stride() != dk.stride():
                dq, dk = dq.contiguous(), dk.contiguous()
            if dv.stride(-1) == 1 and all(st % 8 == 0 for st in dv.stride()[:3]):
                # dV rides along in the same launch (its own vectorised strided copy was 1.4 % of the step)
                rotary_oop(dq, dk, dqkv[:, :, 0], dqkv[:, :, 1], rot_dim, base, inverse=True, v_in=dv,
                           v_out=dqkv[:, :, 2])
            else:
                rotary_oop(dq, dk, dqkv[:, :, 0], dqkv[:, :, 1], rot_dim, base, inverse=True)
                dqkv[:, :, 2].copy_(dv)
        else:
            dqkv[:, :, 0].copy_(dq)
            dqkv[:, :, 1].copy_(dk)
            dqkv[:, :, 2].copy_(dv)
        return dqkv.view(B, S, 3 * H * D), None, None, None, None


def causal_self_attention(qkv, H, D, rot_dim, base=10000.0):
    """qkv: [B, S, 3*H*D] laid out (3, H, D). Returns the attention context [B, S, H*D]."""
    if qkv.is_cuda and qkv.dtype in (torch.float16, torch.bfloat16) and D % 8 == 0 and (rot_dim % 16 == 0):
        N.cuda_lib()
        if _BACKEND == "native" and D == 128 and qkv.shape[1] % 128 == 0:
            return NativeCausalSelfAttentionFn.apply(qkv.contiguous(), H, D, rot_dim, base)
        return CausalSelfAttentionFn.apply(qkv.contiguous(), H, D, rot_dim, base)
    B, S = qkv.shape[0], qkv.shape[1]
    t = qkv.view(B, S, 3, H, D)
    q, k, v = t[:, :, 0], t[:, :, 1], t[:, :, 2]
    if rot_dim > 0:
        q, k = K._rotary_ref(q, k, rot_dim, base, False)
    ctx = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
    return ctx.transpose(1, 2).reshape(B, S, H * D)
