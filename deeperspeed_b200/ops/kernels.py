"""Functional + autograd wrappers over the sm_100a kernel library.

CUDA tensors always take the native path (missing library => hard error). CPU tensors take an equivalent plain
PyTorch path so the engine/pipeline/ZeRO logic is testable without a GPU; the same torch code is the oracle the
GPU numerics tests compare against.
"""
import math
import os

import torch
import torch.nn.functional as F

from . import native as N

_launches = 0  # kernels of ours launched (bench.py reports this as `gpu_launches`)


def launch_count():
    return _launches


def _count(n=1):
    global _launches
    _launches += n


# ------------------------------------------------------------------------------------------------ GEMM
# Tile policy: "auto" lets the native dispatcher choose (autotuned variant if the problem was tuned, else CTA-pair
# 256 x 256 tiles for large outputs and single-CTA 128 x 256 / 128 x 128 tiles otherwise); "pair" / "single" force
# one (DSB_GEMM_KERNEL env, tools/bench_gemm.py).
GEMM_KERNEL = os.environ.get("DSB_GEMM_KERNEL", "auto")


def _use_pair(M, Nn, K, batch):
    if GEMM_KERNEL == "pair":
        return True
    if GEMM_KERNEL == "single":
        return False
    return M >= 256 and Nn >= 128 and ((M + 255) // 256) * ((Nn + 127) // 128) * batch >= 64


def gemm(a, b, *, trans_a=False, trans_b=True, epilogue=N.EPI_NONE, alpha=1.0, bias=None, aux=None, aux_out=None,
         out=None):
    """D = epi(alpha * op(a) @ op(b)).

    Layout convention (row-major 2-D or batched 3-D tensors, last dim contiguous):
      trans_a=False: a is [M, K];  trans_a=True: a is [K, M]  (the kernel reads it MN-major, no copy)
      trans_b=True : b is [N, K] (nn.Linear weight layout); trans_b=False: b is [K, N]
    """
    batched = a.dim() == 3
    if trans_a:
        K, M = a.shape[-2], a.shape[-1]
    else:
        M, K = a.shape[-2], a.shape[-1]
    if trans_b:
        Nn = b.shape[-2]
        assert b.shape[-1] == K, (a.shape, b.shape, trans_a, trans_b)
    else:
        Nn = b.shape[-1]
        assert b.shape[-2] == K, (a.shape, b.shape, trans_a, trans_b)
    batch = a.shape[0] if batched else 1
    if not N.use_native(a):
        return _gemm_ref(a, b, trans_a, trans_b, epilogue, alpha, bias, aux, aux_out, out)
    assert a.stride(-1) == 1 and b.stride(-1) == 1
    if out is None:
        out = torch.empty((batch, M, Nn) if batched else (M, Nn), dtype=a.dtype, device=a.device)
    lib = N.cuda_lib()
    sa = a.stride(0) if batched else 0
    sb = (b.stride(0) if b.dim() == 3 else 0) if batched else 0
    sd = out.stride(0) if batched else 0
    if batched and b.dim() == 2:
        sb = 0
    s_aux = aux.stride(0) if (aux is not None and aux.dim() == 3) else 0
    s_auxo = aux_out.stride(0) if (aux_out is not None and aux_out.dim() == 3) else 0
    if GEMM_KERNEL == "auto":
        fn = lib.dsb_gemm_auto      # tuned table (ops/gemm_autotune.py) first, static rule otherwise
    else:
        fn = lib.dsb_gemm_pair if _use_pair(M, Nn, K, batch) else lib.dsb_gemm
    rc = fn(N.ptr(a), N.ptr(b), N.ptr(out), M, Nn, K, batch, a.stride(-2), b.stride(-2), out.stride(-2),
            sa, sb, sd, int(trans_a), int(not trans_b), N.dtype_code(a.dtype), epilogue, float(alpha),
            N.ptr(bias), N.ptr(aux), aux.stride(-2) if aux is not None else 0, s_aux, N.ptr(aux_out),
            aux_out.stride(-2) if aux_out is not None else 0, s_auxo, N.stream_ptr())
    N.check(rc, "dsb_gemm")
    _count()
    return out


def gelu_tanh(x):
    return 0.5 * x * (1.0 + torch.tanh(0.7978845608028654 * x * (1.0 + 0.044715 * x * x)))


def gelu_tanh_grad(x):
    k0, k1 = 0.7978845608028654, 0.044715
    u = k0 * x * (1.0 + k1 * x * x)
    t = torch.tanh(u)
    return 0.5 * (1.0 + t) + 0.5 * x * (1.0 - t * t) * k0 * (1.0 + 3.0 * k1 * x * x)


def _gemm_ref(a, b, trans_a, trans_b, epilogue, alpha, bias, aux, aux_out, out):
    dt = a.dtype
    af = a.float().transpose(-1, -2) if trans_a else a.float()
    bf = b.float().transpose(-1, -2) if trans_b else b.float()
    acc = torch.matmul(af, bf) * alpha
    if bias is not None and epilogue != N.EPI_DGELU:
        acc = acc + bias.float()
    if epilogue == N.EPI_BIAS_GELU:
        if aux_out is not None:
            aux_out.copy_(acc.to(dt))
        acc = gelu_tanh(acc)
    elif epilogue == N.EPI_BIAS_RESIDUAL:
        acc = acc + aux.float()
    elif epilogue == N.EPI_DGELU:
        acc = acc * gelu_tanh_grad(aux.float())
    res = acc.to(dt)
    if out is not None:
        out.copy_(res)
        return out
    return res


ACCUMULATE_INTO_GRAD = True   # wgrad GEMMs add straight into an existing .grad (arena view) via the epilogue


def _wgrad(dy2, x2, w):
    """dW = dy^T x. If the parameter already owns a gradient buffer (a view of the flat gradient arena), the GEMM
    epilogue accumulates into it in place (no dW temporary, no separate add pass) and autograd gets None."""
    g = w.grad
    if ACCUMULATE_INTO_GRAD and g is not None and g.dtype == dy2.dtype and g.is_contiguous() and g.shape == w.shape:
        gemm(dy2, x2, trans_a=True, trans_b=False, epilogue=N.EPI_BIAS_RESIDUAL, aux=g, out=g)
        return None
    return gemm(dy2, x2, trans_a=True, trans_b=False)


def _bgrad(dy2, b):
    g = b.grad
    if ACCUMULATE_INTO_GRAD and g is not None and g.dtype == dy2.dtype and g.is_contiguous():
        column_sum(dy2, out=g, accumulate=True)
        return None
    return column_sum(dy2)


class LinearFn(torch.autograd.Function):
    """y = x @ W^T (+ b) with tcgen05 GEMMs for fwd / dgrad / wgrad; wgrad accumulates straight into `w.grad`
    storage when the engine has pre-allocated it in the flat gradient arena (no flatten/unflatten copies)."""

    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.reshape(-1, x.shape[-1])
        y = gemm(x2, w, trans_b=True, epilogue=N.EPI_BIAS if b is not None else N.EPI_NONE, bias=b)
        ctx.save_for_backward(x2, w, b)
        ctx.has_bias = b is not None
        ctx.x_shape = x.shape
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w, b = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm(dy2, w, trans_b=False).view(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            dw = _wgrad(dy2, x2, w)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = _bgrad(dy2, b)
        return dx, dw, db


def linear(x, w, b=None):
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16):
        return LinearFn.apply(x, w, b)
    return F.linear(x, w, b)


class LinearGeluFn(torch.autograd.Function):
    """h = gelu(x @ W^T + b): bias+GeLU run in the GEMM epilogue (the pre-activation is kept for backward, where
    gelu' is applied by a bandwidth kernel before the dgrad/wgrad GEMMs)."""

    @staticmethod
    def forward(ctx, x, w, b):
        x2 = x.reshape(-1, x.shape[-1])
        pre = torch.empty((x2.shape[0], w.shape[0]), dtype=x.dtype, device=x.device)
        h = gemm(x2, w, trans_b=True, epilogue=N.EPI_BIAS_GELU, bias=b, aux_out=pre)
        ctx.save_for_backward(x2, w, pre, b)
        ctx.x_shape = x.shape
        return h.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dh):
        x2, w, pre, b = ctx.saved_tensors
        dh2 = dh.reshape(-1, dh.shape[-1])
        dpre = bias_gelu_bwd(dh2.contiguous(), pre, None)
        dx = gemm(dpre, w, trans_b=False).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = _wgrad(dpre, x2, w) if ctx.needs_input_grad[1] else None
        db = _bgrad(dpre, b) if ctx.needs_input_grad[2] else None
        return dx, dw, db


def linear_gelu(x, w, b):
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16):
        return LinearGeluFn.apply(x, w, b)
    return gelu_tanh(F.linear(x, w, b))


# ------------------------------------------------------------------------------------------------ reductions / misc
def column_sum(x, out=None, accumulate=False):
    rows, cols = x.shape
    if not N.use_native(x):
        r = x.float().sum(0).to(x.dtype)
        if out is not None:
            out.copy_(out + r if accumulate else r)
            return out
        return r
    if out is None:
        out = torch.empty(cols, dtype=x.dtype, device=x.device)
    sc = N.scratch(cols * 4, x.device, "colsum")
    N.check(N.cuda_lib().dsb_column_sum(N.ptr(x), N.ptr(out), rows, cols, N.dtype_code(x.dtype), int(accumulate),
                                        N.ptr(sc), N.stream_ptr()), "dsb_column_sum")
    _count(3)
    return out


def fused_add(a, b, c=None, d=None, out=None):
    if not N.use_native(a):
        r = a + b
        if c is not None:
            r = r + c
        if d is not None:
            r = r + d
        return r
    if out is None:
        out = torch.empty_like(a)
    N.check(N.cuda_lib().dsb_fused_add(N.ptr(a), N.ptr(b), N.ptr(c), N.ptr(d), N.ptr(out), a.numel(),
                                       N.dtype_code(a.dtype), N.stream_ptr()), "dsb_fused_add")
    _count()
    return out


def sumsq(x, out=None, flag=None, accumulate=False):
    """Sum of squares of a flat tensor into a device scalar (+ non-finite flag), no host sync."""
    if out is None:
        out = torch.zeros(1, dtype=torch.float32, device=x.device)
    if not N.use_native(x):
        xf = x.float()
        s = (xf * xf).sum()
        out.copy_(out + s if accumulate else s)
        if flag is not None and not bool(torch.isfinite(s)):
            flag.fill_(1)
        return out
    sc = N.scratch(N.cuda_lib().dsb_sumsq_scratch_bytes(), x.device, "sumsq")
    N.check(N.cuda_lib().dsb_sumsq(N.ptr(x), N.dtype_code(x.dtype), x.numel(), N.ptr(out), N.ptr(flag),
                                   int(accumulate), N.ptr(sc), N.stream_ptr()), "dsb_sumsq")
    _count(2)
    return out


def scale_(x, scale=1.0, scale_dev=None):
    if not N.use_native(x):
        s = scale * (float(scale_dev) if scale_dev is not None else 1.0)
        return x.mul_(s)
    N.check(N.cuda_lib().dsb_scale(N.ptr(x), N.dtype_code(x.dtype), x.numel(), float(scale), N.ptr(scale_dev),
                                   N.stream_ptr()), "dsb_scale")
    _count()
    return x


def cast(src, dst):
    if not N.use_native(src):
        dst.copy_(src)
        return dst
    N.check(N.cuda_lib().dsb_cast(N.ptr(src), N.dtype_code(src.dtype), N.ptr(dst), N.dtype_code(dst.dtype),
                                  src.numel(), N.stream_ptr()), "dsb_cast")
    _count()
    return dst


# ------------------------------------------------------------------------------------------------ optimizers
def adam_flat(p, m, v, g, *, lp=None, lr, beta1, beta2, eps, weight_decay, adamw=True, bias_correction=True, step=1,
              inv_scale=1.0, inv_scale_dev=None, skip_dev=None):
    """One fused pass over a flat fp32 master/m/v partition (+ optional low-precision write-back)."""
    if not N.use_native(p):
        if skip_dev is not None and int(skip_dev) != 0:
            return
        s = inv_scale * (float(inv_scale_dev) if inv_scale_dev is not None else 1.0)
        gf = g.float() * s
        if not adamw:
            gf = gf + weight_decay * p
        m.mul_(beta1).add_(gf, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(gf, gf, value=1 - beta2)
        bc1 = 1 - beta1 ** step if bias_correction else 1.0
        bc2 = 1 - beta2 ** step if bias_correction else 1.0
        upd = (m / bc1) / ((v / bc2).sqrt() + eps)
        if adamw:
            upd = upd + weight_decay * p
        p.add_(upd, alpha=-lr)
        if lp is not None:
            lp.copy_(p)
        return
    N.check(N.cuda_lib().dsb_adam_flat(N.ptr(p), N.ptr(m), N.ptr(v), N.ptr(g), N.dtype_code(g.dtype), N.ptr(lp),
                                       N.dtype_code(lp.dtype) if lp is not None else 0, p.numel(), float(lr),
                                       float(beta1), float(beta2), float(eps), float(weight_decay), int(adamw),
                                       int(bias_correction), int(step), float(inv_scale), N.ptr(inv_scale_dev),
                                       N.ptr(skip_dev), N.stream_ptr()), "dsb_adam_flat")
    _count()


# ------------------------------------------------------------------------------------------------ LayerNorm
def _ln_ref(x, res, gamma, beta, eps):
    s = x.float() + (res.float() if res is not None else 0.0)
    if res is not None:
        s = s.to(x.dtype).float()
    mu = s.mean(-1, keepdim=True)
    var = ((s - mu) ** 2).mean(-1, keepdim=True)
    rstd = torch.rsqrt(var + eps)
    y = (s - mu) * rstd * gamma.float() + beta.float()
    return y.to(x.dtype), s.to(x.dtype), mu.squeeze(-1), rstd.squeeze(-1)


class LayerNormFn(torch.autograd.Function):
    """y = LN(x [+ residual]); returns (y, s) with s = x + residual produced by the same pass. Without a residual, s
    is an alias of x: routing the residual stream through it hands the stream's gradient to THIS backward, which
    adds it inside the LN-backward kernel (`dres`) instead of leaving a separate accumulation add to autograd.
    gamma/beta gradients are accumulated straight into existing `.grad` buffers by the finalize kernel."""

    @staticmethod
    def forward(ctx, x, residual, gamma, beta, eps):
        hidden = x.shape[-1]
        x2 = x.reshape(-1, hidden)
        rows = x2.shape[0]
        r2 = residual.reshape(-1, hidden) if residual is not None else None
        y = torch.empty_like(x2)
        ssum = torch.empty_like(x2) if residual is not None else x2
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        N.check(N.cuda_lib().dsb_layernorm_fwd(N.ptr(x2), N.ptr(r2), N.ptr(gamma), N.ptr(beta), N.ptr(y),
                                               N.ptr(ssum) if residual is not None else 0, N.ptr(mean), N.ptr(rstd),
                                               rows, hidden, float(eps), N.dtype_code(x.dtype), N.stream_ptr()),
                "dsb_layernorm_fwd")
        _count()
        ctx.save_for_backward(ssum, mean, rstd, gamma, beta)
        ctx.has_res = residual is not None
        ctx.shape = x.shape
        ctx.set_materialize_grads(False)
        if residual is not None:
            return y.view(x.shape), ssum.view(x.shape)
        return y.view(x.shape), x.detach()

    @staticmethod
    def backward(ctx, dy, dsum):
        ssum, mean, rstd, gamma, beta = ctx.saved_tensors
        hidden = ssum.shape[-1]
        rows = ssum.shape[0]
        if dy is None:   # only the pass-through stream was used downstream
            return dsum, (dsum if ctx.has_res else None), None, None, None
        dy2 = dy.reshape(-1, hidden).contiguous()
        dres = dsum.reshape(-1, hidden).contiguous() if dsum is not None else None
        dx = torch.empty_like(ssum)
        gg, gb = gamma.grad, beta.grad
        acc = (ACCUMULATE_INTO_GRAD and gg is not None and gb is not None and gg.dtype == ssum.dtype
               and gb.dtype == ssum.dtype and gg.is_contiguous() and gb.is_contiguous())
        if acc:
            dgamma, dbeta = gg, gb
        else:
            dgamma = torch.empty_like(gamma)
            dbeta = torch.empty_like(gamma)
        lib = N.cuda_lib()
        sc = N.scratch(lib.dsb_layernorm_bwd_scratch_bytes(hidden), ssum.device, "ln_bwd")
        N.check(lib.dsb_layernorm_bwd_acc(N.ptr(dy2), N.ptr(ssum), N.ptr(mean), N.ptr(rstd), N.ptr(gamma), N.ptr(dres),
                                          N.ptr(dx), N.ptr(dgamma), N.ptr(dbeta), rows, hidden,
                                          N.dtype_code(ssum.dtype), int(acc), N.ptr(sc), N.stream_ptr()),
                "dsb_layernorm_bwd")
        _count(2)
        dx = dx.view(ctx.shape)
        if acc:
            dgamma = dbeta = None
        return dx, (dx if ctx.has_res else None), dgamma, dbeta, None


class LayerNormInvertibleFn(torch.autograd.Function):
    """y = LN(x) keeping only y (and rstd) for backward: the kernel recovers x-hat = (y - beta) / gamma. Saves the
    activation-sized input tensor per LayerNorm (DeepSpeedTransformerConfig.normalize_invertible)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        hidden = x.shape[-1]
        x2 = x.reshape(-1, hidden)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        N.check(N.cuda_lib().dsb_layernorm_fwd(N.ptr(x2), 0, N.ptr(gamma), N.ptr(beta), N.ptr(y), 0, N.ptr(mean),
                                               N.ptr(rstd), rows, hidden, float(eps), N.dtype_code(x.dtype),
                                               N.stream_ptr()), "dsb_layernorm_fwd")
        _count()
        ctx.save_for_backward(y, rstd, gamma, beta)
        ctx.shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        y, rstd, gamma, beta = ctx.saved_tensors
        rows, hidden = y.shape
        dy2 = dy.reshape(-1, hidden).contiguous()
        dx = torch.empty_like(y)
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        lib = N.cuda_lib()
        sc = N.scratch(lib.dsb_layernorm_bwd_scratch_bytes(hidden), y.device, "ln_bwd")
        N.check(lib.dsb_layernorm_bwd_inv(N.ptr(dy2), N.ptr(y), N.ptr(rstd), N.ptr(gamma), N.ptr(beta), 0, N.ptr(dx),
                                          N.ptr(dgamma), N.ptr(dbeta), rows, hidden, N.dtype_code(y.dtype), 0,
                                          N.ptr(sc), N.stream_ptr()), "dsb_layernorm_bwd_inv")
        _count(2)
        return dx.view(ctx.shape), dgamma, dbeta, None


def layer_norm_invertible(x, gamma, beta, eps=1e-5):
    """LayerNorm whose backward needs only its output (see LayerNormInvertibleFn); plain LN where the kernel is absent."""
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16) and x.shape[-1] % 8 == 0 and x.shape[-1] <= 8192:
        N.cuda_lib()
        return LayerNormInvertibleFn.apply(x, gamma, beta, eps)
    return F.layer_norm(x, (x.shape[-1],), gamma, beta, eps)


def layer_norm(x, gamma, beta, eps=1e-5, residual=None):
    """Returns (LN(x + residual), x + residual). With residual=None the second output is an alias of x: use it (not
    x) as the residual stream downstream so the stream's gradient is folded into the LN backward kernel."""
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16) and x.shape[-1] % 8 == 0 and x.shape[-1] <= 8192:
        N.cuda_lib()
        return LayerNormFn.apply(x.contiguous(), residual.contiguous() if residual is not None else None, gamma,
                                 beta, eps)
    s = x if residual is None else x + residual
    return F.layer_norm(s, (s.shape[-1],), gamma, beta, eps), s


# ------------------------------------------------------------------------------------------------ softmax / gelu / dropout
def softmax_fwd(x, mask=None, scale=1.0, causal=False, out=None):
    """x: [B, H, Sq, Sk]; mask additive [B, Sk]."""
    B, H, Sq, Sk = x.shape
    if not N.use_native(x):
        s = x.float() * scale
        if mask is not None:
            s = s + mask.float().view(B, 1, 1, Sk)
        if causal:
            idx = torch.arange(Sk, device=x.device).view(1, 1, 1, Sk)
            qi = torch.arange(Sq, device=x.device).view(1, 1, Sq, 1) + (Sk - Sq)
            s = s.masked_fill(idx > qi, float("-inf"))
        return torch.softmax(s, -1).to(x.dtype)
    if out is None:
        out = torch.empty_like(x)
    N.check(N.cuda_lib().dsb_softmax_fwd(N.ptr(x), N.ptr(mask), N.ptr(out), B, H, Sq, Sk, float(scale), int(causal),
                                         N.dtype_code(x.dtype), N.stream_ptr()), "dsb_softmax_fwd")
    _count()
    return out


def softmax_bwd(dy, y, scale=1.0, out=None):
    if not N.use_native(y):
        yf, gf = y.float(), dy.float()
        return (yf * (gf - (yf * gf).sum(-1, keepdim=True)) * scale).to(y.dtype)
    if out is None:
        out = torch.empty_like(y)
    rows = y.numel() // y.shape[-1]
    N.check(N.cuda_lib().dsb_softmax_bwd(N.ptr(dy), N.ptr(y), N.ptr(out), rows, y.shape[-1], float(scale),
                                         N.dtype_code(y.dtype), N.stream_ptr()), "dsb_softmax_bwd")
    _count()
    return out


def bias_gelu_fwd(x, bias):
    if not N.use_native(x):
        return gelu_tanh(x.float() + (bias.float() if bias is not None else 0.0)).to(x.dtype)
    y = torch.empty_like(x)
    cols = x.shape[-1]
    N.check(N.cuda_lib().dsb_bias_gelu_fwd(N.ptr(x), N.ptr(bias), N.ptr(y), x.numel() // cols, cols,
                                           N.dtype_code(x.dtype), N.stream_ptr()), "dsb_bias_gelu_fwd")
    _count()
    return y


def bias_gelu_bwd(dy, x, bias):
    if not N.use_native(x):
        return (dy.float() * gelu_tanh_grad(x.float() + (bias.float() if bias is not None else 0.0))).to(x.dtype)
    dx = torch.empty_like(x)
    cols = x.shape[-1]
    N.check(N.cuda_lib().dsb_bias_gelu_bwd(N.ptr(dy), N.ptr(x), N.ptr(bias), N.ptr(dx), x.numel() // cols, cols,
                                           N.dtype_code(x.dtype), N.stream_ptr()), "dsb_bias_gelu_bwd")
    _count()
    return dx


def dropout_fwd(x, ratio, seed, offset, bias=None, residual=None):
    """Returns (out, mask uint8). out = dropout(x + bias) + residual."""
    cols = x.shape[-1]
    if not N.use_native(x):
        g = torch.Generator(device="cpu").manual_seed(int(seed) + int(offset))
        keep = (torch.rand(x.shape, generator=g) >= ratio)
        v = x.float() + (bias.float() if bias is not None else 0.0)
        v = torch.where(keep, v / (1.0 - ratio), torch.zeros_like(v))
        if residual is not None:
            v = v + residual.float()
        return v.to(x.dtype), keep.to(torch.uint8)
    out = torch.empty_like(x)
    mask = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    N.check(N.cuda_lib().dsb_dropout_fwd(N.ptr(x), N.ptr(bias), N.ptr(residual), N.ptr(out), N.ptr(mask),
                                         x.numel() // cols, cols, float(ratio), int(seed), int(offset),
                                         N.dtype_code(x.dtype), N.stream_ptr()), "dsb_dropout_fwd")
    _count()
    return out, mask


def dropout_bwd(dy, mask, ratio):
    if not N.use_native(dy):
        return (dy.float() * mask.float() / (1.0 - ratio)).to(dy.dtype)
    dx = torch.empty_like(dy)
    N.check(N.cuda_lib().dsb_dropout_bwd(N.ptr(dy), N.ptr(mask), N.ptr(dx), dy.numel(), float(ratio),
                                         N.dtype_code(dy.dtype), N.stream_ptr()), "dsb_dropout_bwd")
    _count()
    return dx


def qkv_bias_split_transpose(qkv, bias, B, S, H, D):
    """qkv [B, S, 3*H*D] laid out (3, H, D) -> q, k, v [B, H, S, D] with bias added."""
    if not N.use_native(qkv):
        t = qkv.view(B, S, 3, H, D)
        if bias is not None:
            t = t + bias.view(1, 1, 3, H, D)
        q, k, v = (t[:, :, i].permute(0, 2, 1, 3).contiguous() for i in range(3))
        return q, k, v
    q = torch.empty((B, H, S, D), dtype=qkv.dtype, device=qkv.device)
    k, v = torch.empty_like(q), torch.empty_like(q)
    N.check(N.cuda_lib().dsb_qkv_bias_split_transpose(N.ptr(qkv), N.ptr(bias), N.ptr(q), N.ptr(k), N.ptr(v), B, S, H,
                                                      D, N.dtype_code(qkv.dtype), N.stream_ptr()), "dsb_qkv_split")
    _count()
    return q, k, v


def transpose_0213(x):
    """[B, d1, d2, D] -> [B, d2, d1, D] contiguous."""
    B, d1, d2, D = x.shape
    if not N.use_native(x):
        return x.permute(0, 2, 1, 3).contiguous()
    y = torch.empty((B, d2, d1, D), dtype=x.dtype, device=x.device)
    N.check(N.cuda_lib().dsb_transpose_0213(N.ptr(x), N.ptr(y), B, d1, d2, D, N.dtype_code(x.dtype), N.stream_ptr()),
            "dsb_transpose_0213")
    _count()
    return y


# ------------------------------------------------------------------------------------------------ cross entropy
class EmbeddingFn(torch.autograd.Function):
    """Row gather whose backward adds the output gradient straight into the table's existing `.grad` (a view of the
    flat gradient arena) with packed 16-bit atomics: no dense [vocab, dim] temporary, no sort, no AccumulateGrad pass
    (csrc/transformer/elementwise.cu). Falls back to torch where that does not apply."""

    @staticmethod
    def forward(ctx, ids, table):
        ids_c = ids.contiguous()
        out = torch.empty(*ids.shape, table.shape[1], dtype=table.dtype, device=table.device)
        N.check(N.cuda_lib().dsb_embedding_fwd(N.ptr(table), N.ptr(ids_c), N.ptr(out), ids_c.numel(), table.shape[1],
                                               table.shape[0], N.dtype_code(table.dtype), N.stream_ptr()),
                "dsb_embedding_fwd")
        _count()
        ctx.save_for_backward(ids_c)
        # the table's shape is remembered, not re-read in backward: under ZeRO-3 the parameter may be back to its
        # partitioned placeholder by then (its gradient is then returned whole and the ZeRO hooks scatter it)
        ctx.table, ctx.table_shape = table, tuple(table.shape)
        return out

    @staticmethod
    def backward(ctx, dout):
        (ids,) = ctx.saved_tensors
        vocab, dim = ctx.table_shape
        g = ctx.table.grad
        into_arena = (ACCUMULATE_INTO_GRAD and g is not None and g.dtype == dout.dtype and g.is_contiguous()
                      and tuple(g.shape) == (vocab, dim))
        if not into_arena:
            g = torch.zeros(vocab, dim, dtype=dout.dtype, device=dout.device)
        d2 = dout.reshape(-1, dim)
        if not d2.is_contiguous():
            d2 = d2.contiguous()
        N.check(N.cuda_lib().dsb_embedding_bwd_acc(N.ptr(d2), N.ptr(ids), N.ptr(g), ids.numel(), dim, vocab,
                                                   N.dtype_code(dout.dtype), N.stream_ptr()),
                "dsb_embedding_bwd_acc")
        _count()
        return None, (None if into_arena else g)


def embedding(ids, table):
    """table[ids]; on CUDA in fp16 / bf16 with a 16-byte-aligned row size our gather / scatter-add kernels."""
    if (table.is_cuda and table.dtype in (torch.float16, torch.bfloat16) and table.dim() == 2 and table.shape[1] % 8 == 0
            and table.is_contiguous() and ids.dtype == torch.int64):
        return EmbeddingFn.apply(ids, table)
    return torch.nn.functional.embedding(ids, table)


class CrossEntropyFn(torch.autograd.Function):
    """Mean token cross entropy over a [tokens, vocab] half/bf16 logits matrix. Backward overwrites the logits
    buffer with the gradient (no second [tokens, vocab] allocation, no fp32 copy of the logits)."""

    @staticmethod
    def forward(ctx, logits, targets, ignore_index):
        tokens, vocab = logits.shape
        loss = torch.empty(tokens, dtype=torch.float32, device=logits.device)
        lse = torch.empty(tokens, dtype=torch.float32, device=logits.device)
        N.check(N.cuda_lib().dsb_cross_entropy_fwd(N.ptr(logits), N.ptr(targets), N.ptr(loss), N.ptr(lse), tokens,
                                                   vocab, logits.stride(0), int(ignore_index),
                                                   N.dtype_code(logits.dtype), N.stream_ptr()), "dsb_ce_fwd")
        _count()
        valid = (targets != ignore_index).sum().clamp(min=1).float()
        ctx.save_for_backward(logits, targets, lse, valid)
        ctx.ignore_index = ignore_index
        return loss.sum() / valid

    @staticmethod
    def backward(ctx, gout):
        logits, targets, lse, valid = ctx.saved_tensors
        tokens, vocab = logits.shape
        gs = (gout.float() / valid).reshape(1).contiguous()
        N.check(N.cuda_lib().dsb_cross_entropy_bwd(N.ptr(logits), N.ptr(targets), N.ptr(lse), N.ptr(gs), 1.0, tokens,
                                                   vocab, logits.stride(0), int(ctx.ignore_index),
                                                   N.dtype_code(logits.dtype), N.stream_ptr()), "dsb_ce_bwd")
        _count()
        return logits, None, None


def cross_entropy(logits, targets, ignore_index=-100):
    lg = logits.reshape(-1, logits.shape[-1])
    tg = targets.reshape(-1)
    if lg.is_cuda and lg.dtype in (torch.float16, torch.bfloat16) and lg.stride(0) % 8 == 0 and lg.stride(1) == 1:
        return CrossEntropyFn.apply(lg, tg.contiguous(), ignore_index)
    return F.cross_entropy(lg.float(), tg, ignore_index=ignore_index)


# ------------------------------------------------------------------------------------------------ rotary
def _rotary_ref(q, k, rot_dim, base, inverse):
    S, D = q.shape[1], q.shape[-1]
    half = rot_dim // 2
    inv_freq = 1.0 / (base ** (torch.arange(0, half, device=q.device, dtype=torch.float32) * 2 / rot_dim))
    ang = torch.arange(S, device=q.device, dtype=torch.float32)[:, None] * inv_freq[None]
    cs, sn = ang.cos()[None, :, None, :], ang.sin()[None, :, None, :]
    if inverse:
        sn = -sn

    def rot(t):
        tf = t.float()
        x0, x1 = tf[..., :half], tf[..., half:rot_dim]
        out = tf.clone()
        out[..., :half] = x0 * cs - x1 * sn
        out[..., half:rot_dim] = x1 * cs + x0 * sn
        return out.to(t.dtype)

    return rot(q), rot(k)


class RotaryFn(torch.autograd.Function):
    """In-place NeoX-style rotary on q, k views of shape [B, S, H, D] (any strides with unit last stride)."""

    @staticmethod
    def forward(ctx, q, k, rot_dim, base):
        ctx.rot_dim, ctx.base = rot_dim, base
        _rotary_inplace(q, k, rot_dim, base, False)
        ctx.mark_dirty(q, k)
        return q, k

    @staticmethod
    def backward(ctx, dq, dk):
        dq, dk = dq.contiguous(), dk.contiguous()
        _rotary_inplace(dq, dk, ctx.rot_dim, ctx.base, True)
        return dq, dk, None, None


def _rotary_inplace(q, k, rot_dim, base, inverse):
    B, S, H, D = q.shape
    assert q.stride() == k.stride() and q.stride(-1) == 1
    N.check(N.cuda_lib().dsb_rotary(N.ptr(q), N.ptr(k), B, S, H, D, rot_dim, q.stride(0), q.stride(1), q.stride(2),
                                    float(base), int(inverse), N.dtype_code(q.dtype), N.stream_ptr()), "dsb_rotary")
    _count()


def rotary(q, k, rot_dim, base=10000.0):
    """q, k: [B, S, H, D]. Returns rotated (q, k)."""
    if q.is_cuda and q.dtype in (torch.float16, torch.bfloat16):
        N.cuda_lib()
        q = q.contiguous() if q.stride(-1) != 1 else q
        k = k.contiguous() if k.stride(-1) != 1 else k
        if q.stride() != k.stride():
            q, k = q.contiguous(), k.contiguous()
        return RotaryFn.apply(q.clone(), k.clone(), rot_dim, base)
    return _rotary_ref(q, k, rot_dim, base, False)


class LinearResidualFn(torch.autograd.Function):
    """y = x @ W^T + b + r with the bias and residual add fused into the GEMM epilogue."""

    @staticmethod
    def forward(ctx, x, w, b, r):
        x2 = x.reshape(-1, x.shape[-1])
        r2 = r.reshape(-1, r.shape[-1])
        if not r2.is_contiguous():
            r2 = r2.contiguous()
        y = gemm(x2, w, trans_b=True, epilogue=N.EPI_BIAS_RESIDUAL, bias=b, aux=r2)
        ctx.save_for_backward(x2, w, b)
        ctx.has_bias = b is not None
        ctx.x_shape = x.shape
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w, b = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = gemm(dy2, w, trans_b=False).view(ctx.x_shape) if ctx.needs_input_grad[0] else None
        dw = _wgrad(dy2, x2, w) if ctx.needs_input_grad[1] else None
        db = _bgrad(dy2, b) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return dx, dw, db, (dy if ctx.needs_input_grad[3] else None)


def linear_residual(x, w, b, r):
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16):
        return LinearResidualFn.apply(x, w, b, r)
    return F.linear(x, w, b) + r


class MlpResidualFn(torch.autograd.Function):
    """y = gelu(x @ W1^T + b1) @ W2^T + b2 + r as four GEMMs with every elementwise step in an epilogue:
    forward bias+GeLU (keeps the pre-activation) and bias+residual; backward gelu'(pre) is applied in the epilogue
    of the FF2 dgrad GEMM, so the 4h-wide gradient is written once and never re-read by a separate kernel
    (the reference runs d_gelu as its own pass, csrc/transformer/gelu_kernels.cu:187-330)."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, r):
        x2 = x.reshape(-1, x.shape[-1])
        r2 = r.reshape(-1, r.shape[-1])
        if not r2.is_contiguous():
            r2 = r2.contiguous()
        pre = torch.empty((x2.shape[0], w1.shape[0]), dtype=x.dtype, device=x.device)
        h = gemm(x2, w1, trans_b=True, epilogue=N.EPI_BIAS_GELU, bias=b1, aux_out=pre)
        y = gemm(h, w2, trans_b=True, epilogue=N.EPI_BIAS_RESIDUAL, bias=b2, aux=r2)
        ctx.save_for_backward(x2, w1, b1, w2, b2, pre, h)
        ctx.x_shape = x.shape
        return y.view(*x.shape[:-1], w2.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w1, b1, w2, b2, pre, h = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        need = ctx.needs_input_grad
        dw2 = _wgrad(dy2, h, w2) if need[3] else None
        db2 = _bgrad(dy2, b2) if (b2 is not None and need[4]) else None
        dx = dw1 = db1 = None
        if need[0] or need[1] or need[2]:
            dpre = gemm(dy2, w2, trans_b=False, epilogue=N.EPI_DGELU, aux=pre)
            if need[0]:
                dx = gemm(dpre, w1, trans_b=False).view(ctx.x_shape)
            if need[1]:
                dw1 = _wgrad(dpre, x2, w1)
            if b1 is not None and need[2]:
                db1 = _bgrad(dpre, b1)
        return dx, dw1, db1, dw2, db2, (dy if need[5] else None)


def mlp_residual(x, w1, b1, w2, b2, r):
    if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16):
        return MlpResidualFn.apply(x, w1, b1, w2, b2, r)
    return F.linear(gelu_tanh(F.linear(x, w1, b1)), w2, b2) + r
