"""Block-sparse softmax (API of reference deepspeed/ops/sparse_attention/softmax.py:17-315):
softmax(x * scale + rpe + key_padding_mask + attn_mask) over the non-zero blocks of each row."""
import torch

from .. import kernels as K
from .. import native as N
from .matmul import get_lut, sparse_to_dense, dense_to_sparse


class _SparseSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale, rpe, key_padding_mask, attn_mask, kp_mask_mode, attn_mask_mode, layout, block, lut):
        B = x.shape[0]
        H, nb = layout.shape[0], layout.shape[1]
        S = nb * block
        if x.is_cuda and x.dtype in (torch.float16, torch.bfloat16):
            y = x.contiguous().clone()
            _, row_ptr, row_idx, _, _ = lut.on(x.device)

            def prep(t):
                return None if t is None else t.to(x.dtype).contiguous()

            rpe_, kpm_, am_ = prep(rpe), prep(key_padding_mask), prep(attn_mask)
            N.check(N.cuda_lib().dsb_bs_softmax_fwd(N.ptr(y), N.ptr(row_ptr), N.ptr(row_idx), B, H, S, block, lut.nnz,
                                                    float(scale), N.ptr(rpe_), N.ptr(kpm_), int(kp_mask_mode == "mul"),
                                                    N.ptr(am_), int(attn_mask_mode == "mul"), N.dtype_code(y.dtype),
                                                    N.stream_ptr()), "dsb_bs_softmax_fwd")
            K._count()
        else:
            d = sparse_to_dense(x.float(), layout, block, zero=float("-inf")) * scale
            if rpe is not None:
                d = d + rpe.float()
            if key_padding_mask is not None:
                m = key_padding_mask.float()[:, None, None, :]
                d = d.masked_fill(m == 0, float("-inf")) if kp_mask_mode == "mul" else d + m
            if attn_mask is not None:
                m = attn_mask.float()[None, None, :, :]
                d = d.masked_fill(m == 0, float("-inf")) if attn_mask_mode == "mul" else d + m
            p = torch.softmax(d, -1)
            p = torch.where(torch.isnan(p), torch.zeros_like(p), p)
            y = dense_to_sparse(p, layout, block).to(x.dtype)
        ctx.save_for_backward(y)
        ctx.cfg = (scale, layout, block, lut)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        scale, layout, block, lut = ctx.cfg
        B = y.shape[0]
        H, nb = layout.shape[0], layout.shape[1]
        if y.is_cuda and y.dtype in (torch.float16, torch.bfloat16):
            dx = dy.contiguous().clone()
            _, row_ptr, row_idx, _, _ = lut.on(y.device)
            N.check(N.cuda_lib().dsb_bs_softmax_bwd(N.ptr(dx), N.ptr(y), N.ptr(row_ptr), N.ptr(row_idx), B, H,
                                                    nb * block, block, lut.nnz, float(scale), N.dtype_code(y.dtype),
                                                    N.stream_ptr()), "dsb_bs_softmax_bwd")
            K._count()
        else:
            yd = sparse_to_dense(y.float(), layout, block)
            gd = sparse_to_dense(dy.float(), layout, block)
            dxd = yd * (gd - (yd * gd).sum(-1, keepdim=True)) * scale
            dx = dense_to_sparse(dxd, layout, block).to(y.dtype)
        return dx, None, None, None, None, None, None, None, None, None


class Softmax:
    """Block-Sparse Softmax over a fixed layout."""

    def __init__(self, layout, block, bench=False):
        self.num_blocks = layout.sum().item()
        self.spdims = layout.shape
        self.layout = layout
        self.block = block
        self.bench = bench
        self.lut = get_lut(layout, block)

    def make_lut(self, device=None):
        """The look-up table of this layout (reference softmax.py:241-249); the same native table as MatMul's."""
        return self.lut

    def __call__(self, x, scale=1., rpe=None, key_padding_mask=None, attn_mask=None, key_padding_mask_mode="add",
                 attn_mask_mode="add"):
        if rpe is not None and rpe.dtype != x.dtype:
            raise ValueError("relative position embedding must be %s" % x.dtype)
        if attn_mask is not None and attn_mask.dtype != x.dtype:
            raise ValueError("Attention mask must be %s" % x.dtype)
        if key_padding_mask is not None and key_padding_mask.dtype != x.dtype:
            raise ValueError("Key padding mask must be %s" % x.dtype)
        return _SparseSoftmax.apply(x, scale, rpe, key_padding_mask, attn_mask, key_padding_mask_mode, attn_mask_mode,
                                    self.layout, self.block, self.lut)
