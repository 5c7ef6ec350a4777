"""Block-sparse matrix multiplication `MatMul(layout, block, mode, trans_a, trans_b)` with mode in
{'sdd', 'dsd', 'dds'} (API of reference deepspeed/ops/sparse_attention/matmul.py:615-749).

  sdd: sparse = dense x dense     dsd: dense = sparse x dense     dds: dense = dense x sparse

Sparse tensors are stored [batch, nnz, block, block] in layout order. On CUDA fp16/bf16 the products run on the
sm_100a tensor-core kernels (csrc/sparse_attn/bs_kernels.cu) driven by LUTs built natively (csrc/sparse_attn/lut.cpp);
other dtypes / CPU go through a dense-emulation path, which is also the test oracle."""
import ctypes as C

import torch

from .. import kernels as K
from .. import native as N

_p, _i, _l, _f = C.c_void_p, C.c_int, C.c_long, C.c_float
N.register_sigs({
    "dsb_bs_sdd": [_p, _p, _p, _p, _i, _i, _i, _i, _i, _l, _f, _i, _p],
    "dsb_bs_dsd": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _l, _i, _i, _p],
    "dsb_bs_softmax_fwd": [_p, _p, _p, _i, _i, _i, _i, _l, _f, _p, _p, _i, _p, _i, _i, _p],
    "dsb_bs_softmax_bwd": [_p, _p, _p, _p, _i, _i, _i, _i, _l, _f, _i, _p],
})


class BlockSparseLUT:
    """Native LUTs for one layout (cached per device)."""
    _host_bound = False

    def __init__(self, layout, block):
        self.layout = layout.to(torch.int64).cpu().contiguous()
        self.block = block
        self.H, self.nb = self.layout.shape[0], self.layout.shape[1]
        lib = N.host_lib()
        if not BlockSparseLUT._host_bound:
            lib.dsb_bs_count_nnz.restype = C.c_long
            lib.dsb_bs_count_nnz.argtypes = [_p, _i, _i]
            lib.dsb_bs_build_luts.argtypes = [_p, _i, _i, _p, _p, _p, _p, _p]
            lib.dsb_sdd_segment.argtypes = [_p, _i, _i, _i, _p, _p, _p]
            BlockSparseLUT._host_bound = True
        self.nnz = int(lib.dsb_bs_count_nnz(self.layout.data_ptr(), self.H, self.nb))
        self.blocks = torch.zeros(max(self.nnz, 1), 4, dtype=torch.int32)
        self.row_ptr = torch.zeros(self.H * self.nb + 1, dtype=torch.int32)
        self.row_idx = torch.zeros(max(self.nnz, 1), 2, dtype=torch.int32)
        self.col_ptr = torch.zeros(self.H * self.nb + 1, dtype=torch.int32)
        self.col_idx = torch.zeros(max(self.nnz, 1), 2, dtype=torch.int32)
        lib.dsb_bs_build_luts(self.layout.data_ptr(), self.H, self.nb, self.blocks.data_ptr(), self.row_ptr.data_ptr(),
                              self.row_idx.data_ptr(), self.col_ptr.data_ptr(), self.col_idx.data_ptr())
        self._dev = {}

    def on(self, device):
        key = str(device)
        if key not in self._dev:
            self._dev[key] = tuple(t.to(device) for t in (self.blocks, self.row_ptr, self.row_idx, self.col_ptr,
                                                           self.col_idx))
        return self._dev[key]

    def sdd_segment(self, max_width):
        """Greedy super-block packing (kept for parity with the reference's C++ `sdd_segment`)."""
        lib = N.host_lib()
        ids = torch.empty(self.H, self.nb, self.nb, dtype=torch.int32)
        widths = torch.empty_like(ids)
        levels = max_width.bit_length()
        counts = torch.zeros(levels, dtype=torch.int64)
        lib.dsb_sdd_segment(self.layout.data_ptr(), self.H, self.nb, max_width, ids.data_ptr(), widths.data_ptr(),
                            counts.data_ptr())
        return ids, widths, counts


_lut_cache = {}


def get_lut(layout, block):
    key = (layout.shape, block, layout.to(torch.int8).cpu().numpy().tobytes())
    if key not in _lut_cache:
        _lut_cache[key] = BlockSparseLUT(layout, block)
    return _lut_cache[key]


# ------------------------------------------------------------------------------------------------ dense emulation
def dense_to_sparse(w, layout, block):
    """[B, H, M, N] dense -> [B, nnz, block, block] in layout order."""
    B = w.shape[0]
    idx = layout.nonzero(as_tuple=False)
    out = torch.empty(B, idx.shape[0], block, block, dtype=w.dtype, device=w.device)
    for z, (h, i, j) in enumerate(idx.tolist()):
        out[:, z] = w[:, h, i * block:(i + 1) * block, j * block:(j + 1) * block]
    return out


def sparse_to_dense(w, layout, block, zero=0.0):
    B = w.shape[0]
    H, nb_r, nb_c = layout.shape
    out = torch.full((B, H, nb_r * block, nb_c * block), zero, dtype=w.dtype, device=w.device)
    for z, (h, i, j) in enumerate(layout.nonzero(as_tuple=False).tolist()):
        out[:, h, i * block:(i + 1) * block, j * block:(j + 1) * block] = w[:, z]
    return out


def _native_ok(*ts):
    return all(t.is_cuda and t.dtype in (torch.float16, torch.bfloat16) for t in ts)


def _sdd_native(a, b, lut, alpha=1.0):
    Bz, H, S, D = a.shape
    blocks = lut.on(a.device)[0]
    c = torch.empty(Bz, lut.nnz, lut.block, lut.block, dtype=a.dtype, device=a.device)
    N.check(N.cuda_lib().dsb_bs_sdd(N.ptr(a), N.ptr(b), N.ptr(c), N.ptr(blocks), Bz, H, S, D, lut.block, lut.nnz,
                                    float(alpha), N.dtype_code(a.dtype), N.stream_ptr()), "dsb_bs_sdd")
    K._count()
    return c


def _dsd_native(p, x, lut, transposed):
    Bz, H, S, D = x.shape
    _, row_ptr, row_idx, col_ptr, col_idx = lut.on(x.device)
    ptr, idx = (col_ptr, col_idx) if transposed else (row_ptr, row_idx)
    o = torch.empty_like(x)
    N.check(N.cuda_lib().dsb_bs_dsd(N.ptr(p), N.ptr(x), N.ptr(o), N.ptr(ptr), N.ptr(idx), Bz, H, S, D, lut.block,
                                    lut.nnz, int(transposed), N.dtype_code(x.dtype), N.stream_ptr()), "dsb_bs_dsd")
    K._count()
    return o


class _SparseMatMul(torch.autograd.Function):
    """All three modes reduce to two native primitives: sdd(A, B) = A B^T on the layout, dsd(P, X, transposed)."""

    @staticmethod
    def forward(ctx, a, b, mode, trans_a, trans_b, layout, block, lut):
        ctx.save_for_backward(a, b)
        ctx.cfg = (mode, trans_a, trans_b, layout, block, lut)
        return _SparseMatMul._run(a, b, mode, trans_a, trans_b, layout, block, lut)

    @staticmethod
    def _run(a, b, mode, trans_a, trans_b, layout, block, lut):
        if mode == "sdd":
            A = a.transpose(-1, -2).contiguous() if trans_a else a.contiguous()
            Bt = b.contiguous() if trans_b else b.transpose(-1, -2).contiguous()   # rows = output columns
            if _native_ok(A, Bt) and A.shape[-1] % 16 == 0 and block % 16 == 0 and block <= 64 and A.shape[-2] == Bt.shape[-2]:
                return _sdd_native(A, Bt, lut)
            return dense_to_sparse(torch.matmul(A.float(), Bt.float().transpose(-1, -2)).to(a.dtype), layout, block)
        if mode == "dsd":
            X = b.transpose(-1, -2).contiguous() if trans_b else b.contiguous()
            if _native_ok(a, X) and block % 16 == 0 and block <= 64 and X.shape[-1] % 8 == 0 and layout.shape[1] == layout.shape[2]:
                return _dsd_native(a.contiguous(), X, lut, transposed=trans_a)
            P = sparse_to_dense(a, layout, block).float()
            if trans_a:
                P = P.transpose(-1, -2)
            return torch.matmul(P, X.float()).to(b.dtype)
        if mode == "dds":
            # C = op(A) op(B_sparse)  =  (op(B_sparse)^T op(A)^T)^T
            A = a.transpose(-1, -2).contiguous() if trans_a else a.contiguous()
            At = A.transpose(-1, -2).contiguous()
            if _native_ok(At, b) and block % 16 == 0 and block <= 64 and At.shape[-1] % 8 == 0 and layout.shape[1] == layout.shape[2]:
                return _dsd_native(b.contiguous(), At, lut, transposed=not trans_b).transpose(-1, -2).contiguous()
            P = sparse_to_dense(b, layout, block).float()
            if trans_b:
                P = P.transpose(-1, -2)
            return torch.matmul(A.float(), P).to(a.dtype)
        raise ValueError(mode)

    @staticmethod
    def backward(ctx, dc):
        a, b = ctx.saved_tensors
        mode, trans_a, trans_b, layout, block, lut = ctx.cfg
        run = _SparseMatMul._run
        da = db = None
        dc = dc.contiguous()
        if mode == "sdd":
            # C = op(A) op(B);  dA' = dC op(B)^T (dsd),  dB' = op(A)^T dC (dsd with transposed sparse)
            if ctx.needs_input_grad[0]:
                da = run(dc, b, "dsd", False, not trans_b, layout, block, lut)
                if trans_a:
                    da = da.transpose(-1, -2)
            if ctx.needs_input_grad[1]:
                A = a.transpose(-1, -2) if trans_a else a
                db = run(dc, A, "dsd", True, False, layout, block, lut)      # = dC^T op(A)
                if not trans_b:
                    db = db.transpose(-1, -2)
        elif mode == "dsd":
            # C = op(P) op(X)
            if ctx.needs_input_grad[0]:
                X = b.transpose(-1, -2) if trans_b else b
                if trans_a:
                    da = run(X, dc, "sdd", False, True, layout, block, lut)   # dP = op(X) dC^T
                else:
                    da = run(dc, X, "sdd", False, True, layout, block, lut)   # dP = dC op(X)^T
            if ctx.needs_input_grad[1]:
                db = run(a, dc, "dsd", not trans_a, False, layout, block, lut)
                if trans_b:
                    db = db.transpose(-1, -2)
        else:  # dds: C = op(A) op(P)
            if ctx.needs_input_grad[0]:
                da = run(dc, b, "dds", False, not trans_b, layout, block, lut)
                if trans_a:
                    da = da.transpose(-1, -2)
            if ctx.needs_input_grad[1]:
                A = a.transpose(-1, -2) if trans_a else a
                At = A.transpose(-1, -2)
                if trans_b:
                    db = run(dc.transpose(-1, -2), At, "sdd", False, True, layout, block, lut)   # dP^T ... = dC^T A
                else:
                    db = run(At, dc.transpose(-1, -2), "sdd", False, True, layout, block, lut)   # dP = A^T dC
        return da, db, None, None, None, None, None, None


class MatMul:
    """Block-Sparse MatMul: C = op(A) op(B) with one sparse operand / result (mode 'sdd' | 'dsd' | 'dds')."""

    def __init__(self, layout, block, mode, trans_a=False, trans_b=False, bench=False):
        if mode not in ["sdd", "dsd", "dds"]:
            raise NotImplementedError("Supported modes are: sdd, dsd, dds")
        self.layout = layout
        self.block = block
        self.mode = mode
        self.trans_a = trans_a
        self.trans_b = trans_b
        self.spdims = layout.shape
        self.bench = bench
        self.lut = get_lut(layout, block)

    def make_lut(self, dtype=None, device=None):
        """The look-up tables of this layout (reference matmul.py:623-660 builds Triton-specific tables per dtype and
        device; ours are built once, natively, and are dtype-independent). Returns the cached BlockSparseLUT."""
        return self.lut

    def __call__(self, a, b):
        a, b = self._pad(a), self._pad(b)
        return _SparseMatMul.apply(a, b, self.mode, self.trans_a, self.trans_b, self.layout, self.block, self.lut)

    @staticmethod
    def _pad(x):
        while x.dim() < 4:
            x = x.unsqueeze(0)
        return x
