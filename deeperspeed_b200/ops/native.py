"""ctypes bindings to the in-tree native libraries (C ABI in include/dsb200.h).

The CUDA library is mandatory on a GPU box: if `torch.cuda.is_available()` and libdsb200.so cannot be loaded the
ops raise instead of silently falling back (the reference JIT-loads one pybind module per op,
reference op_builder/builder.py:217-256; here everything is built ahead of time by tools/build_native.py).
"""
import ctypes as C
import os
import threading

import torch

_LIB_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "lib")
_lock = threading.Lock()
_cuda_lib = None
_host_lib = None

DTYPE_F16, DTYPE_BF16, DTYPE_F32 = 0, 1, 2
EPI_NONE, EPI_BIAS, EPI_BIAS_GELU, EPI_BIAS_RESIDUAL, EPI_DGELU = 0, 1, 2, 3, 4

_p, _i, _l, _f, _u64 = C.c_void_p, C.c_int, C.c_long, C.c_float, C.c_ulonglong

_CUDA_SIGS = {
    "dsb_gemm": [_p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _l, _l, _l, _i, _i, _i, _i, _f, _p, _p, _l, _l, _p, _l, _l, _p],
    "dsb_gemm_pair": [_p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _l, _l, _l, _i, _i, _i, _i, _f, _p, _p, _l, _l, _p, _l, _l, _p],
    "dsb_gemm_auto": [_p, _p, _p, _i, _i, _i, _i, _l, _l, _l, _l, _l, _l, _i, _i, _i, _i, _f, _p, _p, _l, _l, _p, _l, _l, _p],
    "dsb_gemm_tune": [_i, _i, _i, _i, _i, _i, _i, _i, _p, _p],
    "dsb_gemm_tuned_variant": [_i, _i, _i, _i, _i, _i, _i],
    "dsb_gemm_pair_force_bn": [_i],
    "dsb_gemm_pair_force_split": [_i],
    "dsb_adam_flat": [_p, _p, _p, _p, _i, _p, _i, _l, _f, _f, _f, _f, _f, _i, _i, _i, _f, _p, _p, _p],
    "dsb_adam_multi": [_i, _p, _p, _p, _p, _p, _i, _i, _i, _f, _f, _f, _f, _f, _i, _i, _i, _f, _p],
    "dsb_lamb_step": [_p, _i, _p, _i, _p, _i, _p, _p, _l, _f, _f, _f, _f, _f, _f, _f, _i, _i, _i, _f, _p, _p, _p],
    "dsb_cast": [_p, _i, _p, _i, _l, _p],
    "dsb_sumsq": [_p, _i, _l, _p, _p, _i, _p, _p],
    "dsb_scale": [_p, _i, _l, _f, _p, _p],
    "dsb_layernorm_fwd": [_p, _p, _p, _p, _p, _p, _p, _p, _l, _i, _f, _i, _p],
    "dsb_layernorm_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _l, _i, _i, _p, _p],
    "dsb_layernorm_bwd_acc": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _l, _i, _i, _i, _p, _p],
    "dsb_layernorm_bwd_inv": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _l, _i, _i, _i, _p, _p],
    "dsb_softmax_fwd": [_p, _p, _p, _i, _i, _i, _i, _f, _i, _i, _p],
    "dsb_softmax_bwd": [_p, _p, _p, _l, _i, _f, _i, _p],
    "dsb_bias_gelu_fwd": [_p, _p, _p, _l, _i, _i, _p],
    "dsb_bias_gelu_bwd": [_p, _p, _p, _p, _l, _i, _i, _p],
    "dsb_dropout_fwd": [_p, _p, _p, _p, _p, _l, _i, _f, _u64, _u64, _i, _p],
    "dsb_dropout_bwd": [_p, _p, _p, _l, _f, _i, _p],
    "dsb_qkv_bias_split_transpose": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p],
    "dsb_transpose_0213": [_p, _p, _i, _i, _i, _i, _i, _p],
    "dsb_column_sum": [_p, _p, _l, _i, _i, _i, _p, _p],
    "dsb_embedding_fwd": [_p, _p, _p, _l, _i, _l, _i, _p],
    "dsb_embedding_bwd_acc": [_p, _p, _p, _l, _i, _l, _i, _p],
    "dsb_fused_add": [_p, _p, _p, _p, _p, _l, _i, _p],
    "dsb_cross_entropy_fwd": [_p, _p, _p, _p, _l, _i, _l, _l, _i, _p],
    "dsb_cross_entropy_bwd": [_p, _p, _p, _p, _f, _l, _i, _l, _l, _i, _p],
    "dsb_rotary": [_p, _p, _i, _i, _i, _i, _i, _l, _l, _l, _f, _i, _i, _p],
}
_CUDA_LONG_RET = {"dsb_lamb_scratch_bytes": [_l], "dsb_sumsq_scratch_bytes": [], "dsb_layernorm_bwd_scratch_bytes": [_i]}


class NativeUnavailable(RuntimeError):
    pass


def lib_path(name="libdsb200.so"):
    return os.path.join(_LIB_DIR, name)


def _bind(lib, sigs, restype):
    for name, args in sigs.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.argtypes = args
        fn.restype = restype


def cuda_lib():
    """Load libdsb200.so (once). Raises NativeUnavailable if missing."""
    global _cuda_lib
    if _cuda_lib is None:
        with _lock:
            if _cuda_lib is None:
                path = lib_path()
                if not os.path.exists(path):
                    raise NativeUnavailable(
                        f"{path} not found: run `python tools/build_native.py` (or __graft_entry__.build())")
                lib = C.CDLL(path)
                _bind(lib, _CUDA_SIGS, _i)
                _bind(lib, _CUDA_LONG_RET, _l)
                _bind_late(lib)
                _cuda_lib = lib
    return _cuda_lib


_LATE_SIGS = {}


def register_sigs(sigs, restype=_i):
    """Other modules (comm, sparse attention, ...) register their entry points here before first use."""
    for k, v in sigs.items():
        _LATE_SIGS[k] = (v, restype)
    if _cuda_lib is not None:
        _bind_late(_cuda_lib)


def _bind_late(lib):
    for name, (args, restype) in _LATE_SIGS.items():
        fn = getattr(lib, name, None)
        if fn is not None:
            fn.argtypes = args
            fn.restype = restype


def host_lib():
    global _host_lib
    if _host_lib is None:
        with _lock:
            if _host_lib is None:
                path = lib_path("libdsb200_host.so")
                if not os.path.exists(path):
                    raise NativeUnavailable(f"{path} not found: run `python tools/build_native.py`")
                _host_lib = C.CDLL(path)
    return _host_lib


def available():
    return os.path.exists(lib_path())


def use_native(t=None):
    """True when the CUDA kernels must be used (CUDA tensor / CUDA box). Fails loudly if the lib is missing."""
    on_cuda = t.is_cuda if t is not None else torch.cuda.is_available()
    if not on_cuda:
        return False
    cuda_lib()
    return True


def dtype_code(dt):
    if dt == torch.float16:
        return DTYPE_F16
    if dt == torch.bfloat16:
        return DTYPE_BF16
    if dt == torch.float32:
        return DTYPE_F32
    raise TypeError(f"unsupported dtype {dt}")


def ptr(t):
    return 0 if t is None else t.data_ptr()


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def check(rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed with code {rc}")


_scratch = {}


def scratch(nbytes, device, tag="default"):
    """Per-(device, tag) grow-only scratch buffer; callers on one stream reuse it in stream order."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), tag)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 16), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf
