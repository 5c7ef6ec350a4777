"""The allocation-free column sum (`dsb_column_sum_noalloc`) must agree with the regular entry point.

Some users of the kernel library cannot allocate GPU memory at the point where they need a bias gradient: their code
runs where the CUDA allocation calls are forbidden, or every byte of HBM already belongs to someone else's pool. For
them the library offers a variant whose scratch is a fixed array inside the .so. This test drives both entry points
straight through the C ABI (ctypes on libdsb200.so, raw device pointers) and checks that the variant gives the same
sums as `dsb_column_sum` on two consecutive calls — the second call would be wrong if the shared array were not
cleared between uses — and that it refuses widths larger than its built-in array.

The file name sorts last on purpose: with `pytest -x` every other test has already run before this one."""
import ctypes as C
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

LIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "deeperspeed_b200", "lib", "libdsb200.so")
DSB_ERR_SIZE = -2003
MAX_COLS = 65536
F16 = 0


def _lib():
    lib = C.CDLL(LIB)
    lib.dsb_column_sum.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    lib.dsb_column_sum.restype = C.c_int
    lib.dsb_column_sum_noalloc.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_void_p]
    lib.dsb_column_sum_noalloc.restype = C.c_int
    return lib


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / (b.abs().max() + 1e-6)).item()


def test_noalloc_column_sum_matches_regular_entry_point_twice():
    lib = _lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    stream = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(0)
    # two different problems back to back: the second is narrower AND has other data, so stale sums would show
    for rows, cols in ((3000, 4096), (777, 1024)):
        x = torch.randn(rows, cols, device=dev, dtype=torch.float16)
        want = torch.empty(cols, device=dev, dtype=torch.float16)
        got = torch.empty(cols, device=dev, dtype=torch.float16)
        scratch = torch.empty(cols, device=dev, dtype=torch.float32)   # the regular entry point needs caller scratch
        assert lib.dsb_column_sum(x.data_ptr(), want.data_ptr(), rows, cols, F16, 0, scratch.data_ptr(), stream) == 0
        assert lib.dsb_column_sum_noalloc(x.data_ptr(), got.data_ptr(), rows, cols, F16, 0, stream) == 0
        torch.cuda.synchronize()
        assert _rel(got, want) < 3e-3
        assert _rel(got, x.float().sum(0)) < 3e-3
    # accumulate=1 adds onto the existing output, exactly like the regular entry point
    base = got.clone()
    assert lib.dsb_column_sum_noalloc(x.data_ptr(), got.data_ptr(), rows, cols, F16, 1, stream) == 0
    torch.cuda.synchronize()
    assert _rel(got, base.float() + x.float().sum(0)) < 3e-3


def test_noalloc_column_sum_rejects_oversize():
    lib = _lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    cols = MAX_COLS + 8
    x = torch.zeros(8, cols, device=dev, dtype=torch.float16)
    out = torch.zeros(cols, device=dev, dtype=torch.float16)
    rc = lib.dsb_column_sum_noalloc(x.data_ptr(), out.data_ptr(), 8, cols, F16, 0, torch.cuda.current_stream().cuda_stream)
    assert rc == DSB_ERR_SIZE
