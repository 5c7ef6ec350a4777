"""GEMM kernel-variant autotuner (reference: the cuBLAS algo search `GemmTest` / `StridedGemmTest`,
csrc/includes/gemm_test.h:27-293, run once per transformer-layer config).

The timing loop and the result table live in the native library (csrc/gemm/gemm_autotune.cu); this module only
enumerates the problems of a model: for every 2-D weight [out, in] and `tokens` rows per micro-batch the forward
(tokens x out x in), dgrad (tokens x in x out, B MN-major) and wgrad (out x in x tokens, both MN-major) GEMMs.
"""
import ctypes as C

import torch

from . import native as N

VARIANTS = ("single", "pair/whole", "pair/split2", "pair/split3", "pair/split4", "pair/auto-split")


def tune(M, Nn, K, *, batch=1, trans_a=False, trans_b=True, dtype=torch.float16, iters=7):
    """Time every variant for one problem; returns (best_variant_name, [ms per variant])."""
    lib = N.cuda_lib()
    ms = (C.c_float * len(VARIANTS))()
    rc = lib.dsb_gemm_tune(M, Nn, K, batch, int(trans_a), int(not trans_b), N.dtype_code(dtype), iters, ms,
                           N.stream_ptr())
    if rc < 0:
        raise RuntimeError(f"dsb_gemm_tune failed with code {rc}")
    return VARIANTS[rc], [float(x) for x in ms]


def tuned_variant(M, Nn, K, *, batch=1, trans_a=False, trans_b=True, dtype=torch.float16):
    rc = N.cuda_lib().dsb_gemm_tuned_variant(M, Nn, K, batch, int(trans_a), int(not trans_b), N.dtype_code(dtype))
    return None if rc < 0 else VARIANTS[rc]


def linear_problems(tokens, in_features, out_features):
    """(M, N, K, trans_a, trans_b) of the three GEMMs of y = x W^T."""
    return [(tokens, out_features, in_features, False, True),     # forward
            (tokens, in_features, out_features, False, False),    # dgrad: dy [tokens, out] @ W [out, in]
            (out_features, in_features, tokens, True, False)]     # wgrad: dy^T @ x


def tune_module(module, tokens, dtype=torch.float16, iters=7, verbose=False):
    """Tune the linear-layer GEMMs of every 2-D parameter of `module` for `tokens` rows per micro-batch."""
    if not torch.cuda.is_available():
        return {}
    seen, report = set(), {}
    for _, p in module.named_parameters():
        if p.dim() != 2 or min(p.shape) < 128:
            continue
        out_f, in_f = p.shape
        for prob in linear_problems(tokens, in_f, out_f):
            if prob in seen:
                continue
            seen.add(prob)
            M, Nn, K, ta, tb = prob
            best, ms = tune(M, Nn, K, trans_a=ta, trans_b=tb, dtype=dtype, iters=iters)
            report[prob] = (best, ms)
            if verbose:
                print(f"[gemm_autotune] M={M} N={Nn} K={K} ta={ta} tb={tb}: {best} {[round(x, 4) for x in ms]}",
                      flush=True)
    return report
