"""Op builders with the reference's names (deepspeed/ops/op_builder, op_builder/__init__.py:16-26):
CPUAdamBuilder, FusedAdamBuilder, FusedLambBuilder, SparseAttnBuilder, TransformerBuilder,
StochasticTransformerBuilder, UtilsBuilder, AsyncIOBuilder — `.NAME`, `.is_compatible()`, `.load()`.

The reference JIT-compiles one pybind module per op (and cannot even spell the sm_100 arch string,
op_builder/builder.py:306). Here every op is part of two ahead-of-time libraries built for sm_100a only
(tools/build_native.py); `load()` makes sure they exist (building them if nvcc is available) and returns a small
module-like object exposing the entry points that the reference's Python wrappers used."""
import os
import shutil
import types

import torch

from .. import native as N


def _ensure_built():
    if not (os.path.exists(N.lib_path()) and os.path.exists(N.lib_path("libdsb200_host.so"))):
        import importlib.util
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
        spec = importlib.util.spec_from_file_location("dsb_build_native", os.path.join(root, "tools", "build_native.py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        mod.build(verbose=False)


class OpBuilder:
    NAME = "op"
    CUDA = False

    def __init__(self, name=None):
        self.name = name or self.NAME

    def absolute_name(self):
        return f"deepspeed.ops.{self.NAME}_op"

    def sources(self):
        return []

    def include_paths(self):
        return ["include", "csrc/common"]

    def is_compatible(self):
        if self.CUDA and shutil.which("nvcc") is None and not os.path.exists(N.lib_path()):
            return False
        return True

    def compute_capability_args(self, cross_compile_archs=None):
        """This framework targets exactly one architecture."""
        return ["-gencode", "arch=compute_100a,code=sm_100a"]

    def command_exists(self, cmd):
        return shutil.which(cmd) is not None

    def builder(self):
        return self

    def jit_load(self, verbose=True):
        return self.load(verbose)

    def load(self, verbose=True):
        _ensure_built()
        return self._module()

    def _module(self):
        raise NotImplementedError


class CUDAOpBuilder(OpBuilder):
    CUDA = True


class CPUAdamBuilder(OpBuilder):
    NAME = "cpu_adam"

    def sources(self):
        return ["csrc/host/cpu_adam.cpp"]

    def _module(self):
        from ..adam import cpu_adam
        lib = cpu_adam._lib()
        m = types.SimpleNamespace()
        m.create_adam = lambda opt_id, lr, b1, b2, eps, wd, adamw=True: lib.dsb_cpu_adam_create(opt_id, lr, b1, b2, eps, wd, int(adamw))
        m.destroy_adam = lambda opt_id: lib.dsb_cpu_adam_destroy(opt_id)

        def adam_update(opt_id, step, lr, b1, b2, eps, wd, bias_correction, p, g, m_, v):
            return lib.dsb_cpu_adam_step(opt_id, step, lr, b1, b2, eps, wd, int(bias_correction), p.data_ptr(),
                                         g.data_ptr(), m_.data_ptr(), v.data_ptr(), p.numel(), None, None, 0)

        def adam_update_copy(opt_id, step, lr, b1, b2, eps, wd, bias_correction, p, g, m_, v, gpu_p):
            return lib.dsb_cpu_adam_step(opt_id, step, lr, b1, b2, eps, wd, int(bias_correction), p.data_ptr(),
                                         g.data_ptr(), m_.data_ptr(), v.data_ptr(), p.numel(), gpu_p.data_ptr(), None,
                                         int(gpu_p.dtype == torch.bfloat16))

        m.adam_update, m.adam_update_copy = adam_update, adam_update_copy
        return m


class FusedAdamBuilder(CUDAOpBuilder):
    NAME = "fused_adam"

    def sources(self):
        return ["csrc/optim/optimizers.cu"]

    def _module(self):
        from ..adam.fused_adam import FusedAdam
        return types.SimpleNamespace(FusedAdam=FusedAdam, lib=N.cuda_lib() if torch.cuda.is_available() else None)


class FusedLambBuilder(CUDAOpBuilder):
    NAME = "fused_lamb"

    def sources(self):
        return ["csrc/optim/optimizers.cu"]

    def _module(self):
        from ..lamb.fused_lamb import FusedLamb
        return types.SimpleNamespace(FusedLamb=FusedLamb, lib=N.cuda_lib() if torch.cuda.is_available() else None)


class TransformerBuilder(CUDAOpBuilder):
    NAME = "transformer"

    def sources(self):
        return ["csrc/gemm/gemm_tcgen05.cu", "csrc/transformer/norm_softmax.cu", "csrc/transformer/elementwise.cu"]

    def _module(self):
        from .. import kernels
        return kernels


class StochasticTransformerBuilder(TransformerBuilder):
    NAME = "stochastic_transformer"


class SparseAttnBuilder(OpBuilder):
    NAME = "sparse_attn"

    def sources(self):
        return ["csrc/sparse_attn/lut.cpp", "csrc/sparse_attn/bs_kernels.cu"]

    def _module(self):
        from ..sparse_attention import matmul
        return matmul


class UtilsBuilder(OpBuilder):
    NAME = "utils"

    def sources(self):
        return ["csrc/host/flatten.cpp"]

    def _module(self):
        from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors
        return types.SimpleNamespace(flatten=_flatten_dense_tensors, unflatten=_unflatten_dense_tensors)


class AsyncIOBuilder(OpBuilder):
    NAME = "async_io"

    def sources(self):
        return ["csrc/host/aio.cpp"]

    def _module(self):
        from .. import aio
        return aio


__op_builders__ = [CPUAdamBuilder(), FusedAdamBuilder(), FusedLambBuilder(), SparseAttnBuilder(), TransformerBuilder(),
                   StochasticTransformerBuilder(), UtilsBuilder(), AsyncIOBuilder()]
ALL_OPS = {op.name: op for op in __op_builders__}
