"""Single-GPU harness for the fused ZeRO step kernels (R = 1: the only "peer" is this GPU). The cross-GPU barriers,
the peer-pointer tables and the LAMB norm exchange run exactly as in the multi-GPU case, which lets one GPU check the
kernels' numerics (tests/test_kernels_gpu.py) and lets `ncu` profile them (tools/ncu_fused_zero.py; ncu must not be run
on multi-rank commands)."""
import ctypes as C

import torch

from ..ops import native as N
from . import fused_zero  # noqa: F401  (registers the ctypes signatures)


class SingleRankFusedZero:
    def __init__(self, n, dtype=torch.float16, tensor_sizes=None, device="cuda"):
        assert n % 8 == 0
        self.n, self.dtype, self.dev = n, dtype, torch.device(device)
        self.lib = N.cuda_lib()
        self.flags = torch.zeros(1024, dtype=torch.int32, device=self.dev)
        self.grad = torch.zeros(n, dtype=dtype, device=self.dev)
        self.param = torch.zeros(n, dtype=dtype, device=self.dev)
        self.red = torch.empty(n, dtype=torch.float32, device=self.dev)
        self.master = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.m = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.v = torch.zeros(n, dtype=torch.float32, device=self.dev)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=self.dev)
        self.flag = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.inv = torch.ones(1, dtype=torch.float32, device=self.dev)
        self.skip = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.scratch = torch.zeros(self.lib.dsb_fused_zero_scratch_bytes(), dtype=torch.uint8, device=self.dev)
        self.epoch = 0
        self._arr = lambda t: (C.c_void_p * 1)(t.data_ptr())
        # LAMB: tensors laid back to back (sizes are multiples of 8), last id = padding
        sizes = list(tensor_sizes or [n])
        assert sum(sizes) <= n and all(s % 8 == 0 for s in sizes)
        starts, tids, pos = [], [], 0
        for i, s in enumerate(sizes):
            starts.append(pos)
            tids.append(i)
            pos += s
        if pos < n:
            starts.append(pos)
            tids.append(len(sizes))
        starts.append(n)
        self.sizes = sizes
        self.nt = len(sizes) + 1
        self.seg_start = torch.tensor(starts, dtype=torch.int64, device=self.dev)
        self.seg_tid = torch.tensor(tids, dtype=torch.int32, device=self.dev)
        self.norms = torch.zeros(2 * self.nt, dtype=torch.float32, device=self.dev)
        self.coeff = torch.ones(self.nt, dtype=torch.float32, device=self.dev)

    def _next(self):
        self.epoch += 1
        return self.epoch

    def rs_norm(self, scale=1.0):
        self.sumsq.zero_()
        self.flag.zero_()
        N.check(self.lib.dsb_fused_rs_norm(self._arr(self.grad), self._arr(self.flags), 1, 0, 0, self.n,
                                           N.dtype_code(self.dtype), float(scale), N.ptr(self.red), N.ptr(self.sumsq),
                                           N.ptr(self.flag), N.ptr(self.scratch), self._next(), N.stream_ptr()),
                "dsb_fused_rs_norm")

    def adam_ag(self, lr, betas, eps, wd, adamw, step):
        N.check(self.lib.dsb_fused_adam_ag(N.ptr(self.red), N.ptr(self.master), N.ptr(self.m), N.ptr(self.v),
                                           self._arr(self.param), self._arr(self.flags), 1, 0, 0, self.n,
                                           N.dtype_code(self.dtype), lr, betas[0], betas[1], eps, wd, int(adamw), 1,
                                           step, N.ptr(self.inv), N.ptr(self.skip), self._next(), N.stream_ptr()),
                "dsb_fused_adam_ag")

    def lamb_ag(self, lr, betas, eps, wd, max_coeff, min_coeff, step):
        N.check(self.lib.dsb_fused_lamb_ag(N.ptr(self.red), N.ptr(self.master), N.ptr(self.m), N.ptr(self.v),
                                           self._arr(self.param), self._arr(self.norms), self._arr(self.flags), 1, 0,
                                           self.n, self.n, N.ptr(self.seg_start), N.ptr(self.seg_tid),
                                           self.seg_tid.numel(), self.nt, N.dtype_code(self.dtype), lr, betas[0],
                                           betas[1], eps, wd, max_coeff, min_coeff, 0, 1, step, N.ptr(self.inv),
                                           N.ptr(self.skip), N.ptr(self.coeff), self._next(), N.stream_ptr()),
                "dsb_fused_lamb_ag")

    def whole_step(self, lr, betas, eps, wd, adamw, step, loss_scale=1.0, clip=0.0, phases=7):
        """The one-kernel step (csrc/comm/fused_zero_step.cu) with this GPU as its only peer: reduce (here: average of
        one copy) + norm, in-kernel scalar exchange -> unscale / clip factor + skip flag, Adam + cast + gather + clearing
        of the gradient bucket. Returns the device tensor [norm, overflow] the kernel wrote for the host."""
        if not hasattr(self, "sumsq2"):
            self.sumsq2 = torch.zeros(2, dtype=torch.float32, device=self.dev)
            self.host_out = torch.zeros(2, dtype=torch.float32, device=self.dev)
            self.red16 = torch.empty(self.n, dtype=self.dtype, device=self.dev)
        self.sumsq2.zero_()
        N.check(self.lib.dsb_fused_zero_step(
            self._arr(self.grad), self._arr(self.param), self._arr(self.flags), None, None, 1, 0, int(phases), 1, 0, 0,
            self.n, self.n, N.dtype_code(self.dtype), N.ptr(self.red16), N.ptr(self.master), N.ptr(self.m),
            N.ptr(self.v), float(lr), float(betas[0]), float(betas[1]), float(eps), float(wd), int(adamw), 1, int(step),
            float(loss_scale), float(clip), N.ptr(self.sumsq2), N.ptr(self.inv), N.ptr(self.skip), N.ptr(self.host_out),
            N.ptr(self.scratch), self._next(), N.stream_ptr()), "dsb_fused_zero_step")
        return self.host_out

