"""FusedAdam: Adam/AdamW with one multi-tensor kernel launch per 24 tensors (sm_100a, fp32 math in registers).
API per reference deepspeed/ops/adam/fused_adam.py:16-182. On CPU tensors it runs the same math in PyTorch so the
engine is testable without a GPU. The ZeRO / FP16 wrappers bypass the per-tensor path and call `step_flat` on
arena slices (one launch for a whole partition)."""
import ctypes as C

import torch

from .. import kernels as K
from .. import native as N


class FusedAdam(torch.optim.Optimizer):
    supports_flat_fused = True

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0.0, amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError("FusedAdam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group["params"]:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def step_flat(self, group, p32, m, v, grad, lp=None, inv_scale=1.0, inv_scale_dev=None, skip_dev=None):
        """Fused update of one flat fp32 partition (used by the arena-based mixed precision wrappers)."""
        group["step"] = group.get("step", 0) + 1
        beta1, beta2 = group["betas"]
        K.adam_flat(p32, m, v, grad, lp=lp, lr=group["lr"], beta1=beta1, beta2=beta2, eps=group["eps"],
                    weight_decay=group["weight_decay"], adamw=bool(self.adam_w_mode),
                    bias_correction=bool(group["bias_correction"]), step=group["step"], inv_scale=inv_scale,
                    inv_scale_dev=inv_scale_dev, skip_dev=skip_dev)

    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None):
        if any(a is not None for a in (grads, output_params, scale, grad_norms)):
            raise RuntimeError("FusedAdam has been updated. Simply initialize it identically to torch.optim.Adam, "
                               "and call step() with no arguments.")
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            group["step"] = group.get("step", 0) + 1
            buckets = {}
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.data.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    state["exp_avg"] = torch.zeros_like(p.data)
                    state["exp_avg_sq"] = torch.zeros_like(p.data)
                buckets.setdefault((p.dtype, p.grad.dtype, p.is_cuda), []).append(p)
            for (pdt, gdt, cuda), plist in buckets.items():
                if cuda:
                    self._launch(plist, group, beta1, beta2)
                else:
                    for p in plist:
                        st = self.state[p]
                        K.adam_flat(p.data.view(-1), st["exp_avg"].view(-1), st["exp_avg_sq"].view(-1),
                                    p.grad.data.view(-1), lr=group["lr"], beta1=beta1, beta2=beta2, eps=group["eps"],
                                    weight_decay=group["weight_decay"], adamw=bool(self.adam_w_mode),
                                    bias_correction=bool(group["bias_correction"]), step=group["step"]) \
                            if p.dtype == torch.float32 else self._cpu_lowp(p, st, group, beta1, beta2)
        return loss

    def _cpu_lowp(self, p, st, group, beta1, beta2):
        p32, m32, v32 = p.data.float().view(-1), st["exp_avg"].float().view(-1), st["exp_avg_sq"].float().view(-1)
        K.adam_flat(p32, m32, v32, p.grad.data.view(-1), lr=group["lr"], beta1=beta1, beta2=beta2, eps=group["eps"],
                    weight_decay=group["weight_decay"], adamw=bool(self.adam_w_mode),
                    bias_correction=bool(group["bias_correction"]), step=group["step"])
        p.data.copy_(p32.view_as(p.data))
        st["exp_avg"].copy_(m32.view_as(p.data))
        st["exp_avg_sq"].copy_(v32.view_as(p.data))

    def _launch(self, plist, group, beta1, beta2):
        n = len(plist)
        arr = C.c_void_p * n
        ps = arr(*[p.data.data_ptr() for p in plist])
        gs = arr(*[(p.grad.data if p.grad.data.is_contiguous() else p.grad.data.contiguous()).data_ptr() for p in plist])
        ms = arr(*[self.state[p]["exp_avg"].data_ptr() for p in plist])
        vs = arr(*[self.state[p]["exp_avg_sq"].data_ptr() for p in plist])
        ns = (C.c_long * n)(*[p.numel() for p in plist])
        rc = N.cuda_lib().dsb_adam_multi(n, C.cast(ps, C.c_void_p), C.cast(gs, C.c_void_p), C.cast(ms, C.c_void_p),
                                         C.cast(vs, C.c_void_p), C.cast(ns, C.c_void_p),
                                         N.dtype_code(plist[0].dtype), N.dtype_code(plist[0].grad.dtype),
                                         N.dtype_code(self.state[plist[0]]["exp_avg"].dtype), float(group["lr"]),
                                         float(beta1), float(beta2), float(group["eps"]), float(group["weight_decay"]),
                                         int(self.adam_w_mode), int(group["bias_correction"]), int(group["step"]), 1.0,
                                         N.stream_ptr())
        N.check(rc, "dsb_adam_multi")
        K._count((n + 23) // 24)
