"""DeepSpeedCPUAdam: host SIMD Adam/AdamW (AVX-512/AVX2/scalar picked at run time, OpenMP) with an optional fused
fp16/bf16 copy-back to the GPU through a pinned double-buffered pipeline.
API per reference deepspeed/ops/adam/cpu_adam.py:13-152; native code in csrc/host/cpu_adam.cpp."""
import ctypes as C

import torch

from .. import native as N

_f, _i, _l, _p = C.c_float, C.c_int, C.c_long, C.c_void_p


def _lib():
    lib = N.host_lib()
    if not getattr(lib, "_cpu_adam_bound", False):
        lib.dsb_cpu_adam_create.argtypes = [_i, _f, _f, _f, _f, _f, _i]
        lib.dsb_cpu_adam_destroy.argtypes = [_i]
        lib.dsb_cpu_adam_step.argtypes = [_i, _l, _f, _f, _f, _f, _f, _i, _p, _p, _p, _p, _l, _p, _p, _i]
        lib.dsb_cpu_cast_lp.argtypes = [_p, _p, _l, _i]
        lib._cpu_adam_bound = True
    return lib


class DeepSpeedCPUAdam(torch.optim.Optimizer):
    optimizer_id = 0

    def __init__(self, model_params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0,
                 amsgrad=False, adamw_mode=True):
        default_args = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, bias_correction=bias_correction,
                            amsgrad=amsgrad)
        super().__init__(model_params, default_args)
        self.opt_id = DeepSpeedCPUAdam.optimizer_id
        DeepSpeedCPUAdam.optimizer_id += 1
        self.adam_w_mode = adamw_mode
        self.ds_opt_adam = _lib()
        self.ds_opt_adam.dsb_cpu_adam_create(self.opt_id, lr, betas[0], betas[1], eps, weight_decay, int(adamw_mode))

    def __del__(self):
        try:
            self.ds_opt_adam.dsb_cpu_adam_destroy(self.opt_id)
        except Exception:
            pass

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault("amsgrad", False)

    @torch.no_grad()
    def step(self, closure=None, fp16_param_groups=None):
        """fp16_param_groups: optional list (per group) of lists of low-precision GPU tensors that receive the
        updated parameters (fused copy-back)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group_id, group in enumerate(self.param_groups):
            for param_id, p in enumerate(group["params"]):
                if p.grad is None:
                    continue
                assert p.device.type == "cpu", "DeepSpeedCPUAdam expects parameters on the CPU"
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32, device="cpu")
                    state["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32, device="cpu")
                state["step"] += 1
                beta1, beta2 = group["betas"]
                dev_lp = host_lp = None
                lp_bf16 = 0
                if fp16_param_groups is not None:
                    tgt = fp16_param_groups[group_id][param_id]
                    lp_bf16 = int(tgt.dtype == torch.bfloat16)
                    if tgt.is_cuda:
                        dev_lp = tgt.data
                    else:
                        host_lp = tgt.data
                pd = p.data if p.data.dtype == torch.float32 else None
                assert pd is not None and pd.is_contiguous(), "CPU Adam works on contiguous fp32 masters"
                g = p.grad.data
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.float().contiguous()
                rc = self.ds_opt_adam.dsb_cpu_adam_step(
                    self.opt_id, state["step"], group["lr"], beta1, beta2, group["eps"], group["weight_decay"],
                    int(group["bias_correction"]), pd.data_ptr(), g.data_ptr(), state["exp_avg"].data_ptr(),
                    state["exp_avg_sq"].data_ptr(), pd.numel(), N.ptr(dev_lp), N.ptr(host_lp), lp_bf16)
                if rc != 0:
                    raise RuntimeError(f"dsb_cpu_adam_step failed: {rc}")
        return loss
