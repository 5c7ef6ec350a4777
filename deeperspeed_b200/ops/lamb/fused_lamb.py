"""FusedLamb: LAMB with per-tensor trust ratio computed on the device in two kernels (moments + |w|,|u| partials,
then ratio + apply). API per reference deepspeed/ops/lamb/fused_lamb.py:38-189; math per SURVEY C.6."""
import torch

from .. import kernels as K
from .. import native as N


class FusedLamb(torch.optim.Optimizer):
    # under ZeRO-1/2 on one NVLink domain the sharded step runs as the fused reduce-scatter + LAMB + all-gather
    # kernels of parallel/fused_zero.py (per-tensor trust ratio summed over the shards)
    supports_fused_zero_lamb = True

    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, eps_inside_sqrt=False,
                 weight_decay=0.0, max_grad_norm=0.0, max_coeff=10.0, min_coeff=0.01, amsgrad=False):
        if amsgrad:
            raise RuntimeError("FusedLamb does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm, max_coeff=max_coeff, min_coeff=min_coeff)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.lamb_coeffs = []

    def step(self, closure=None, grads=None, output_params=None, scale=1.0, grad_norms=None):
        loss = closure() if closure is not None else None

        def _norm_list(x):
            if x is None:
                return [None] * len(self.param_groups)
            if isinstance(x, (list, tuple)) and len(x) > 0 and not isinstance(x[0], (list, tuple)):
                return [x]
            return x

        grads_group, output_params_group = _norm_list(grads), _norm_list(output_params)
        if grad_norms is None:
            grad_norms = [None] * len(self.param_groups)
        del self.lamb_coeffs[:]
        for group, grads_this, outs_this, grad_norm_group in zip(self.param_groups, grads_group, output_params_group,
                                                                 grad_norms):
            if grads_this is None:
                grads_this = [None] * len(group["params"])
            if outs_this is None:
                outs_this = [None] * len(group["params"])
            if grad_norm_group is None:
                grad_norm_group = [None] * len(group["params"])
            elif not isinstance(grad_norm_group, list):
                grad_norm_group = [grad_norm_group]
            bias_correction = 1 if group["bias_correction"] else 0
            for p, grad, output_param, grad_norm in zip(group["params"], grads_this, outs_this, grad_norm_group):
                combined_scale = scale
                if group["max_grad_norm"] > 0 and grad_norm is not None:
                    clip = ((grad_norm / scale) + 1e-6) / group["max_grad_norm"]
                    if clip > 1:
                        combined_scale = clip * scale
                if p.grad is None and grad is None:
                    continue
                if grad is None:
                    grad = p.grad.data
                if grad.is_sparse:
                    raise RuntimeError("FusedLamb does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p.data, dtype=torch.float32)
                    state["exp_avg_sq"] = torch.zeros_like(p.data, dtype=torch.float32)
                beta1, beta2 = group["betas"]
                state["step"] += 1
                coeff = self._one(p.data, output_param, grad, state, group, beta1, beta2, combined_scale,
                                  bias_correction)
                self.lamb_coeffs.append(coeff)
        return loss

    def _one(self, p, out_p, grad, state, group, beta1, beta2, combined_scale, bias_correction):
        m, v = state["exp_avg"], state["exp_avg_sq"]
        step = state["step"]
        if p.is_cuda:
            lib = N.cuda_lib()
            coeff = torch.empty(1, dtype=torch.float32, device=p.device)
            sc = N.scratch(lib.dsb_lamb_scratch_bytes(p.numel()), p.device, "lamb")
            g = grad if grad.is_contiguous() else grad.contiguous()
            rc = lib.dsb_lamb_step(N.ptr(p), N.dtype_code(p.dtype), N.ptr(out_p),
                                   N.dtype_code(out_p.dtype) if out_p is not None else 0, N.ptr(g),
                                   N.dtype_code(g.dtype), N.ptr(m), N.ptr(v), p.numel(), float(group["lr"]),
                                   float(beta1), float(beta2), float(group["eps"]), float(group["weight_decay"]),
                                   float(group["max_coeff"]), float(group["min_coeff"]), int(self.eps_mode == 0),
                                   bias_correction, int(step), float(combined_scale), N.ptr(coeff), N.ptr(sc),
                                   N.stream_ptr())
            N.check(rc, "dsb_lamb_step")
            K._count(2)
            return coeff
        # CPU reference path (same math)
        pf = p.float()
        gf = grad.float() / combined_scale
        m.mul_(beta1).add_(gf, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(gf, gf, value=1 - beta2)
        denom = (v + group["eps"]).sqrt() if self.eps_mode == 0 else v.sqrt() + group["eps"]
        update = m / denom + group["weight_decay"] * pf
        wn, un = pf.norm(), update.norm()
        coeff = 1.0
        if wn > 0 and un > 0:
            coeff = float(min(max(wn / un, group["min_coeff"]), group["max_coeff"]))
        if bias_correction:
            step_size = group["lr"] * (1 - beta2 ** step) ** 0.5 / (1 - beta1 ** step)
        else:
            step_size = group["lr"]
        pf = pf - step_size * coeff * update
        p.copy_(pf)
        if out_p is not None:
            out_p.copy_(pf)
        return torch.tensor([coeff])

    def get_lamb_coeffs(self):
        return [float(x) for c in self.lamb_coeffs for x in torch.as_tensor(c).reshape(-1).tolist()]
