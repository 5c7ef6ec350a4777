"""OnebitLamb: LAMB whose (scaled) momentum is communicated with the 1-bit compressed all-reduce after warm-up.

Semantics per reference deepspeed/runtime/fp16/onebit/lamb.py:11-471: warm-up = plain LAMB (per-tensor trust ratio
`lamb_coeff`, clipped to [min_coeff, max_coeff]) while an EMA of the coefficients (`lamb_coeff_freeze`, coeff_beta) is
tracked; at `freeze_step` the variance and the coefficients are frozen, per-tensor `scaling_coeff` (mean momentum RMS
over the tensor's RMS) equalises momentum magnitudes so all tensors can share ONE fused flat buffer, the engine's
gradient all-reduce is disabled and the flat momentum goes through `compressed_allreduce`. In the compression phase
the effective coefficient is `lamb_coeff_freeze * clamp(factor, ...)` where factor comes from the ratio of a freshly
re-estimated variance to the frozen one (factor_min/max, factor_threshold limit its step-to-step change)."""
import types

import numpy as np
import torch
import torch.distributed as dist

from ....utils.logging import logger
from .adam import _gcd


class OnebitLamb(torch.optim.Optimizer):
    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., max_coeff=10.0, min_coeff=0.01,
                 amsgrad=False, cuda_aware=False, comm_backend_name="nccl", coeff_beta=0.9, factor_max=4.0,
                 factor_min=0.5, factor_threshold=0.1):
        if amsgrad:
            raise RuntimeError("1-bit Lamb does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm, max_coeff=max_coeff, min_coeff=min_coeff)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        assert dist.is_initialized() is True, "Please initialize the torch distributed backend."
        self.deepspeed = deepspeed
        self.lamb_freeze_key = False
        self.initialize = False
        self.freeze_step = freeze_step
        self.cuda_aware = cuda_aware
        self.coeff_beta = coeff_beta
        self.factor_max = factor_max
        self.factor_min = factor_min
        self.factor_threshold = factor_threshold
        self.using_pipeline = False
        self.comm_backend_name = comm_backend_name
        self.comm_backend_handle = None
        if self.comm_backend_name == "nccl":
            from ...comm.nccl import NcclBackend
            self.using_pipeline = hasattr(self.deepspeed, "pipeline_enable_backward_allreduce")
            self.comm_backend_handle = NcclBackend(self.deepspeed.mpu if self.deepspeed is not None else None)
        elif self.comm_backend_name == "mpi":
            from ...comm.mpi import MpiBackend
            self.comm_backend_handle = MpiBackend(cuda_aware)
        self.size = self.comm_backend_handle.size
        self.divider = int(self.size * 32 / _gcd(self.size, 32))
        self.exp_avg_flat = []
        self.dummy_exp_avg = {}
        self.corrected_tensor_sizes = []
        self.server_chunk_sizes = []
        self.worker_errors = []
        self.server_errors = []
        self.lamb_coeffs = []

    def _all_params(self):
        for group in self.param_groups:
            for p in group["params"]:
                yield group, p

    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        if grads is None:
            grads_group = [None] * len(self.param_groups)
        elif isinstance(grads, types.GeneratorType):
            grads_group = [grads]
        elif type(grads[0]) != list:
            grads_group = [grads]
        else:
            grads_group = grads
        del self.lamb_coeffs[:]

        if self.lamb_freeze_key:
            exp_avg_last_step = [[self.state[p]["exp_avg"].detach().clone() for p in group["params"]]
                                 for group in self.param_groups]
            if "scaling_coeff" not in self.state[self.param_groups[0]["params"][0]]:
                # first compression step: per-tensor momentum scaling so tensors share one error-feedback scale
                momentum_scales = [[(torch.norm(self.state[p]["exp_avg"]) /
                                     np.sqrt(torch.numel(self.state[p]["exp_avg"]))).item() for p in group["params"]]
                                   for group in self.param_groups]
                united_scale = sum(sum(x) for x in momentum_scales) / sum(len(x) for x in momentum_scales)
                for i, group in enumerate(self.param_groups):
                    for j, p in enumerate(group["params"]):
                        self.state[p]["scaling_coeff"] = united_scale / max(momentum_scales[i][j], 1e-20)

        for group, grads_this_group in zip(self.param_groups, grads_group):
            if grads_this_group is None:
                grads_this_group = [None] * len(group["params"])
            for p, grad in zip(group["params"], grads_this_group):
                if p.grad is None and grad is None:
                    continue
                if grad is None:
                    grad = p.grad.data
                if grad.is_sparse:
                    raise RuntimeError("1-bit Lamb does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0 or (len(state) == 1 and "scaling_coeff" in state):
                    state["step"] = 0
                    state["lamb_coeff_freeze"] = 0.0
                    state["last_factor"] = 1.0
                    state["exp_avg"] = torch.zeros_like(p.data)
                    state["exp_avg_sq"] = torch.zeros_like(p.data)
                    state["exp_avg_sq_fresh"] = torch.zeros_like(p.data)
                if not self.initialize:
                    self.lamb_freeze_key = False
                exp_avg, exp_avg_sq, exp_avg_sq_fresh = state["exp_avg"], state["exp_avg_sq"], state["exp_avg_sq_fresh"]
                beta1, beta2 = group["betas"]
                max_coeff, min_coeff = group["max_coeff"], group["min_coeff"]
                state["step"] += 1
                if self.lamb_freeze_key is False:
                    # warm-up: vanilla LAMB
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                    if state["step"] == self.freeze_step:
                        exp_avg_sq_fresh.data = exp_avg_sq.detach().clone()
                    grad = None
                    if self.initialize:
                        weight_norm = p.data.pow(2).sum().sqrt()
                        update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                        if group["weight_decay"] > 0.0:
                            update += group["weight_decay"] * p.data
                        update_norm = update.pow(2).sum().sqrt()
                        lamb_coeff = 1.0
                        if weight_norm != 0 and update_norm != 0:
                            lamb_coeff = (weight_norm / update_norm).item()
                            lamb_coeff = min(max(lamb_coeff, min_coeff), max_coeff)
                        if lamb_coeff != 1.0:
                            state["lamb_coeff_freeze"] = self.coeff_beta * state["lamb_coeff_freeze"] + \
                                (1 - self.coeff_beta) * lamb_coeff
                        self.lamb_coeffs.append(lamb_coeff)
                        with torch.no_grad():
                            p.add_(-group["lr"] * lamb_coeff * update)
                else:
                    # compression stage: local momentum, scaled so all tensors share one magnitude
                    if self.initialize:
                        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                        exp_avg.mul_(self.state[p]["scaling_coeff"])
                    grad = None

        # one-time setup of the fused flat momentum buffer + error-feedback buffers
        if len(self.exp_avg_flat) == 0:
            momentum_groups = []
            tensor_size = 0
            for group in self.param_groups:
                for p in group["params"]:
                    momentum_groups.append(self.state[p]["exp_avg"])
                    tensor_size += torch.numel(p.data)
            corrected = tensor_size
            if tensor_size % (self.size * self.divider) != 0:
                corrected += (self.size * self.divider) - (tensor_size % (self.size * self.divider))
            self.corrected_tensor_sizes.append(corrected)
            self.server_chunk_sizes.append(corrected // self.size)
            flat = torch.cat([m.detach().float().view(-1) for m in momentum_groups])
            self.exp_avg_flat.append(flat)
            off = 0
            for group in self.param_groups:
                for p in group["params"]:
                    n = p.numel()
                    self.state[p]["exp_avg"] = flat.narrow(0, off, n).view_as(p.data)
                    off += n

        if self.initialize and len(self.worker_errors) == 0:
            dev = self.exp_avg_flat[0].device
            for i in range(len(self.exp_avg_flat)):
                self.worker_errors.append(torch.zeros(self.corrected_tensor_sizes[i], device=dev))
                self.server_errors.append(torch.zeros(self.server_chunk_sizes[i], device=dev))

        if self.lamb_freeze_key:
            if self.size > 1:
                for i in range(len(self.exp_avg_flat)):
                    if not self.initialize:
                        self.exp_avg_flat[i].zero_()
                        self.worker_errors[i].zero_()
                        self.server_errors[i].zero_()
                    res = self.comm_backend_handle.compressed_allreduce(
                        self.exp_avg_flat[i], self.worker_errors[i], self.server_errors[i],
                        self.deepspeed.local_rank if self.deepspeed else 0)
                    if res.data_ptr() != self.exp_avg_flat[i].data_ptr():
                        self.exp_avg_flat[i].copy_(res)

        if self.lamb_freeze_key and self.initialize:
            for i, group in enumerate(self.param_groups):
                bias_correction = 1 if group["bias_correction"] else 0
                for j, p in enumerate(group["params"]):
                    state = self.state[p]
                    exp_avg, exp_avg_sq, exp_avg_sq_fresh = state["exp_avg"], state["exp_avg_sq"], state["exp_avg_sq_fresh"]
                    beta1, beta2 = group["betas"]
                    exp_avg.div_(self.state[p]["scaling_coeff"])
                    # momentum change approximates the averaged gradient -> refresh a "fresh" variance estimate
                    if "exp_avg_mask" in group:
                        if exp_avg.device != group["exp_avg_mask"].device:
                            group["exp_avg_mask"] = group["exp_avg_mask"].to(device=exp_avg.device)
                        exp_avg.mul_(group["exp_avg_mask"])
                    grad_reconstruct = ((exp_avg - exp_avg_last_step[i][j] * beta1) / (1 - beta1))
                    exp_avg_sq_fresh.mul_(beta2).addcmul_(grad_reconstruct, grad_reconstruct, value=1 - beta2)
                    denom = exp_avg_sq.sqrt() + group["eps"]
                    update_prelim = exp_avg / denom
                    if group["weight_decay"] > 0.0:
                        update = update_prelim + group["weight_decay"] * p.data
                    else:
                        update = update_prelim
                    lamb_coeff = 1.0
                    update_norm = update.pow(2).sum().sqrt()
                    denom_real = exp_avg_sq_fresh.sqrt() + group["eps"]
                    factor = (denom / denom_real).max().item()
                    if group["weight_decay"] > 0.0:
                        update_ratio = min(1.0, (update_prelim.pow(2).sum().sqrt() / update_norm).item())
                        factor = factor * update_ratio + (1.0 - update_ratio)
                    factor = min(max(factor, self.factor_min), self.factor_max)
                    factor = min(factor, state["last_factor"] * (1.0 + self.factor_threshold))
                    factor = max(factor, state["last_factor"] * (1.0 - self.factor_threshold))
                    state["last_factor"] = factor
                    lamb_coeff = state["lamb_coeff_freeze"] * factor
                    self.lamb_coeffs.append(lamb_coeff)
                    with torch.no_grad():
                        p.add_(-group["lr"] * lamb_coeff * update)
            del exp_avg_last_step[:]
            exp_avg_last_step = None

        if not self.initialize:
            self.lamb_freeze_key = False
            self.initialize = True
            logger.info(f"Finished the initialization step at rank {dist.get_rank()}")
            return loss

        if self.lamb_freeze_key is False:
            first = self.state[self.param_groups[0]["params"][0]]
            if first["step"] >= self.freeze_step:
                logger.info("OnebitLamb - starting compressed communication")
                self.lamb_freeze_key = True
                if self.deepspeed is not None:
                    if self.using_pipeline:
                        self.deepspeed.pipeline_enable_backward_allreduce = False
                    else:
                        self.deepspeed.enable_backward_allreduce = False
        return loss

    def load_state_dict(self, state_dict):
        for i, group in enumerate(self.param_groups):
            if "exp_avg_mask" in group:
                state_dict["param_groups"][i]["exp_avg_mask"] = group["exp_avg_mask"]
            elif "exp_avg_mask" in state_dict["param_groups"][i]:
                state_dict["param_groups"][i].pop("exp_avg_mask")
        super().load_state_dict(state_dict)
        # the fused flat buffer and the error buffers are rebuilt on the next step
        del self.exp_avg_flat[:]
        self.dummy_exp_avg.clear()
        del self.corrected_tensor_sizes[:]
        del self.server_chunk_sizes[:]
        first = self.state[self.param_groups[0]["params"][0]]
        if first["step"] < self.freeze_step:
            if dist.get_rank() == 0:
                logger.info("Checkpoint loaded and OnebitLamb warmup stage starts/continues.")
            if self.lamb_freeze_key is True:
                self.lamb_freeze_key = False
                if self.deepspeed is not None:
                    if self.using_pipeline:
                        self.deepspeed.pipeline_enable_backward_allreduce = True
                    else:
                        self.deepspeed.enable_backward_allreduce = True
            for group in self.param_groups:
                for p in group["params"]:
                    self.state[p]["lamb_coeff_freeze"] = 0.0
                    self.state[p]["last_factor"] = 1.0
                    if "scaling_coeff" in self.state[p]:
                        self.state[p].pop("scaling_coeff")
        else:
            if dist.get_rank() == 0:
                logger.info("Checkpoint loaded and OnebitLamb compression stage starts/continues.")
            if self.lamb_freeze_key is False:
                self.lamb_freeze_key = True
                if self.deepspeed is not None:
                    if self.using_pipeline:
                        self.deepspeed.pipeline_enable_backward_allreduce = False
                    else:
                        self.deepspeed.enable_backward_allreduce = False
        del self.worker_errors[:]
        del self.server_errors[:]

    def get_lamb_coeffs(self):
        return self.lamb_coeffs
