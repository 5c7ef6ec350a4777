"""OnebitAdam: Adam whose momentum is communicated with the error-compensated 1-bit all-reduce after a warm-up.

Semantics per reference deepspeed/runtime/fp16/onebit/adam.py:14-322: during `freeze_step` warm-up steps it is plain
Adam (gradients all-reduced by the engine); afterwards the variance is frozen, the engine's gradient all-reduce is
switched off and each rank updates its LOCAL momentum, which is then averaged with `compressed_allreduce`
(fused sm_100a kernel on GPUs). No bias correction in either phase. `exp_avg_mask` zeroes masked momentum entries.
Checkpoint load drops the error buffers and re-derives the phase from `state['step']` vs `freeze_step`."""
import types

import torch
import torch.distributed as dist

from ....utils.logging import logger


class OnebitAdam(torch.optim.Optimizer):
    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("1-bit Adam does not support the AMSGrad variant.")
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self.eps_mode = 0 if eps_inside_sqrt else 1
        assert dist.is_initialized() is True, "Please initialize the torch distributed backend."
        self.comm_time = 0.0
        self.step_time = 0.0
        self.ave_step = 1
        self.bk_time = 0.0
        self.deepspeed = deepspeed
        self.adam_freeze_key = False
        self.initialize = False
        self.freeze_step = freeze_step
        self.cuda_aware = cuda_aware
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
        # pad so numel splits evenly into `size` server chunks whose sign words are whole 32-bit words
        self.divider = int(self.size * 32 / _gcd(self.size, 32))

    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        gather_time = allgather_time = 0
        if grads is None:
            grads_group = [None] * len(self.param_groups)
        elif isinstance(grads, types.GeneratorType):
            grads_group = [grads]
        elif type(grads[0]) != list:
            grads_group = [grads]
        else:
            grads_group = grads
        for group, grads_this_group in zip(self.param_groups, grads_group):
            if grads_this_group is None:
                grads_this_group = [None] * len(group["params"])
            for p, grad in zip(group["params"], grads_this_group):
                if p.grad is None and grad is None:
                    continue
                if grad is None:
                    grad = p.grad.data
                if grad.is_sparse:
                    raise RuntimeError("1-bit Adam does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p.data)
                    state["exp_avg_sq"] = torch.zeros_like(p.data)
                if not self.initialize or (self.adam_freeze_key and "worker_error" not in state):
                    state["tensor_size"] = torch.numel(p.data)
                    state["corrected_tensor_size"] = state["tensor_size"]
                    if state["tensor_size"] % (self.size * self.divider) != 0:
                        state["corrected_tensor_size"] += (self.size * self.divider) - (state["tensor_size"] % (self.size * self.divider))
                    state["server_chunk_size"] = state["corrected_tensor_size"] // self.size
                    if torch.cuda.is_available() and p.is_cuda:
                        torch.cuda.empty_cache()
                    state["worker_error"] = torch.zeros(state["corrected_tensor_size"], device=p.device, dtype=torch.float32)
                    state["server_error"] = torch.zeros(state["server_chunk_size"], device=p.device, dtype=torch.float32)
                    if torch.cuda.is_available() and p.is_cuda:
                        torch.cuda.empty_cache()
                    self.adam_freeze_key = False if not self.initialize else self.adam_freeze_key
                    self.initialize = True
                exp_avg, exp_avg_sq = state["exp_avg"], state["exp_avg_sq"]
                beta1, beta2 = group["betas"]
                state["step"] += 1
                if self.adam_freeze_key is False:
                    exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                    exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                    grad = None
                    update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                else:
                    if "non_freeze" in group.keys() and group["non_freeze"] is True:
                        dist.all_reduce(grad)
                        grad.mul_(1 / dist.get_world_size())
                        exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                        exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
                        grad = None
                    else:
                        if self.size > 1:
                            exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                            grad = None
                            m32 = exp_avg.float() if exp_avg.dtype != torch.float32 else exp_avg
                            res = self.comm_backend_handle.compressed_allreduce(m32, state["worker_error"],
                                                                                state["server_error"],
                                                                                self.deepspeed.local_rank if self.deepspeed else 0)
                            exp_avg.copy_(res.view_as(exp_avg))
                        else:
                            exp_avg.mul_(beta1).add_(grad, alpha=1 - beta1)
                        if "exp_avg_mask" in group:
                            if exp_avg.device != group["exp_avg_mask"].device:
                                group["exp_avg_mask"] = group["exp_avg_mask"].to(device=exp_avg.device)
                            exp_avg.mul_(group["exp_avg_mask"])
                    update = exp_avg / (exp_avg_sq.sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update += group["weight_decay"] * p.data
                with torch.no_grad():
                    p.add_(-group["lr"] * update)
            if not self.initialize:
                self.initialize = True
        if self.adam_freeze_key is False:
            any_state = next((self.state[p] for g in self.param_groups for p in g["params"] if p in self.state), None)
            if any_state is not None and any_state["step"] >= self.freeze_step:
                logger.info("OnebitAdam - starting compressed communication")
                self.adam_freeze_key = True
                if self.deepspeed is not None:
                    if self.using_pipeline:
                        self.deepspeed.pipeline_enable_backward_allreduce = False
                    else:
                        self.deepspeed.enable_backward_allreduce = False
        return loss

    def load_state_dict(self, state_dict):
        """Restore, dropping stale error-feedback buffers and re-deriving the phase (warm-up vs compression)."""
        for i, group in enumerate(self.param_groups):
            if "exp_avg_mask" in group:
                state_dict["param_groups"][i]["exp_avg_mask"] = group["exp_avg_mask"]
            elif "exp_avg_mask" in state_dict["param_groups"][i]:
                state_dict["param_groups"][i].pop("exp_avg_mask")
        super().load_state_dict(state_dict)
        first = next((self.state[p] for g in self.param_groups for p in g["params"] if p in self.state), None)
        step = first["step"] if first is not None else 0
        if step < self.freeze_step:
            if dist.get_rank() == 0:
                logger.info("Checkpoint loaded and OnebitAdam warmup stage starts/continues.")
            if self.adam_freeze_key is True:
                self.adam_freeze_key = False
                if self.deepspeed is not None:
                    if self.using_pipeline:
                        self.deepspeed.pipeline_enable_backward_allreduce = True
                    else:
                        self.deepspeed.enable_backward_allreduce = True
        else:
            if dist.get_rank() == 0:
                logger.info("Checkpoint loaded and OnebitAdam compression stage starts/continues.")
            if self.adam_freeze_key is False:
                self.adam_freeze_key = True
                if self.deepspeed is not None:
                    if self.using_pipeline:
                        self.deepspeed.pipeline_enable_backward_allreduce = False
                    else:
                        self.deepspeed.enable_backward_allreduce = False
        for group in self.param_groups:
            for p in group["params"]:
                if p in self.state:
                    self.state[p].pop("worker_error", None)
                    self.state[p].pop("server_error", None)


def _gcd(a, b):
    while b:
        a, b = b, a % b
    return a
