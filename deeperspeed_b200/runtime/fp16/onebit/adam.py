"""OnebitAdam: Adam that, after a warm-up, freezes its variance and communicates only the MOMENTUM, through the
error-compensated 1-bit all-reduce (one fused sm_100a kernel on GPUs, runtime/comm/nccl.py).

Semantics per reference deepspeed/runtime/fp16/onebit/adam.py:14-322:
  warm-up (step <= freeze_step): plain Adam without bias correction; the ENGINE averages the gradients.
  compression (afterwards)     : the engine's gradient all-reduce is switched off; every rank folds its LOCAL gradient
                                 into its momentum, the momenta are averaged with `compressed_allreduce`, the frozen
                                 variance scales the update. Groups flagged `non_freeze` keep exact Adam with an
                                 ordinary all-reduce; `exp_avg_mask` zeroes momentum entries after the exchange.
  checkpoints                  : error-feedback buffers are dropped on load and the phase is re-derived from `step`.

The two phases are separate methods and the engine switch lives in one place (`_engine_reduces_gradients`), instead of
one long branchy loop.
"""
import math
import types

import torch
import torch.distributed as dist

from ....utils.logging import logger


def _gcd(a, b):
    return math.gcd(a, b)


class OnebitAdam(torch.optim.Optimizer):
    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., amsgrad=False, cuda_aware=False,
                 comm_backend_name="nccl"):
        if amsgrad:
            raise RuntimeError("1-bit Adam does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm))
        assert dist.is_initialized() is True, "Please initialize the torch distributed backend."
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.comm_time = self.step_time = self.bk_time = 0.0
        self.ave_step = 1
        self.deepspeed = deepspeed
        self.freeze_step = freeze_step
        self.adam_freeze_key = False      # True once the compression phase has started
        self.initialize = False           # True once error-feedback buffers exist
        self.cuda_aware = cuda_aware
        self.comm_backend_name = comm_backend_name
        self.using_pipeline = False
        self.comm_backend_handle = None
        if comm_backend_name == "nccl":
            from ...comm.nccl import NcclBackend
            self.using_pipeline = hasattr(deepspeed, "pipeline_enable_backward_allreduce")
            self.comm_backend_handle = NcclBackend(deepspeed.mpu if deepspeed is not None else None)
        elif comm_backend_name == "mpi":
            from ...comm.mpi import MpiBackend
            self.comm_backend_handle = MpiBackend(cuda_aware)
        self.size = self.comm_backend_handle.size
        # tensors are padded so that each of the `size` server chunks packs into whole 32-bit sign words
        self.divider = self.size * 32 // math.gcd(self.size, 32)

    # ------------------------------------------------------------------ helpers
    def _engine_reduces_gradients(self, enabled):
        """Who averages gradients: the engine (warm-up) or nobody (compression: momenta are exchanged instead)."""
        if self.deepspeed is None:
            return
        attr = "pipeline_enable_backward_allreduce" if self.using_pipeline else "enable_backward_allreduce"
        setattr(self.deepspeed, attr, enabled)

    def _first_state(self):
        for group in self.param_groups:
            for p in group["params"]:
                if p in self.state:
                    return self.state[p]
        return None

    def _ensure_buffers(self, p, state):
        """Moments, and the error-feedback buffers of the compressed exchange (padded to the chunk granularity)."""
        if not state:
            state["step"] = 0
            state["exp_avg"] = torch.zeros_like(p.data)
            state["exp_avg_sq"] = torch.zeros_like(p.data)
        if self.initialize and not (self.adam_freeze_key and "worker_error" not in state):
            return
        n = p.data.numel()
        unit = self.size * self.divider
        padded = -(-n // unit) * unit
        state["tensor_size"], state["corrected_tensor_size"] = n, padded
        state["server_chunk_size"] = padded // self.size
        state["worker_error"] = torch.zeros(padded, device=p.device, dtype=torch.float32)
        state["server_error"] = torch.zeros(padded // self.size, device=p.device, dtype=torch.float32)
        if not self.initialize:
            self.adam_freeze_key = False
        self.initialize = True

    @staticmethod
    def _grads_per_group(grads, n_groups):
        if grads is None:
            return [None] * n_groups
        if isinstance(grads, types.GeneratorType) or type(grads[0]) != list:
            return [grads]
        return grads

    # ------------------------------------------------------------------ the two update rules
    @staticmethod
    def _exact_moments(state, grad, beta1, beta2):
        state["exp_avg"].mul_(beta1).add_(grad, alpha=1 - beta1)
        state["exp_avg_sq"].mul_(beta2).addcmul_(grad, grad, value=1 - beta2)

    def _compressed_momentum(self, state, grad, beta1, group):
        m = state["exp_avg"]
        m.mul_(beta1).add_(grad, alpha=1 - beta1)
        if self.size > 1:
            local_rank = self.deepspeed.local_rank if self.deepspeed else 0
            averaged = self.comm_backend_handle.compressed_allreduce(m if m.dtype == torch.float32 else m.float(),
                                                                     state["worker_error"], state["server_error"],
                                                                     local_rank)
            m.copy_(averaged.view_as(m))
        mask = group.get("exp_avg_mask")
        if mask is not None:
            if mask.device != m.device:
                mask = group["exp_avg_mask"] = mask.to(device=m.device)
            m.mul_(mask)

    def _fused_step(self, p, grad, state, group, beta1):
        """Compression-stage step of one tensor inside the 1-bit all-reduce kernel, when the backend offers it and the
        tensors qualify (flat fp32 on a GPU, no momentum mask). Returns False to fall back to the separate passes."""
        backend = self.comm_backend_handle
        fused = getattr(backend, "fused_adam_supported", None)
        if (fused is None or self.size == 1 or group.get("exp_avg_mask") is not None or self.eps_mode != 1
                or p.data.dtype != torch.float32 or grad.dtype != torch.float32 or not p.data.is_contiguous()
                or not grad.is_contiguous() or not fused(state["exp_avg"], state["worker_error"])):
            return False
        backend.compressed_adam_step(state["exp_avg"], grad, p.data, state["exp_avg_sq"], state["worker_error"],
                                     state["server_error"], beta1, group["lr"], group["eps"], group["weight_decay"])
        return True

    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        for group, given in zip(self.param_groups, self._grads_per_group(grads, len(self.param_groups))):
            given = given if given is not None else [None] * len(group["params"])
            beta1, beta2 = group["betas"]
            for p, grad in zip(group["params"], given):
                if grad is None:
                    if p.grad is None:
                        continue
                    grad = p.grad.data
                if grad.is_sparse:
                    raise RuntimeError("1-bit Adam does not support sparse gradients")
                state = self.state[p]
                self._ensure_buffers(p, state)
                state["step"] += 1
                if not self.adam_freeze_key:
                    self._exact_moments(state, grad, beta1, beta2)
                elif group.get("non_freeze", False) is True:
                    dist.all_reduce(grad)
                    grad.mul_(1 / dist.get_world_size())
                    self._exact_moments(state, grad, beta1, beta2)
                elif self._fused_step(p, grad, state, group, beta1):
                    continue      # momentum, exchange and parameter update ran as one kernel
                else:
                    self._compressed_momentum(state, grad, beta1, group)
                update = state["exp_avg"] / (state["exp_avg_sq"].sqrt() + group["eps"])
                if group["weight_decay"] > 0.0:
                    update += group["weight_decay"] * p.data
                with torch.no_grad():
                    p.add_(update, alpha=-group["lr"])
            self.initialize = True
        if not self.adam_freeze_key:
            first = self._first_state()
            if first is not None and first["step"] >= self.freeze_step:
                logger.info("OnebitAdam - starting compressed communication")
                self.adam_freeze_key = True
                self._engine_reduces_gradients(False)
        return loss

    def load_state_dict(self, state_dict):
        """Restore; masks come from the live groups, error buffers are dropped, the phase follows the loaded step."""
        for live, saved in zip(self.param_groups, state_dict["param_groups"]):
            if "exp_avg_mask" in live:
                saved["exp_avg_mask"] = live["exp_avg_mask"]
            else:
                saved.pop("exp_avg_mask", None)
        super().load_state_dict(state_dict)
        first = self._first_state()
        compressing = (first["step"] if first is not None else 0) >= self.freeze_step
        if dist.get_rank() == 0:
            logger.info("Checkpoint loaded and OnebitAdam %s stage starts/continues.",
                        "compression" if compressing else "warmup")
        if compressing != self.adam_freeze_key:
            self.adam_freeze_key = compressing
            self._engine_reduces_gradients(not compressing)
        for state in self.state.values():
            state.pop("worker_error", None)
            state.pop("server_error", None)
