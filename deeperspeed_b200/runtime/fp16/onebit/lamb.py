"""OnebitLamb: LAMB whose momentum travels through the 1-bit compressed all-reduce after a warm-up.

Semantics per reference deepspeed/runtime/fp16/onebit/lamb.py:11-471:
  warm-up      plain LAMB (trust ratio ||w|| / ||update|| clipped to [min_coeff, max_coeff]); an EMA of the ratios
               (`lamb_coeff_freeze`, rate `coeff_beta`) is tracked for later.
  freeze_step  variance and ratios freeze. A per-tensor `scaling_coeff` (mean momentum RMS over all tensors divided by
               this tensor's RMS) equalises momentum magnitudes so every tensor can share ONE flat buffer and one
               error-feedback scale; the engine's gradient all-reduce is switched off.
  compression  local gradient -> local momentum -> scale -> `compressed_allreduce` of the flat buffer -> unscale. The
               step uses `lamb_coeff_freeze * factor`, where factor = max(frozen denom / fresh denom) tracks how the
               variance would have moved (the gradient is reconstructed from the momentum change), limited to
               [factor_min, factor_max] and to +-factor_threshold per step.

The momenta of all parameters live in one flat arena from the first step on (views per parameter), so the exchange
needs no gather / scatter; phases are separate methods.
"""
import math
import types

import torch
import torch.distributed as dist

from ....utils.logging import logger
from .adam import _gcd  # noqa: F401  (kept importable from here)


class OnebitLamb(torch.optim.Optimizer):
    def __init__(self, params, deepspeed=None, lr=1e-3, freeze_step=100000, bias_correction=True, betas=(0.9, 0.999),
                 eps=1e-8, eps_inside_sqrt=False, weight_decay=0., max_grad_norm=0., max_coeff=10.0, min_coeff=0.01,
                 amsgrad=False, cuda_aware=False, comm_backend_name="nccl", coeff_beta=0.9, factor_max=4.0,
                 factor_min=0.5, factor_threshold=0.1):
        if amsgrad:
            raise RuntimeError("1-bit Lamb does not support the AMSGrad variant.")
        super().__init__(params, dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps,
                                      weight_decay=weight_decay, max_grad_norm=max_grad_norm, max_coeff=max_coeff,
                                      min_coeff=min_coeff))
        assert dist.is_initialized() is True, "Please initialize the torch distributed backend."
        self.eps_mode = 0 if eps_inside_sqrt else 1
        self.deepspeed = deepspeed
        self.freeze_step = freeze_step
        self.lamb_freeze_key = False      # True in the compression phase
        self.initialize = False           # the very first call only sets state up
        self.cuda_aware = cuda_aware
        self.coeff_beta = coeff_beta
        self.factor_max, self.factor_min, self.factor_threshold = factor_max, factor_min, factor_threshold
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
        self.divider = self.size * 32 // math.gcd(self.size, 32)
        # flat momentum arena(s) and their error-feedback buffers (lists: the reference keeps one entry per arena)
        self.exp_avg_flat = []
        self.dummy_exp_avg = {}
        self.corrected_tensor_sizes = []
        self.server_chunk_sizes = []
        self.worker_errors = []
        self.server_errors = []
        self.lamb_coeffs = []

    # ------------------------------------------------------------------ small helpers
    def _all_params(self):
        for group in self.param_groups:
            for p in group["params"]:
                yield group, p

    def _engine_reduces_gradients(self, enabled):
        if self.deepspeed is None:
            return
        attr = "pipeline_enable_backward_allreduce" if self.using_pipeline else "enable_backward_allreduce"
        setattr(self.deepspeed, attr, enabled)

    def _first_state(self):
        return self.state[self.param_groups[0]["params"][0]]

    def _resolve_grads(self, grads):
        """[(group, p, grad)] for every parameter that has a gradient this step."""
        if grads is None:
            per_group = [None] * len(self.param_groups)
        elif isinstance(grads, types.GeneratorType) or type(grads[0]) != list:
            per_group = [grads]
        else:
            per_group = grads
        out = []
        for group, given in zip(self.param_groups, per_group):
            given = given if given is not None else [None] * len(group["params"])
            for p, g in zip(group["params"], given):
                if g is None:
                    if p.grad is None:
                        continue
                    g = p.grad.data
                if g.is_sparse:
                    raise RuntimeError("1-bit Lamb does not support sparse gradients")
                out.append((group, p, g))
        return out

    def _ensure_state(self, p):
        state = self.state[p]
        if not state or (len(state) == 1 and "scaling_coeff" in state):
            state["step"] = 0
            state["lamb_coeff_freeze"] = 0.0
            state["last_factor"] = 1.0
            for key in ("exp_avg", "exp_avg_sq", "exp_avg_sq_fresh"):
                state[key] = torch.zeros_like(p.data)
        return state

    def _ensure_scaling_coeffs(self):
        """First compression step: scaling_coeff = (mean RMS of all momenta) / (RMS of this momentum)."""
        if "scaling_coeff" in self._first_state():
            return
        rms = {p: (torch.norm(self.state[p]["exp_avg"]) / math.sqrt(self.state[p]["exp_avg"].numel())).item()
               for _, p in self._all_params()}
        mean_rms = sum(rms.values()) / len(rms)
        for p, r in rms.items():
            self.state[p]["scaling_coeff"] = mean_rms / max(r, 1e-20)

    def _ensure_flat_momentum(self):
        """Re-home every exp_avg as a view of one flat fp32 arena (the unit of the compressed exchange)."""
        if self.exp_avg_flat:
            return
        params = [p for _, p in self._all_params()]
        flat = torch.cat([self.state[p]["exp_avg"].detach().float().view(-1) for p in params])
        unit = self.size * self.divider
        padded = -(-flat.numel() // unit) * unit
        self.corrected_tensor_sizes.append(padded)
        self.server_chunk_sizes.append(padded // self.size)
        self.exp_avg_flat.append(flat)
        offset = 0
        for p in params:
            self.state[p]["exp_avg"] = flat.narrow(0, offset, p.numel()).view_as(p.data)
            offset += p.numel()

    def _ensure_error_buffers(self):
        if not self.initialize or self.worker_errors:
            return
        dev = self.exp_avg_flat[0].device
        for padded, chunk in zip(self.corrected_tensor_sizes, self.server_chunk_sizes):
            self.worker_errors.append(torch.zeros(padded, device=dev))
            self.server_errors.append(torch.zeros(chunk, device=dev))

    # ------------------------------------------------------------------ phases
    def _warmup_update(self, group, p, state, grad):
        beta1, beta2 = group["betas"]
        m, v = state["exp_avg"], state["exp_avg_sq"]
        m.mul_(beta1).add_(grad, alpha=1 - beta1)
        v.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
        if state["step"] == self.freeze_step:
            state["exp_avg_sq_fresh"].data = v.detach().clone()
        if not self.initialize:
            return
        update = m / (v.sqrt() + group["eps"])
        if group["weight_decay"] > 0.0:
            update += group["weight_decay"] * p.data
        w_norm, u_norm = p.data.pow(2).sum().sqrt(), update.pow(2).sum().sqrt()
        coeff = 1.0
        if w_norm != 0 and u_norm != 0:
            coeff = min(max((w_norm / u_norm).item(), group["min_coeff"]), group["max_coeff"])
        if coeff != 1.0:
            state["lamb_coeff_freeze"] = self.coeff_beta * state["lamb_coeff_freeze"] + (1 - self.coeff_beta) * coeff
        self.lamb_coeffs.append(coeff)
        with torch.no_grad():
            p.add_(update, alpha=-group["lr"] * coeff)

    def _exchange_momentum(self):
        if self.size == 1:
            return
        rank = self.deepspeed.local_rank if self.deepspeed else 0
        for flat, werr, serr in zip(self.exp_avg_flat, self.worker_errors, self.server_errors):
            averaged = self.comm_backend_handle.compressed_allreduce(flat, werr, serr, rank)
            if averaged.data_ptr() != flat.data_ptr():
                flat.copy_(averaged)

    def _compressed_update(self, previous):
        """Parameter update of the compression phase; `previous[p]` is the momentum before this step."""
        for group, p in self._all_params():
            state = self.state[p]
            beta1, beta2 = group["betas"]
            m, v, v_fresh = state["exp_avg"], state["exp_avg_sq"], state["exp_avg_sq_fresh"]
            m.div_(state["scaling_coeff"])
            mask = group.get("exp_avg_mask")
            if mask is not None:
                if mask.device != m.device:
                    mask = group["exp_avg_mask"] = mask.to(device=m.device)
                m.mul_(mask)
            # the averaged gradient is not available any more: recover it from the momentum recursion
            g_hat = (m - previous[p] * beta1) / (1 - beta1)
            v_fresh.mul_(beta2).addcmul_(g_hat, g_hat, value=1 - beta2)
            denom = v.sqrt() + group["eps"]
            direction = m / denom
            update = direction + group["weight_decay"] * p.data if group["weight_decay"] > 0.0 else direction
            factor = (denom / (v_fresh.sqrt() + group["eps"])).max().item()
            if group["weight_decay"] > 0.0:
                share = min(1.0, (direction.pow(2).sum().sqrt() / update.pow(2).sum().sqrt()).item())
                factor = factor * share + (1.0 - share)
            factor = min(max(factor, self.factor_min), self.factor_max)
            last = state["last_factor"]
            factor = max(min(factor, last * (1.0 + self.factor_threshold)), last * (1.0 - self.factor_threshold))
            state["last_factor"] = factor
            coeff = state["lamb_coeff_freeze"] * factor
            self.lamb_coeffs.append(coeff)
            with torch.no_grad():
                p.add_(update, alpha=-group["lr"] * coeff)

    # ------------------------------------------------------------------ step
    def step(self, closure=None, grads=None):
        loss = closure() if closure is not None else None
        work = self._resolve_grads(grads)
        del self.lamb_coeffs[:]
        if not self.initialize:
            self.lamb_freeze_key = False
        compressing = self.lamb_freeze_key
        previous = None
        if compressing:
            previous = {p: self.state[p]["exp_avg"].detach().clone() for _, p in self._all_params()}
            self._ensure_scaling_coeffs()

        for group, p, grad in work:
            state = self._ensure_state(p)
            state["step"] += 1
            if not compressing:
                self._warmup_update(group, p, state, grad)
            elif self.initialize:
                beta1 = group["betas"][0]
                state["exp_avg"].mul_(beta1).add_(grad, alpha=1 - beta1).mul_(state["scaling_coeff"])

        self._ensure_flat_momentum()
        self._ensure_error_buffers()
        if compressing:
            self._exchange_momentum()
            if self.initialize:
                self._compressed_update(previous)

        if not self.initialize:
            self.lamb_freeze_key = False
            self.initialize = True
            logger.info(f"Finished the initialization step at rank {dist.get_rank()}")
            return loss
        if not self.lamb_freeze_key and self._first_state()["step"] >= self.freeze_step:
            logger.info("OnebitLamb - starting compressed communication")
            self.lamb_freeze_key = True
            self._engine_reduces_gradients(False)
        return loss

    def load_state_dict(self, state_dict):
        """Restore; the flat arena and the error buffers are rebuilt on the next step, the phase follows `step`."""
        for live, saved in zip(self.param_groups, state_dict["param_groups"]):
            if "exp_avg_mask" in live:
                saved["exp_avg_mask"] = live["exp_avg_mask"]
            else:
                saved.pop("exp_avg_mask", None)
        super().load_state_dict(state_dict)
        for buffers in (self.exp_avg_flat, self.corrected_tensor_sizes, self.server_chunk_sizes, self.worker_errors,
                        self.server_errors):
            del buffers[:]
        self.dummy_exp_avg.clear()
        compressing = self._first_state()["step"] >= self.freeze_step
        if dist.get_rank() == 0:
            logger.info("Checkpoint loaded and OnebitLamb %s stage starts/continues.",
                        "compression" if compressing else "warmup")
        if compressing != self.lamb_freeze_key:
            self.lamb_freeze_key = compressing
            self._engine_reduces_gradients(not compressing)
        if not compressing:
            for _, p in self._all_params():   # the frozen statistics are re-learned during the rest of the warm-up
                self.state[p]["lamb_coeff_freeze"] = 0.0
                self.state[p]["last_factor"] = 1.0
                self.state[p].pop("scaling_coeff", None)

    def get_lamb_coeffs(self):
        return self.lamb_coeffs
