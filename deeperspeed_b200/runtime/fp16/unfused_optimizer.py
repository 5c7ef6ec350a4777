"""FP16_UnfusedOptimizer: fp32 master weights kept PER PARAMETER, for optimizers whose update is not element-wise
over a flat buffer (LAMB's per-tensor trust ratio) or for arbitrary client optimizers.

Protocol (attribute names, `state_dict` keys, step / backward / overflow semantics) per reference
deepspeed/runtime/fp16/unfused_optimizer.py:21-407. How it is done here:
  * the norm of every group and the non-finite check come from the device-side sum-of-squares kernel accumulated over the
    gradients; ONE host read per step decides skip / clip (the reference synchronises per tensor);
  * unscale + clip is a single factor folded into the fp16 -> fp32 gradient conversion (multi-tensor `_foreach` ops), and
    the fp32 -> fp16 weight copy-back is one multi-tensor copy.
"""
import torch
import torch.distributed as dist

from ...ops import kernels as K
from ...utils.logging import logger
from .loss_scaler import INITIAL_LOSS_SCALE, SCALE_WINDOW, MIN_LOSS_SCALE


class FP16_UnfusedOptimizer:
    def __init__(self, init_optimizer, static_loss_scale=1.0, dynamic_loss_scale=False, dynamic_loss_args=None,
                 verbose=True, mpu=None, clip_grad=0.0, fused_lamb_legacy=False, deepspeed=None):
        self.deepspeed = deepspeed          # the engine, when it builds the wrapper (reference unfused_optimizer.py:31)
        self.fused_lamb_legacy = fused_lamb_legacy
        self.optimizer = init_optimizer
        self.verbose, self.clip_grad, self.mpu = verbose, clip_grad, mpu
        self.norm_type = 2
        self.overflow = False

        # the wrapped optimizer is re-pointed at fp32 copies; the model keeps the 16-bit tensors
        self.fp16_groups, self.fp32_groups = [], []
        for group in self.optimizer.param_groups:
            model_params = group["params"]
            masters = [p.detach().clone().float().requires_grad_() for p in model_params]
            self.fp16_groups.append(model_params)
            self.fp32_groups.append(masters)
            group["params"] = masters

        self.cur_iter = 0
        self.dynamic_loss_scale = bool(dynamic_loss_scale)
        if self.dynamic_loss_scale:
            args = dynamic_loss_args or {INITIAL_LOSS_SCALE: 2.0 ** 16, SCALE_WINDOW: 1000, MIN_LOSS_SCALE: 0.25}
            self.cur_scale = args[INITIAL_LOSS_SCALE]
            self.scale_window = args[SCALE_WINDOW]
            self.min_loss_scale = args[MIN_LOSS_SCALE]
            self.scale_factor = 2.0
            self.last_overflow_iter = -1
        else:
            self.cur_scale = static_loss_scale
        self.initialize_optimizer_states()

    # ------------------------------------------------------------------ gradients
    def zero_grad(self, set_grads_to_None=True):
        for p in (p for group in self.fp16_groups for p in group):
            if set_grads_to_None:
                p.grad = None
            elif p.grad is not None:
                p.grad.detach_()
                p.grad.zero_()

    def _counts_here(self, p):
        """Replicated (non-model-parallel) parameters are counted by model-parallel rank 0 only."""
        return self.mpu is None or getattr(p, "model_parallel", False) or self.mpu.get_model_parallel_rank() == 0

    def _norms_and_overflow(self):
        """([norm of each group], overflow). Sums of squares and a non-finite flag are accumulated on the device."""
        dev = self.fp16_groups[0][0].device
        flag = torch.zeros(1, dtype=torch.int32, device=dev)
        sums = []
        for group in self.fp16_groups:
            acc = torch.zeros(1, dtype=torch.float32, device=dev)
            for p in group:
                if p.grad is not None and self._counts_here(p):
                    K.sumsq(p.grad.data.contiguous().view(-1), out=acc, flag=flag, accumulate=True)
            sums.append(acc)
        packed = torch.cat(sums + [flag.float()])
        if self.mpu is not None and dist.is_initialized():
            dist.all_reduce(packed, group=self.mpu.get_model_parallel_group())
        *squares, bad = packed.tolist()
        finite = all(v == v and v != float("inf") for v in squares)
        return [v ** 0.5 if v == v else -1 for v in squares], (bad > 0 or not finite)

    def unscale_and_clip_grads(self, norm_groups, apply_scale=True, prev_scale=None):
        """The divisor that undoes the loss scale and, if the global norm exceeds `clip_grad`, clips it as well."""
        scale = self.cur_scale if prev_scale is None else prev_scale
        divisor = scale
        if self.clip_grad > 0.0:
            global_norm = sum(n * n for n in norm_groups) ** 0.5
            excess = (global_norm / scale + 1e-6) / self.clip_grad
            if excess > 1:
                divisor = excess * scale
        if apply_scale:
            grads = [p.grad.data for group in self.fp32_groups for p in group if p.grad is not None]
            if grads:
                torch._foreach_mul_(grads, 1.0 / divisor)
        return divisor

    # ------------------------------------------------------------------ step
    def step_fused_lamb(self, closure=None):
        return self.step(closure)

    def step(self, closure=None):
        norms, self.overflow = self._norms_and_overflow()
        scale_used = self.cur_scale
        self._update_scale(self.overflow)
        if self.overflow:
            if self.verbose:
                logger.info("[deepspeed] OVERFLOW! Skipping step. Attempted loss scale: {}, reducing to {}".format(
                    scale_used, self.cur_scale))
            self.zero_grad()
            return self.overflow

        divisor = self.unscale_and_clip_grads(norms, apply_scale=False, prev_scale=scale_used)
        masters = [m for group in self.fp32_groups for m in group]
        models = [p for group in self.fp16_groups for p in group]
        for m, p in zip(masters, models):
            m.grad = torch.zeros_like(m) if p.grad is None else p.grad.to(m.dtype)
        torch._foreach_mul_([m.grad for m in masters], 1.0 / divisor)
        self.optimizer.step()
        for m in masters:
            m.grad = None
        torch._foreach_copy_([p.data for p in models], [m.data for m in masters])
        self.zero_grad()
        return self.overflow

    def backward(self, loss, retain_graph=False):
        (loss.float() * self.cur_scale).backward(retain_graph=retain_graph)

    def _update_scale(self, skip):
        """Halve on overflow (not below min_loss_scale); double after `scale_window` clean steps since the last one."""
        before = self.cur_scale
        if not self.dynamic_loss_scale:
            if skip:
                logger.info("Grad overflow on iteration %s", self.cur_iter)
                logger.info("Using static loss scale of %s", self.cur_scale)
        elif skip:
            self.cur_scale = max(before / self.scale_factor, self.min_loss_scale)
            self.last_overflow_iter = self.cur_iter
            if self.verbose:
                logger.info("Grad overflow on iteration: %s", self.cur_iter)
                logger.info(f"Reducing dynamic loss scale from {before} to {self.cur_scale}")
        else:
            clean = self.cur_iter - self.last_overflow_iter - 1
            if clean > 0 and clean % self.scale_window == 0:
                self.cur_scale = before * self.scale_factor
                if self.verbose:
                    logger.info(f"No Grad overflow for {self.scale_window} iterations")
                    logger.info(f"Increasing dynamic loss scale from {before} to {self.cur_scale}")
        self.cur_iter += 1

    # ------------------------------------------------------------------ pass-throughs to the wrapped optimizer
    state = property(lambda self: self.optimizer.state, lambda self, v: setattr(self.optimizer, "state", v))
    param_groups = property(lambda self: self.optimizer.param_groups,
                            lambda self, v: setattr(self.optimizer, "param_groups", v))
    loss_scale = property(lambda self: self.cur_scale)

    def __repr__(self):
        return repr(self.optimizer)

    # ------------------------------------------------------------------ checkpoint
    _DYNAMIC_KEYS = ("last_overflow_iter", "scale_factor", "scale_window")

    def state_dict(self):
        sd = {"dynamic_loss_scale": self.dynamic_loss_scale, "cur_scale": self.cur_scale, "cur_iter": self.cur_iter}
        if self.dynamic_loss_scale:
            sd.update({k: getattr(self, k) for k in self._DYNAMIC_KEYS})
        sd["optimizer_state_dict"] = self.optimizer.state_dict()
        sd["fp32_groups"] = self.fp32_groups
        return sd

    def load_state_dict(self, state_dict, load_optimizer_states=True):
        for key in ("dynamic_loss_scale", "cur_scale", "cur_iter"):
            setattr(self, key, state_dict[key])
        if state_dict["dynamic_loss_scale"]:
            for key in self._DYNAMIC_KEYS:
                setattr(self, key, state_dict[key])
        if load_optimizer_states:
            self.optimizer.load_state_dict(state_dict["optimizer_state_dict"])
        # the masters are authoritative (the model's 16-bit weights are a rounding of them)
        for mine, saved in zip(self.fp32_groups, state_dict["fp32_groups"]):
            for m, s in zip(mine, saved):
                m.data.copy_(s.data)

    def refresh_fp32_params(self):
        """Re-derive the masters from the model's current weights (after loading model-only checkpoints)."""
        for masters, models in zip(self.fp32_groups, self.fp16_groups):
            for m, p in zip(masters, models):
                m.data.copy_(p.data)

    def initialize_optimizer_states(self):
        """Create the wrapped optimizer's state tensors with one zero-gradient step that is then undone, so that a
        checkpoint can be loaded before the first real step."""
        masters = [m for group in self.fp32_groups for m in group]
        keep = [m.data.clone() for m in masters]
        for m in masters:
            m.grad = torch.zeros_like(m)
        self.optimizer.step()
        for m, k in zip(masters, keep):
            m.data.copy_(k)
            m.grad = None
        for st in self.optimizer.state.values():   # step counters advanced by the dummy step
            if isinstance(st, dict) and not torch.is_tensor(st.get("step", None)) and "step" in st:
                st["step"] = 0
