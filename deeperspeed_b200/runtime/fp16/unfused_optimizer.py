"""FP16_UnfusedOptimizer: per-parameter fp32 master copies for optimizers whose update is not element-wise
(LAMB's per-tensor trust ratio) or arbitrary client optimizers. Protocol per reference
deepspeed/runtime/fp16/unfused_optimizer.py:21-407. Overflow and norm come from the fused device-side kernel over
each gradient (no per-tensor host sync); one host read per step."""
import torch
import torch.distributed as dist

from ...ops import kernels as K
from ...utils.logging import logger
from .loss_scaler import INITIAL_LOSS_SCALE, SCALE_WINDOW, MIN_LOSS_SCALE


class FP16_UnfusedOptimizer:
    def __init__(self, init_optimizer, static_loss_scale=1.0, dynamic_loss_scale=False, dynamic_loss_args=None,
                 verbose=True, mpu=None, clip_grad=0.0, fused_lamb_legacy=False):
        self.fused_lamb_legacy = fused_lamb_legacy
        self.optimizer = init_optimizer
        self.fp16_groups = []
        self.fp32_groups = []
        for i, param_group in enumerate(self.optimizer.param_groups):
            self.fp16_groups.append(param_group["params"])
            fp32_group = [p.clone().float().detach() for p in param_group["params"]]
            for p in fp32_group:
                p.requires_grad = True
            self.fp32_groups.append(fp32_group)
            param_group["params"] = self.fp32_groups[i]
        if dynamic_loss_scale:
            self.dynamic_loss_scale = True
            self.cur_iter = 0
            self.last_overflow_iter = -1
            self.scale_factor = 2.0
            if dynamic_loss_args is None:
                self.cur_scale = 1.0 * 2 ** 16
                self.scale_window = 1000
                self.min_loss_scale = 0.25
            else:
                self.cur_scale = dynamic_loss_args[INITIAL_LOSS_SCALE]
                self.scale_window = dynamic_loss_args[SCALE_WINDOW]
                self.min_loss_scale = dynamic_loss_args[MIN_LOSS_SCALE]
        else:
            self.dynamic_loss_scale = False
            self.cur_iter = 0
            self.cur_scale = static_loss_scale
        self.verbose = verbose
        self.clip_grad = clip_grad
        self.norm_type = 2
        self.mpu = mpu
        self.overflow = False
        self.initialize_optimizer_states()

    def zero_grad(self, set_grads_to_None=True):
        for group in self.fp16_groups:
            for p in group:
                if set_grads_to_None:
                    p.grad = None
                elif p.grad is not None:
                    p.grad.detach_()
                    p.grad.zero_()

    def _norms_and_overflow(self):
        """Per-group norm^2 on the device + one non-finite flag; a single host read."""
        dev = self.fp16_groups[0][0].device
        flag = torch.zeros(1, dtype=torch.int32, device=dev)
        per_group = []
        for group in self.fp16_groups:
            acc = torch.zeros(1, dtype=torch.float32, device=dev)
            for p in group:
                if p.grad is None:
                    continue
                if self.mpu is not None and not getattr(p, "model_parallel", False) \
                        and self.mpu.get_model_parallel_rank() != 0:
                    continue
                g = p.grad.data
                K.sumsq(g.view(-1) if g.is_contiguous() else g.contiguous().view(-1), out=acc, flag=flag,
                        accumulate=True)
            per_group.append(acc)
        packed = torch.cat(per_group + [flag.float()])
        if self.mpu is not None and dist.is_initialized():
            dist.all_reduce(packed, group=self.mpu.get_model_parallel_group())
        vals = packed.tolist()
        overflow = vals[-1] > 0 or any(v != v or v == float("inf") for v in vals[:-1])
        return [v ** 0.5 if v == v else -1 for v in vals[:-1]], overflow

    def step_fused_lamb(self, closure=None):
        return self.step(closure)

    def step(self, closure=None):
        norm_groups, self.overflow = self._norms_and_overflow()
        prev_scale = self.cur_scale
        self._update_scale(self.overflow)
        if self.overflow:
            if self.verbose:
                logger.info("[deepspeed] OVERFLOW! Skipping step. Attempted loss scale: {}, reducing to {}".format(
                    prev_scale, self.cur_scale))
            self.zero_grad()
            return self.overflow
        combined_scale = self.unscale_and_clip_grads(norm_groups, apply_scale=False, prev_scale=prev_scale)
        inv = 1.0 / combined_scale
        for i, group in enumerate(self.fp16_groups):
            for fp32_param, fp16_param in zip(self.fp32_groups[i], group):
                if fp16_param.grad is None:
                    fp32_param.grad = torch.zeros(fp16_param.size(), dtype=fp32_param.dtype, device=fp32_param.device)
                else:
                    fp32_param.grad = fp16_param.grad.to(fp32_param.dtype).mul_(inv)
        self.optimizer.step()
        for fp32_group, fp16_group in zip(self.fp32_groups, self.fp16_groups):
            for fp32_param, fp16_param in zip(fp32_group, fp16_group):
                fp32_param.grad = None
                fp16_param.data.copy_(fp32_param.data)
        self.zero_grad()
        return self.overflow

    def unscale_and_clip_grads(self, norm_groups, apply_scale=True, prev_scale=None):
        scale = self.cur_scale if prev_scale is None else prev_scale
        total_norm = sum(n ** 2.0 for n in norm_groups) ** 0.5
        combined_scale = scale
        if self.clip_grad > 0.0:
            clip = ((total_norm / scale) + 1e-6) / self.clip_grad
            if clip > 1:
                combined_scale = clip * scale
        if apply_scale:
            for group in self.fp32_groups:
                for param in group:
                    if param.grad is not None:
                        param.grad.data.mul_(1.0 / combined_scale)
        return combined_scale

    def backward(self, loss, retain_graph=False):
        scaled_loss = (loss.float()) * self.cur_scale
        scaled_loss.backward(retain_graph=retain_graph)

    def _update_scale(self, skip):
        if self.dynamic_loss_scale:
            prev_scale = self.cur_scale
            if skip:
                self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_loss_scale)
                self.last_overflow_iter = self.cur_iter
                if self.verbose:
                    logger.info("Grad overflow on iteration: %s", self.cur_iter)
                    logger.info(f"Reducing dynamic loss scale from {prev_scale} to {self.cur_scale}")
            else:
                stable_interval = (self.cur_iter - self.last_overflow_iter) - 1
                if (stable_interval > 0) and (stable_interval % self.scale_window == 0):
                    self.cur_scale *= self.scale_factor
                    if self.verbose:
                        logger.info(f"No Grad overflow for {self.scale_window} iterations")
                        logger.info(f"Increasing dynamic loss scale from {prev_scale} to {self.cur_scale}")
        else:
            if skip:
                logger.info("Grad overflow on iteration %s", self.cur_iter)
                logger.info("Using static loss scale of %s", self.cur_scale)
        self.cur_iter += 1

    @property
    def state(self):
        return self.optimizer.state

    @state.setter
    def state(self, value):
        self.optimizer.state = value

    @property
    def param_groups(self):
        return self.optimizer.param_groups

    @param_groups.setter
    def param_groups(self, value):
        self.optimizer.param_groups = value

    @property
    def loss_scale(self):
        return self.cur_scale

    def state_dict(self):
        sd = {"dynamic_loss_scale": self.dynamic_loss_scale, "cur_scale": self.cur_scale, "cur_iter": self.cur_iter}
        if self.dynamic_loss_scale:
            sd["last_overflow_iter"] = self.last_overflow_iter
            sd["scale_factor"] = self.scale_factor
            sd["scale_window"] = self.scale_window
        sd["optimizer_state_dict"] = self.optimizer.state_dict()
        sd["fp32_groups"] = self.fp32_groups
        return sd

    def refresh_fp32_params(self):
        for current_group, saved_group in zip(self.fp32_groups, self.fp16_groups):
            for current, saved in zip(current_group, saved_group):
                current.data.copy_(saved.data)

    def load_state_dict(self, state_dict, load_optimizer_states=True):
        self.dynamic_loss_scale = state_dict["dynamic_loss_scale"]
        self.cur_scale = state_dict["cur_scale"]
        self.cur_iter = state_dict["cur_iter"]
        if state_dict["dynamic_loss_scale"]:
            self.last_overflow_iter = state_dict["last_overflow_iter"]
            self.scale_factor = state_dict["scale_factor"]
            self.scale_window = state_dict["scale_window"]
        if load_optimizer_states:
            self.optimizer.load_state_dict(state_dict["optimizer_state_dict"])
        for current_group, saved_group in zip(self.fp32_groups, state_dict["fp32_groups"]):
            for current, saved in zip(current_group, saved_group):
                current.data.copy_(saved.data)

    def __repr__(self):
        return repr(self.optimizer)

    def initialize_optimizer_states(self):
        """Run one zero-gradient step so state tensors exist before the first checkpoint load."""
        for i, group in enumerate(self.fp16_groups):
            for param in self.fp32_groups[i]:
                param.grad = torch.zeros(param.size(), dtype=param.dtype, device=param.device)
        saved = [[p.data.clone() for p in g] for g in self.fp32_groups]
        self.optimizer.step()
        for g, sg in zip(self.fp32_groups, saved):
            for p, s in zip(g, sg):
                p.data.copy_(s)
                p.grad = None
        # reset any step counters the dummy step advanced
        for st in self.optimizer.state.values():
            if isinstance(st, dict) and "step" in st and not torch.is_tensor(st["step"]):
                st["step"] = 0
