"""FP16_Optimizer: flat low-precision arena + flat fp32 master per param group, dynamic/static loss scale with the
FP16_Optimizer update rule (no hysteresis), overflow skip, unscale+clip folded into the fused Adam kernel.
Protocol and state-dict keys per reference deepspeed/runtime/fp16/fused_optimizer.py:51-497."""
import torch

from ...utils.logging import logger
from .flat_optimizer import FlatMixedPrecisionOptimizer
from .loss_scaler import INITIAL_LOSS_SCALE, SCALE_WINDOW, MIN_LOSS_SCALE


class FP16_Optimizer(FlatMixedPrecisionOptimizer):
    def __init__(self, init_optimizer, static_loss_scale=1.0, dynamic_loss_scale=False, initial_dynamic_scale=2 ** 32,
                 dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0, fused_adam_legacy=False, timers=None,
                 dp_process_group=None, fp32_allreduce=False):
        self.fused_adam_legacy = fused_adam_legacy
        self.timers = timers
        if dynamic_loss_scale:
            self.dynamic_loss_scale = True
            self.cur_iter = 0
            self.last_overflow_iter = -1
            self.scale_factor = 2
            if dynamic_loss_args is None:
                self.cur_scale = initial_dynamic_scale
                self.scale_window = 1000
                self.min_loss_scale = 1
            else:
                self.cur_scale = dynamic_loss_args[INITIAL_LOSS_SCALE]
                self.scale_window = dynamic_loss_args[SCALE_WINDOW]
                self.min_loss_scale = dynamic_loss_args[MIN_LOSS_SCALE]
        else:
            self.dynamic_loss_scale = False
            self.cur_iter = 0
            self.cur_scale = static_loss_scale
        super().__init__(init_optimizer, dp_process_group=dp_process_group, shard=False, mpu=mpu, clip_grad=clip_grad,
                         fp32_allreduce=fp32_allreduce, verbose=verbose)
        # reference attribute names
        self.fp16_groups = self.group_params
        self.fp16_groups_flat = [a.data for a in self.arenas]
        self.fp32_groups_flat = self.fp32_masters

    def _loss_scale(self):
        return self.cur_scale

    @property
    def loss_scale(self):
        return self.cur_scale

    def _on_step_result(self, overflow):
        self._update_scale(overflow)

    def _update_scale(self, skip):
        if self.dynamic_loss_scale:
            prev_scale = self.cur_scale
            if skip:
                self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_loss_scale)
                self.last_overflow_iter = self.cur_iter
                if self.verbose:
                    logger.info(f"\nGrad overflow on iteration {self.cur_iter}")
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
                logger.info("Grad overflow on iteration: %s", self.cur_iter)
                logger.info("Using static loss scale of: %s", self.cur_scale)
        self.cur_iter += 1

    def step_fused_adam(self, closure=None):
        return self.step(closure)

    def unscale_and_clip_grads(self, grad_groups_flat, norm_groups, apply_scale=True):
        """Reference helper kept for API parity: returns the combined scale."""
        total_norm = sum(n ** 2.0 for n in norm_groups) ** 0.5
        combined_scale = self.cur_scale
        if self.clip_grad > 0.0:
            clip = ((total_norm / self.cur_scale) + 1e-6) / self.clip_grad
            if clip > 1:
                combined_scale = clip * self.cur_scale
        if apply_scale:
            for grad in grad_groups_flat:
                grad.data.mul_(1.0 / combined_scale)
        return combined_scale

    def state_dict(self):
        sd = {
            "dynamic_loss_scale": self.dynamic_loss_scale,
            "cur_scale": self.cur_scale,
            "cur_iter": self.cur_iter,
        }
        if self.dynamic_loss_scale:
            sd["last_overflow_iter"] = self.last_overflow_iter
            sd["scale_factor"] = self.scale_factor
            sd["scale_window"] = self.scale_window
        sd["optimizer_state_dict"] = self.optimizer.state_dict()
        sd["fp32_groups_flat"] = [a.to_unpadded(m.data) if a.numel_used != a.numel_unpadded else
                                  m.data.narrow(0, 0, a.numel_unpadded).clone()
                                  for a, m in zip(self.arenas, self.fp32_masters)]
        sd["clip_grad"] = self.clip_grad
        return sd

    def refresh_fp32_params(self):
        super().refresh_fp32_params()

    def load_state_dict(self, state_dict, load_optimizer_states=True):
        self.dynamic_loss_scale = state_dict["dynamic_loss_scale"]
        self.cur_scale = state_dict["cur_scale"]
        self.cur_iter = state_dict["cur_iter"]
        if state_dict["dynamic_loss_scale"]:
            self.last_overflow_iter = state_dict["last_overflow_iter"]
            self.scale_factor = state_dict["scale_factor"]
            self.scale_window = state_dict["scale_window"]
        if load_optimizer_states:
            self._load_inner_state(state_dict["optimizer_state_dict"])
        self.clip_grad = state_dict["clip_grad"]
        for a, cur, saved in zip(self.arenas, self.fp32_masters, state_dict["fp32_groups_flat"]):
            saved = saved.to(cur.device)
            if saved.numel() == cur.numel():
                cur.data.copy_(saved)
            else:
                a.from_unpadded(saved.float(), out=cur.data)

    def _load_inner_state(self, osd):
        """Inner state tensors are saved unpadded-flat per group; pad them to the arena layout on load."""
        import copy
        osd = copy.deepcopy(osd)
        for gi, (a, m) in enumerate(zip(self.arenas, self.fp32_masters)):
            pid = osd["param_groups"][gi]["params"][0]
            st = osd["state"].get(pid, {})
            for k, v in list(st.items()):
                if torch.is_tensor(v) and v.dim() == 1 and v.numel() == a.numel_unpadded and v.numel() != m.numel():
                    st[k] = a.from_unpadded(v.float().to(m.device))
        self.optimizer.load_state_dict(osd)

    def __repr__(self):
        return repr(self.optimizer)


def is_fp16_fused_supported_optimizer(optimizer):
    """Is `optimizer` one FP16_Optimizer drives through its fused path (reference fp16/fused_optimizer.py:35-51)?"""
    from ...ops.adam import FusedAdam
    if isinstance(optimizer, (FusedAdam, torch.optim.Adam, torch.optim.AdamW)):
        return True
    return optimizer.__class__.__name__.lower() == "onebitadam"
