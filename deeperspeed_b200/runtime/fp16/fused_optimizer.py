"""FP16_Optimizer: flat low-precision arena + flat fp32 master per param group, dynamic/static loss scale with the
FP16_Optimizer update rule (no hysteresis), overflow skip, unscale+clip folded into the fused Adam kernel.
Protocol and state-dict keys per reference deepspeed/runtime/fp16/fused_optimizer.py:51-497."""
import torch

from ...utils.logging import logger
from .flat_optimizer import FlatMixedPrecisionOptimizer
from .loss_scaler import INITIAL_LOSS_SCALE, SCALE_WINDOW, MIN_LOSS_SCALE


# loss-scale attributes that travel in the state dict (the last three only with a dynamic scale)
_SCALE_KEYS = ("dynamic_loss_scale", "cur_scale", "cur_iter")
_DYNAMIC_KEYS = ("last_overflow_iter", "scale_factor", "scale_window")


class FP16_Optimizer(FlatMixedPrecisionOptimizer):
    def __init__(self, init_optimizer, static_loss_scale=1.0, dynamic_loss_scale=False, initial_dynamic_scale=2 ** 32,
                 dynamic_loss_args=None, verbose=True, mpu=None, clip_grad=0.0, fused_adam_legacy=False, timers=None,
                 dp_process_group=None, fp32_allreduce=False, deepspeed=None):
        self.deepspeed = deepspeed          # the engine, when it builds the wrapper (reference fused_optimizer.py:61)
        self.fused_adam_legacy = fused_adam_legacy
        self.timers = timers
        self.dynamic_loss_scale = bool(dynamic_loss_scale)
        self.cur_iter = 0
        self.cur_scale = static_loss_scale
        if self.dynamic_loss_scale:
            knobs = dynamic_loss_args or {INITIAL_LOSS_SCALE: initial_dynamic_scale, SCALE_WINDOW: 1000,
                                          MIN_LOSS_SCALE: 1}
            self.cur_scale = knobs[INITIAL_LOSS_SCALE]
            self.scale_window, self.min_loss_scale = knobs[SCALE_WINDOW], knobs[MIN_LOSS_SCALE]
            self.scale_factor, self.last_overflow_iter = 2, -1
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
        """FP16_Optimizer rule (no hysteresis): halve on overflow, double after `scale_window` clean iterations."""
        say = logger.info if self.verbose else (lambda *a: None)
        before = self.cur_scale
        if not self.dynamic_loss_scale:
            if skip:
                logger.info("Grad overflow on iteration: %s", self.cur_iter)
                logger.info("Using static loss scale of: %s", self.cur_scale)
        elif skip:
            self.cur_scale = max(before / self.scale_factor, self.min_loss_scale)
            self.last_overflow_iter = self.cur_iter
            say(f"\nGrad overflow on iteration {self.cur_iter}")
            say(f"Reducing dynamic loss scale from {before} to {self.cur_scale}")
        else:
            clean = self.cur_iter - self.last_overflow_iter - 1
            if clean > 0 and clean % self.scale_window == 0:
                self.cur_scale = before * self.scale_factor
                say(f"No Grad overflow for {self.scale_window} iterations")
                say(f"Increasing dynamic loss scale from {before} to {self.cur_scale}")
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

    def _scale_keys(self, dynamic):
        return _SCALE_KEYS + (_DYNAMIC_KEYS if dynamic else ())

    def state_dict(self):
        sd = {key: getattr(self, key) for key in self._scale_keys(self.dynamic_loss_scale)}
        sd["optimizer_state_dict"] = self.optimizer.state_dict()
        # masters are saved without the arena's alignment padding, so the file does not depend on the layout
        sd["fp32_groups_flat"] = [a.to_unpadded(m.data) if a.numel_used != a.numel_unpadded else
                                  m.data.narrow(0, 0, a.numel_unpadded).clone()
                                  for a, m in zip(self.arenas, self.fp32_masters)]
        sd["clip_grad"] = self.clip_grad
        return sd

    def refresh_fp32_params(self):
        super().refresh_fp32_params()

    def load_state_dict(self, state_dict, load_optimizer_states=True):
        for key in self._scale_keys(state_dict["dynamic_loss_scale"]):
            setattr(self, key, state_dict[key])
        self.clip_grad = state_dict["clip_grad"]
        if load_optimizer_states:
            self._load_inner_state(state_dict["optimizer_state_dict"])
        for arena, master, saved in zip(self.arenas, self.fp32_masters, state_dict["fp32_groups_flat"]):
            saved = saved.to(master.device)
            if saved.numel() == master.numel():
                master.data.copy_(saved)
            else:
                arena.from_unpadded(saved.float(), out=master.data)

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


def _fused_supported_classes():
    from ...ops.adam import FusedAdam
    return [FusedAdam, torch.optim.Adam, torch.optim.AdamW]


def __getattr__(name):
    # FP16_FUSED_SUPPORTED_OPTIMIZERS (reference fp16/fused_optimizer.py:22): resolved on first use, because
    # ops.adam imports this package's siblings
    if name == "FP16_FUSED_SUPPORTED_OPTIMIZERS":
        return _fused_supported_classes()
    raise AttributeError(name)


def is_fp16_fused_supported_optimizer(optimizer):
    """Is `optimizer` one FP16_Optimizer drives through its fused path (reference fp16/fused_optimizer.py:35-51)?"""
    if isinstance(optimizer, tuple(_fused_supported_classes())):
        return True
    return optimizer.__class__.__name__.lower() == "onebitadam"
