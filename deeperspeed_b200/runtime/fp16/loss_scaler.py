"""Static and dynamic loss scalers.

Contract (reference deepspeed/runtime/fp16/loss_scaler.py:32-168): class names, constructor keywords and attribute
names are fixed because these objects are pickled into ZeRO checkpoints (`deeperspeed_b200.compat` makes the
reference's module path resolve here). The schedule itself is written as a small state machine:

    overflow      -> spend one unit of hysteresis; when none is left, divide the scale (never below `min_scale`)
    clean step    -> `scale_window` clean steps in a row since the last overflow multiply the scale and refill the
                     hysteresis (with `consecutive_hysteresis` every clean step refills it)
"""
import math

import torch

INITIAL_LOSS_SCALE = "init_scale"
SCALE_WINDOW = "scale_window"
DELAYED_SHIFT = "delayed_shift"
MIN_LOSS_SCALE = "min_scale"


def to_python_float(t):
    return t.item() if hasattr(t, "item") else t[0]


class LossScalerBase:
    """Holds the current scale and applies it to a loss."""

    def __init__(self, cur_scale):
        self.cur_scale = cur_scale

    @property
    def loss_scale(self):
        return self.cur_scale

    def scale_gradient(self, module, grad_in, grad_out):
        """Backward-hook form: every incoming gradient times the scale."""
        s = self.loss_scale
        return tuple(g * s for g in grad_in)

    def update_scale(self, overflow):
        """Static scalers ignore overflow reports."""

    def backward(self, loss, retain_graph=False):
        (loss * self.loss_scale).backward(retain_graph=retain_graph)


class LossScaler(LossScalerBase):
    """A fixed scale; never reports an overflow."""

    def __init__(self, scale=1):
        super().__init__(scale)

    def has_overflow(self, params):
        return False

    @staticmethod
    def _has_inf_or_nan(x):
        return False


class DynamicLossScaler(LossScalerBase):
    """Scale that backs off on overflow and grows after `scale_window` clean iterations."""

    def __init__(self, init_scale=2 ** 32, scale_factor=2.0, scale_window=1000, min_scale=1, delayed_shift=1,
                 consecutive_hysteresis=False, raise_error_at_min_scale=True):
        super().__init__(init_scale)
        self.cur_iter = 0
        self.last_overflow_iter = -1
        self.scale_factor = scale_factor
        self.scale_window = scale_window
        self.min_scale = min_scale
        self.delayed_shift = delayed_shift
        self.cur_hysteresis = delayed_shift
        self.consecutive_hysteresis = consecutive_hysteresis
        self.raise_error_at_min_scale = raise_error_at_min_scale

    @staticmethod
    def _has_inf_or_nan(x):
        """True if the sum of `x` is not a finite number (one reduction instead of an element-wise scan)."""
        return not math.isfinite(float(x.float().sum()))

    def has_overflow_serial(self, params):
        return any(p.grad is not None and self._has_inf_or_nan(p.grad.data) for p in params)

    def _on_overflow(self):
        if self.delayed_shift != 1 and self.cur_hysteresis != 1:
            self.cur_hysteresis -= 1          # tolerated: hysteresis left
        else:
            if self.raise_error_at_min_scale and self.cur_scale == self.min_scale:
                raise Exception("Current loss scale already at minimum - cannot decrease scale anymore. Exiting run.")
            self.cur_scale = max(self.cur_scale / self.scale_factor, self.min_scale)
        self.last_overflow_iter = self.cur_iter

    def _on_clean_step(self):
        window_full = (self.cur_iter - self.last_overflow_iter) % self.scale_window == 0
        if self.consecutive_hysteresis or window_full:
            self.cur_hysteresis = self.delayed_shift
        if window_full:
            self.cur_scale *= self.scale_factor

    def update_scale(self, overflow):
        (self._on_overflow if overflow else self._on_clean_step)()
        self.cur_iter += 1


def CreateLossScaler(dtype, static_loss_scale, dynamic_scaling, dynamic_loss_args):
    """fp16 + dynamic -> DynamicLossScaler(**args); fp16 static -> LossScaler(scale); other dtypes need no scaling."""
    if dtype != torch.float16:
        return LossScaler(scale=1.0)
    if not dynamic_scaling:
        return LossScaler(scale=static_loss_scale)
    return DynamicLossScaler(**(dynamic_loss_args or {}))
