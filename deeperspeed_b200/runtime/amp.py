"""`"amp": {"enabled": true, ...}` without apex.

The reference hands model and optimizer to `apex.amp.initialize` and scales the loss with `amp.scale_loss`
(deepspeed/runtime/engine.py:682-693, 1085-1095, 1146-1153). Here the same config block drives PyTorch's own mixed
precision: the forward pass runs under `torch.autocast`, parameters and gradients stay fp32, and a `torch.amp.GradScaler`
does the dynamic loss scaling (unscale before clipping, skip the step on overflow). apex `opt_level`s map as
  O0 -> no autocast (fp32), O1 / O2 -> autocast to fp16, O3 -> autocast to fp16 as well (pure-fp16 weights are what the
  `fp16` block is for); `"cast_model_type"` / a `"dtype"` key of "bfloat16" selects bf16 (no loss scaling needed).
`loss_scale`: "dynamic" (default) or a number.
"""
import torch


class NativeAmp:
    def __init__(self, params, device):
        params = dict(params or {})
        level = str(params.get("opt_level", "O1")).upper()
        wants_bf16 = "bf" in str(params.get("dtype", params.get("cast_model_type", ""))).lower()
        self.dtype = torch.bfloat16 if wants_bf16 else torch.float16
        self.device_type = device.type if isinstance(device, torch.device) else str(device).split(":")[0]
        if self.device_type == "cpu" and self.dtype == torch.float16:
            self.dtype = torch.bfloat16          # CPU autocast is a bf16 feature
        self.autocast_on = level != "O0"
        scale = params.get("loss_scale", "dynamic")
        dynamic = scale in (None, "dynamic") and self.dtype == torch.float16
        static = 1.0 if self.dtype == torch.bfloat16 or scale in (None, "dynamic") else float(scale)
        self.scaler = torch.amp.GradScaler(self.device_type, enabled=self.autocast_on and (dynamic or static != 1.0),
                                           init_scale=2.0 ** 16 if dynamic else static,
                                           growth_interval=2000 if dynamic else 2 ** 62)
        self.skipped = False

    def autocast(self):
        return torch.autocast(self.device_type, dtype=self.dtype, enabled=self.autocast_on)

    def backward(self, loss):
        self.scaler.scale(loss).backward()

    def step(self, optimizer, parameters, max_norm=0.0):
        """Unscale, clip (on true gradients), step unless an overflow was found, update the scale."""
        self.scaler.unscale_(optimizer)
        if max_norm > 0.0:
            torch.nn.utils.clip_grad_norm_(parameters, max_norm)
        before = self.scaler.get_scale()
        self.scaler.step(optimizer)
        self.scaler.update()
        self.skipped = self.scaler.is_enabled() and self.scaler.get_scale() < before
        return self.skipped

    @property
    def loss_scale(self):
        return self.scaler.get_scale() if self.scaler.is_enabled() else 1.0

    def state_dict(self):
        return self.scaler.state_dict()

    def load_state_dict(self, sd):
        if sd:
            self.scaler.load_state_dict(sd)
