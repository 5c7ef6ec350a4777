"""LR schedules: LRRangeTest, OneCycle, WarmupLR, WarmupDecayLR + CLI tuning arguments.

Same names, constructor arguments, config keys and formulas as the reference (deepspeed/runtime/lr_schedules.py,
SURVEY C.10). Organisation is ours: every schedule is a closed-form function of the iteration count (`lr_at` /
`mom_at`), the stepping / state-dict / param-group plumbing lives once in `_Schedule`, and the command-line surface is
generated from one table.
"""
import argparse
import math

from torch.optim import Optimizer

from ..utils.logging import logger

LR_SCHEDULE = "lr_schedule"
LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR = "LRRangeTest", "OneCycle", "WarmupLR", "WarmupDecayLR"
VALID_LR_SCHEDULES = [LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR]

# (flag / config key, type, default, help, schedule it belongs to). `cycle_momentum` is a switch, listed separately.
_TUNING = [
    ("lr_range_test_min_lr", float, 0.001, "Starting lr value.", LR_RANGE_TEST),
    ("lr_range_test_step_rate", float, 1.0, "scaling rate for LR range test.", LR_RANGE_TEST),
    ("lr_range_test_step_size", int, 1000, "training steps per LR change.", LR_RANGE_TEST),
    ("lr_range_test_staircase", bool, False, "use staircase scaling.", LR_RANGE_TEST),
    ("cycle_first_step_size", int, 1000, "size of first step of 1Cycle schedule.", ONE_CYCLE),
    ("cycle_first_stair_count", int, -1, "first stair count for 1Cycle schedule.", ONE_CYCLE),
    ("cycle_second_step_size", int, -1, "size of second step of 1Cycle schedule.", ONE_CYCLE),
    ("cycle_second_stair_count", int, -1, "second stair count for 1Cycle.", ONE_CYCLE),
    ("decay_step_size", int, 1000, "size of intervals for applying post cycle decay.", ONE_CYCLE),
    ("cycle_min_lr", float, 0.01, "1Cycle LR lower bound.", ONE_CYCLE),
    ("cycle_max_lr", float, 0.1, "1Cycle LR upper bound.", ONE_CYCLE),
    ("decay_lr_rate", float, 0.0, "post cycle LR decay rate.", ONE_CYCLE),
    ("cycle_min_mom", float, 0.8, "1Cycle momentum lower bound.", ONE_CYCLE),
    ("cycle_max_mom", float, 0.9, "1Cycle momentum upper bound.", ONE_CYCLE),
    ("decay_mom_rate", float, 0.0, "post cycle momentum decay rate.", ONE_CYCLE),
    ("warmup_min_lr", float, 0, "WarmupLR minimum/initial LR value", WARMUP_LR),
    ("warmup_max_lr", float, 0.001, "WarmupLR maximum LR value.", WARMUP_LR),
    ("warmup_num_steps", int, 1000, "WarmupLR step count for LR warmup.", WARMUP_LR),
]
# the reference's key constants (LR_RANGE_TEST_MIN_LR = "lr_range_test_min_lr", ...)
globals().update({key.upper(): key for key, *_ in _TUNING})
EDGE_VALUE, MID_VALUE, TOTAL_NUM_STEPS = "edge_value", "mid_value", "total_num_steps"
_KEYS_OF = {kind: [key for key, *_, owner in _TUNING if owner == kind] for kind in (LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR)}
_KEYS_OF[WARMUP_DECAY_LR] = _KEYS_OF[WARMUP_LR]
# the learning rate a schedule starts the optimizer from / peaks at, as reported by get_lr_from_config
_HEADLINE_LR = {LR_RANGE_TEST: "lr_range_test_min_lr", ONE_CYCLE: "cycle_max_lr", WARMUP_LR: "warmup_max_lr",
                WARMUP_DECAY_LR: "warmup_max_lr"}


# ------------------------------------------------------------------------------------------------ command line
def add_tuning_arguments(parser):
    group = parser.add_argument_group("Convergence Tuning", "Convergence tuning configurations")
    group.add_argument("--lr_schedule", type=str, default=None, help="LR schedule for training.")
    for key, kind, default, text, _ in _TUNING:
        group.add_argument("--" + key, type=kind, default=default, help=text)
        if key == "decay_lr_rate":      # keeps the reference's flag order in --help
            group.add_argument("--cycle_momentum", default=False, action="store_true",
                               help="Enable 1Cycle momentum schedule.")
    return parser


def parse_arguments():
    return add_tuning_arguments(argparse.ArgumentParser()).parse_known_args()


def override_params(args, params, keys):
    """Copy the attributes of `args` named in `keys` that are set (not None) into `params`."""
    params.update({k: getattr(args, k) for k in keys if getattr(args, k, None) is not None})


def override_lr_range_test_params(args, params):
    override_params(args, params, _KEYS_OF[LR_RANGE_TEST])


def override_1cycle_params(args, params):
    override_params(args, params, _KEYS_OF[ONE_CYCLE])


def override_warmupLR_params(args, params):
    override_params(args, params, _KEYS_OF[WARMUP_LR])


def get_config_from_args(args):
    """({"type", "params"}, None) built from parsed tuning arguments, or (None, reason)."""
    kind = getattr(args, LR_SCHEDULE, None)
    if kind is None:
        return None, "--{} not specified on command line".format(LR_SCHEDULE)
    if kind not in VALID_LR_SCHEDULES:
        return None, "{} is not supported LR schedule".format(kind)
    config = {"type": kind, "params": {}}
    override_params(args, config["params"], _KEYS_OF[kind])
    return config, None


def get_lr_from_config(config):
    """(learning rate the optimizer should be built with, "") or (None, reason)."""
    for field in ("type", "params"):
        if field not in config:
            return None, "LR schedule {} not defined in config".format(field)
    if config["type"] not in VALID_LR_SCHEDULES:
        return None, "{} is not a valid LR schedule".format(config["type"])
    return config["params"][_HEADLINE_LR[config["type"]]], ""


def get_torch_optimizer(optimizer):
    """Unwrap an FP16/ZeRO wrapper down to the torch.optim.Optimizer holding param_groups."""
    for candidate in (optimizer, getattr(optimizer, "optimizer", None)):
        if isinstance(candidate, Optimizer):
            return candidate
    raise TypeError("{} is not a subclass of torch.optim.Optimizer".format(type(optimizer).__name__))


def _per_group(optimizer, value, name):
    """One value per param group: a scalar is repeated, a list must have the right length."""
    groups = len(optimizer.param_groups)
    if not isinstance(value, (list, tuple)):
        return [value] * groups
    if len(value) != groups:
        raise ValueError("expected {} value for {}, got {}".format(groups, name, value))
    return list(value)


# ------------------------------------------------------------------------------------------------ schedules
class _Schedule:
    """step() bookkeeping shared by all schedules. Subclasses give `get_lr()` (one value per param group) for the
    current `last_batch_iteration`, optionally `get_mom()`."""

    def _bind(self, optimizer, last_batch_iteration):
        self.optimizer = get_torch_optimizer(optimizer)
        self.last_batch_iteration = last_batch_iteration

    def _set(self, field, values):
        for group, value in zip(self.optimizer.param_groups, values):
            group[field] = value

    def step(self, batch_iteration=None):
        self.last_batch_iteration = self.last_batch_iteration + 1 if batch_iteration is None else batch_iteration
        self._set("lr", self.get_lr())
        self._last_lr = [group["lr"] for group in self.optimizer.param_groups]
        self._after_step()

    def _after_step(self):
        pass

    def get_last_lr(self):
        assert getattr(self, "_last_lr", None) is not None, "need to call step() first"
        return self._last_lr

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]


class LRRangeTest(_Schedule):
    """lr = min_lr * (1 + step_rate * interval), interval = (it + 1) / step_size (floored if staircase)."""

    def __init__(self, optimizer, lr_range_test_min_lr=1e-3, lr_range_test_step_size=2000,
                 lr_range_test_step_rate=1.0, lr_range_test_staircase=False, last_batch_iteration=-1):
        self._bind(optimizer, last_batch_iteration)
        self.min_lr = _per_group(self.optimizer, lr_range_test_min_lr, "lr_range_test_min_lr")
        self.step_size, self.step_rate, self.staircase = lr_range_test_step_size, lr_range_test_step_rate, \
            lr_range_test_staircase
        if last_batch_iteration == -1:
            self._set("lr", self.min_lr)

    def _interval(self):
        ratio = float(self.last_batch_iteration + 1) / self.step_size
        return math.floor(ratio) if self.staircase else ratio

    def get_lr(self):
        growth = 1 + self.step_rate * self._interval()
        return [lr * growth for lr in self.min_lr]


class OneCycle(_Schedule):
    """1Cycle: lr min->max->min over (first + second) steps, momentum the inverse, then decay."""

    def __init__(self, optimizer, cycle_min_lr, cycle_max_lr, decay_lr_rate=0.0, cycle_first_step_size=2000,
                 cycle_second_step_size=None, cycle_first_stair_count=0, cycle_second_stair_count=None,
                 decay_step_size=0, cycle_momentum=True, cycle_min_mom=0.8, cycle_max_mom=0.9, decay_mom_rate=0.0,
                 last_batch_iteration=-1):
        self._bind(optimizer, last_batch_iteration)
        groups = len(self.optimizer.param_groups)
        up = float(cycle_first_step_size)
        down = up if cycle_second_step_size is None else float(cycle_second_step_size)
        self.total_size, self.step_ratio = up + down, up / (up + down)
        self.first_stair_count = cycle_first_stair_count
        self.second_stair_count = cycle_first_stair_count if cycle_second_stair_count is None else cycle_second_stair_count
        self.decay_step_size, self.decay_lr_rate = decay_step_size, decay_lr_rate
        self.min_lrs, self.max_lrs = [cycle_min_lr] * groups, [cycle_max_lr] * groups
        fresh = last_batch_iteration == -1
        if fresh:
            self._set("lr", self.min_lrs)
        self.cycle_momentum = bool(cycle_momentum)
        if self.cycle_momentum and "betas" not in self.optimizer.defaults:
            logger.warning("cycle_momentum is disabled because optimizer {} does not support momentum".format(
                type(self.optimizer).__name__))
            self.cycle_momentum = False
        if self.cycle_momentum:
            self.decay_mom_rate = decay_mom_rate
            self.min_moms, self.max_moms = [(cycle_min_mom, 0.99)] * groups, [(cycle_max_mom, 0.99)] * groups
            if fresh:
                self._set("betas", self.min_moms)

    def _position(self):
        """0 at the ends of a cycle, 1 at its peak (piecewise linear in the iteration)."""
        x = (self.last_batch_iteration + 1) / self.total_size
        x = x - math.floor(x)                  # phase inside the current cycle, in [0, 1)
        if x <= self.step_ratio:
            return x / self.step_ratio
        return (x - 1) / (self.step_ratio - 1)

    def _decay_interval(self):
        # decay_step_size 0 is documented as "no decay" (the reference divides by it, lr_schedules.py:609,619): the
        # schedule then rests at cycle_min_lr / cycle_max_mom once the cycle is over
        if not self.decay_step_size:
            return 0.0
        return (self.last_batch_iteration - self.total_size + 1) / self.decay_step_size

    def _in_cycle(self):
        return self.last_batch_iteration < self.total_size

    def get_lr(self):
        if self._in_cycle():
            s = self._position()
            return [lo + (hi - lo) * s for lo, hi in zip(self.min_lrs, self.max_lrs)]
        shrink = 1 + self.decay_lr_rate * self._decay_interval()
        return [lo / shrink for lo in self.min_lrs]

    def get_mom(self):
        if not self.cycle_momentum:
            return None
        if self._in_cycle():
            s = self._position()
            return [(hi[0] - (hi[0] - lo[0]) * s, lo[1]) for lo, hi in zip(self.min_moms, self.max_moms)]
        grow = 1 + self.decay_mom_rate * self._decay_interval()
        return [(beta1 * grow, beta2) for beta1, beta2 in self.max_moms]

    def _after_step(self):
        if self.cycle_momentum:
            self._set("betas", self.get_mom())


class WarmupLR(_Schedule):
    """lr = min + (max - min) * log(it + 1) / log(warmup_steps) during warm-up, then max."""

    def __init__(self, optimizer, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 last_batch_iteration=-1):
        self._bind(optimizer, last_batch_iteration)
        self.min_lrs = _per_group(self.optimizer, warmup_min_lr, "min_lr")
        self.max_lrs = _per_group(self.optimizer, warmup_max_lr, "max_lr")
        self.delta_lrs = [hi - lo for hi, lo in zip(self.max_lrs, self.min_lrs)]
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)

    def _after_warmup(self):
        return 1.0

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
        return self._after_warmup()

    def get_lr(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return [0.0]
        gamma = self._get_gamma()
        return [lo + delta * gamma for lo, delta in zip(self.min_lrs, self.delta_lrs)]

    def step(self, last_batch_iteration=None):
        super().step(last_batch_iteration)


class WarmupDecayLR(WarmupLR):
    """WarmupLR followed by linear decay to 0 at total_num_steps."""

    def __init__(self, optimizer, total_num_steps, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 last_batch_iteration=-1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, last_batch_iteration)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning("total_num_steps {} is less than warmup_num_steps {}".format(total_num_steps,
                                                                                        warmup_num_steps))

    def _after_warmup(self):
        remaining = float(self.total_num_steps - self.last_batch_iteration)
        return max(0.0, remaining / float(max(1.0, self.total_num_steps - self.warmup_num_steps)))
