"""LR schedules: LRRangeTest, OneCycle, WarmupLR, WarmupDecayLR + CLI tuning arguments.
Formulas per reference deepspeed/runtime/lr_schedules.py:301-809."""
import argparse
import math

from torch.optim import Optimizer

from ..utils.logging import logger

LR_SCHEDULE = "lr_schedule"
LR_RANGE_TEST = "LRRangeTest"
ONE_CYCLE = "OneCycle"
WARMUP_LR = "WarmupLR"
WARMUP_DECAY_LR = "WarmupDecayLR"
VALID_LR_SCHEDULES = [LR_RANGE_TEST, ONE_CYCLE, WARMUP_LR, WARMUP_DECAY_LR]

LR_RANGE_TEST_MIN_LR = "lr_range_test_min_lr"
LR_RANGE_TEST_STEP_RATE = "lr_range_test_step_rate"
LR_RANGE_TEST_STEP_SIZE = "lr_range_test_step_size"
LR_RANGE_TEST_STAIRCASE = "lr_range_test_staircase"
EDGE_VALUE = "edge_value"
MID_VALUE = "mid_value"
CYCLE_FIRST_STEP_SIZE = "cycle_first_step_size"
CYCLE_FIRST_STAIR_COUNT = "cycle_first_stair_count"
CYCLE_SECOND_STEP_SIZE = "cycle_second_step_size"
CYCLE_SECOND_STAIR_COUNT = "cycle_second_stair_count"
DECAY_STEP_SIZE = "decay_step_size"
CYCLE_MIN_LR = "cycle_min_lr"
CYCLE_MAX_LR = "cycle_max_lr"
DECAY_LR_RATE = "decay_lr_rate"
CYCLE_MIN_MOM = "cycle_min_mom"
CYCLE_MAX_MOM = "cycle_max_mom"
DECAY_MOM_RATE = "decay_mom_rate"
WARMUP_MIN_LR = "warmup_min_lr"
WARMUP_MAX_LR = "warmup_max_lr"
WARMUP_NUM_STEPS = "warmup_num_steps"
TOTAL_NUM_STEPS = "total_num_steps"


def add_tuning_arguments(parser):
    group = parser.add_argument_group("Convergence Tuning", "Convergence tuning configurations")
    group.add_argument("--lr_schedule", type=str, default=None, help="LR schedule for training.")
    group.add_argument("--lr_range_test_min_lr", type=float, default=0.001, help="Starting lr value.")
    group.add_argument("--lr_range_test_step_rate", type=float, default=1.0, help="scaling rate for LR range test.")
    group.add_argument("--lr_range_test_step_size", type=int, default=1000, help="training steps per LR change.")
    group.add_argument("--lr_range_test_staircase", type=bool, default=False, help="use staircase scaling.")
    group.add_argument("--cycle_first_step_size", type=int, default=1000, help="size of first step of 1Cycle schedule.")
    group.add_argument("--cycle_first_stair_count", type=int, default=-1, help="first stair count for 1Cycle schedule.")
    group.add_argument("--cycle_second_step_size", type=int, default=-1, help="size of second step of 1Cycle schedule.")
    group.add_argument("--cycle_second_stair_count", type=int, default=-1, help="second stair count for 1Cycle.")
    group.add_argument("--decay_step_size", type=int, default=1000, help="size of intervals for applying post cycle decay.")
    group.add_argument("--cycle_min_lr", type=float, default=0.01, help="1Cycle LR lower bound.")
    group.add_argument("--cycle_max_lr", type=float, default=0.1, help="1Cycle LR upper bound.")
    group.add_argument("--decay_lr_rate", type=float, default=0.0, help="post cycle LR decay rate.")
    group.add_argument("--cycle_momentum", default=False, action="store_true", help="Enable 1Cycle momentum schedule.")
    group.add_argument("--cycle_min_mom", type=float, default=0.8, help="1Cycle momentum lower bound.")
    group.add_argument("--cycle_max_mom", type=float, default=0.9, help="1Cycle momentum upper bound.")
    group.add_argument("--decay_mom_rate", type=float, default=0.0, help="post cycle momentum decay rate.")
    group.add_argument("--warmup_min_lr", type=float, default=0, help="WarmupLR minimum/initial LR value")
    group.add_argument("--warmup_max_lr", type=float, default=0.001, help="WarmupLR maximum LR value.")
    group.add_argument("--warmup_num_steps", type=int, default=1000, help="WarmupLR step count for LR warmup.")
    return parser


def parse_arguments():
    parser = argparse.ArgumentParser()
    parser = add_tuning_arguments(parser)
    lr_sched_args, unknown_args = parser.parse_known_args()
    return lr_sched_args, unknown_args


def override_params(args, params, keys):
    for k in keys:
        if hasattr(args, k) and getattr(args, k) is not None:
            params[k] = getattr(args, k)


def override_lr_range_test_params(args, params):
    override_params(args, params, [LR_RANGE_TEST_MIN_LR, LR_RANGE_TEST_STEP_RATE, LR_RANGE_TEST_STEP_SIZE,
                                   LR_RANGE_TEST_STAIRCASE])


def override_1cycle_params(args, params):
    override_params(args, params, [CYCLE_FIRST_STEP_SIZE, CYCLE_FIRST_STAIR_COUNT, CYCLE_SECOND_STEP_SIZE,
                                   CYCLE_SECOND_STAIR_COUNT, DECAY_STEP_SIZE, CYCLE_MIN_LR, CYCLE_MAX_LR, DECAY_LR_RATE,
                                   CYCLE_MIN_MOM, CYCLE_MAX_MOM, DECAY_MOM_RATE])


def override_warmupLR_params(args, params):
    override_params(args, params, [WARMUP_MIN_LR, WARMUP_MAX_LR, WARMUP_NUM_STEPS])


def get_config_from_args(args):
    if not hasattr(args, LR_SCHEDULE) or args.lr_schedule is None:
        return None, "--{} not specified on command line".format(LR_SCHEDULE)
    if args.lr_schedule not in VALID_LR_SCHEDULES:
        return None, "{} is not supported LR schedule".format(args.lr_schedule)
    config = {"type": args.lr_schedule, "params": {}}
    if args.lr_schedule == LR_RANGE_TEST:
        override_params(args, config["params"], [LR_RANGE_TEST_MIN_LR, LR_RANGE_TEST_STEP_RATE,
                                                 LR_RANGE_TEST_STEP_SIZE, LR_RANGE_TEST_STAIRCASE])
    elif args.lr_schedule == ONE_CYCLE:
        override_params(args, config["params"], [CYCLE_FIRST_STEP_SIZE, CYCLE_FIRST_STAIR_COUNT, CYCLE_SECOND_STEP_SIZE,
                                                 CYCLE_SECOND_STAIR_COUNT, DECAY_STEP_SIZE, CYCLE_MIN_LR, CYCLE_MAX_LR,
                                                 DECAY_LR_RATE, CYCLE_MIN_MOM, CYCLE_MAX_MOM, DECAY_MOM_RATE])
    else:
        override_params(args, config["params"], [WARMUP_MIN_LR, WARMUP_MAX_LR, WARMUP_NUM_STEPS])
    return config, None


def get_lr_from_config(config):
    if "type" not in config:
        return None, "LR schedule type not defined in config"
    if "params" not in config:
        return None, "LR schedule params not defined in config"
    t, p = config["type"], config["params"]
    if t not in VALID_LR_SCHEDULES:
        return None, "{} is not a valid LR schedule".format(t)
    if t == LR_RANGE_TEST:
        return p[LR_RANGE_TEST_MIN_LR], ""
    if t == ONE_CYCLE:
        return p[CYCLE_MAX_LR], ""
    return p[WARMUP_MAX_LR], ""


def get_torch_optimizer(optimizer):
    """Unwrap an FP16/ZeRO wrapper down to the torch.optim.Optimizer holding param_groups."""
    if isinstance(optimizer, Optimizer):
        return optimizer
    if hasattr(optimizer, "optimizer") and isinstance(optimizer.optimizer, Optimizer):
        return optimizer.optimizer
    raise TypeError("{} is not a subclass of torch.optim.Optimizer".format(type(optimizer).__name__))


class _Schedule:
    def get_last_lr(self):
        assert getattr(self, "_last_lr", None) is not None, "need to call step() first"
        return self._last_lr

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]

    def _apply(self):
        for group, lr in zip(self.optimizer.param_groups, self.get_lr()):
            group["lr"] = lr
        self._last_lr = [g["lr"] for g in self.optimizer.param_groups]


def _format_param(optimizer, value, name):
    if isinstance(value, (list, tuple)):
        if len(value) != len(optimizer.param_groups):
            raise ValueError("expected {} value for {}, got {}".format(len(optimizer.param_groups), name, value))
        return list(value)
    return [value] * len(optimizer.param_groups)


class LRRangeTest(_Schedule):
    """lr = min_lr * (1 + step_rate * interval), interval = (it + 1) / step_size (floored if staircase)."""

    def __init__(self, optimizer, lr_range_test_min_lr=1e-3, lr_range_test_step_size=2000,
                 lr_range_test_step_rate=1.0, lr_range_test_staircase=False, last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.min_lr = _format_param(self.optimizer, lr_range_test_min_lr, "lr_range_test_min_lr")
        self.step_size = lr_range_test_step_size
        self.step_rate = lr_range_test_step_rate
        self.last_batch_iteration = last_batch_iteration
        self.staircase = lr_range_test_staircase
        self.interval_fn = self._staircase_interval if lr_range_test_staircase else self._continuous_interval
        if last_batch_iteration == -1:
            for g, lr in zip(self.optimizer.param_groups, self.min_lr):
                g["lr"] = lr

    def _staircase_interval(self):
        return math.floor(float(self.last_batch_iteration + 1) / self.step_size)

    def _continuous_interval(self):
        return float(self.last_batch_iteration + 1) / self.step_size

    def _get_increase(self):
        return 1 + self.step_rate * self.interval_fn()

    def get_lr(self):
        inc = self._get_increase()
        return [m * inc for m in self.min_lr]

    def step(self, batch_iteration=None):
        if batch_iteration is None:
            batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = batch_iteration
        self._apply()


class OneCycle(_Schedule):
    """1Cycle: lr min->max->min over (first + second) steps, momentum the inverse, then decay."""

    def __init__(self, optimizer, cycle_min_lr, cycle_max_lr, decay_lr_rate=0.0, cycle_first_step_size=2000,
                 cycle_second_step_size=None, cycle_first_stair_count=0, cycle_second_stair_count=None,
                 decay_step_size=0, cycle_momentum=True, cycle_min_mom=0.8, cycle_max_mom=0.9, decay_mom_rate=0.0,
                 last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        first = float(cycle_first_step_size)
        second = float(cycle_second_step_size) if cycle_second_step_size is not None else first
        self.total_size = first + second
        self.step_ratio = first / self.total_size
        self.first_stair_count = cycle_first_stair_count
        self.second_stair_count = cycle_first_stair_count if cycle_second_stair_count is None else cycle_second_stair_count
        self.decay_step_size = decay_step_size
        self.min_lrs = [cycle_min_lr] * len(self.optimizer.param_groups)
        self.max_lrs = [cycle_max_lr] * len(self.optimizer.param_groups)
        self.decay_lr_rate = decay_lr_rate
        if last_batch_iteration == -1:
            for lr, g in zip(self.min_lrs, self.optimizer.param_groups):
                g["lr"] = lr
        self.cycle_momentum = cycle_momentum
        if cycle_momentum:
            if "betas" not in self.optimizer.defaults:
                logger.warning("cycle_momentum is disabled because optimizer {} does not support momentum".format(
                    type(self.optimizer).__name__))
                self.cycle_momentum = False
            else:
                self.decay_mom_rate = decay_mom_rate
                self.min_moms = [(cycle_min_mom, 0.99)] * len(self.optimizer.param_groups)
                self.max_moms = [(cycle_max_mom, 0.99)] * len(self.optimizer.param_groups)
                if last_batch_iteration == -1:
                    for m, g in zip(self.min_moms, self.optimizer.param_groups):
                        g["betas"] = m
        self.last_batch_iteration = last_batch_iteration

    def _get_scale_factor(self):
        it = self.last_batch_iteration + 1
        cycle = math.floor(1 + it / self.total_size)
        x = 1.0 + it / self.total_size - cycle
        return x / self.step_ratio if x <= self.step_ratio else (x - 1) / (self.step_ratio - 1)

    def _get_cycle_lr(self):
        sf = self._get_scale_factor()
        return [lo + (hi - lo) * sf for lo, hi in zip(self.min_lrs, self.max_lrs)]

    def _get_cycle_mom(self):
        sf = self._get_scale_factor()
        return [(hi[0] - (hi[0] - lo[0]) * sf, lo[1]) for lo, hi in zip(self.min_moms, self.max_moms)]

    def _get_decay_lr(self, decay_batch_iteration):
        interval = decay_batch_iteration / self.decay_step_size
        return [lr / (1 + self.decay_lr_rate * interval) for lr in self.min_lrs]

    def _get_decay_mom(self, decay_batch_iteration):
        interval = decay_batch_iteration / self.decay_step_size
        f = 1 + self.decay_mom_rate * interval
        return [(b0 * f, b1) for b0, b1 in self.max_moms]

    def get_lr(self):
        if self.last_batch_iteration < self.total_size:
            return self._get_cycle_lr()
        return self._get_decay_lr(self.last_batch_iteration - self.total_size + 1)

    def get_mom(self):
        if not self.cycle_momentum:
            return None
        if self.last_batch_iteration < self.total_size:
            return self._get_cycle_mom()
        return self._get_decay_mom(self.last_batch_iteration - self.total_size + 1)

    def step(self, batch_iteration=None):
        if batch_iteration is None:
            batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = batch_iteration
        self._apply()
        if self.cycle_momentum:
            for g, m in zip(self.optimizer.param_groups, self.get_mom()):
                g["betas"] = m


class WarmupLR(_Schedule):
    """lr = min + (max - min) * log(it + 1) / log(warmup_steps) during warm-up, then max."""

    def __init__(self, optimizer, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 last_batch_iteration=-1):
        self.optimizer = get_torch_optimizer(optimizer)
        self.min_lrs = _format_param(self.optimizer, warmup_min_lr, "min_lr")
        self.max_lrs = _format_param(self.optimizer, warmup_max_lr, "max_lr")
        self.delta_lrs = [hi - lo for hi, lo in zip(self.max_lrs, self.min_lrs)]
        self.warmup_num_steps = max(2, warmup_num_steps)
        self.inverse_log_warm_up = 1.0 / math.log(self.warmup_num_steps)
        self.last_batch_iteration = last_batch_iteration

    def get_lr(self):
        if self.last_batch_iteration < 0:
            logger.warning("Attempting to get learning rate from scheduler before it has started")
            return [0.0]
        gamma = self._get_gamma()
        return [lo + d * gamma for lo, d in zip(self.min_lrs, self.delta_lrs)]

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        self._apply()

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
        return 1.0


class WarmupDecayLR(WarmupLR):
    """WarmupLR followed by linear decay to 0 at total_num_steps."""

    def __init__(self, optimizer, total_num_steps, warmup_min_lr=0.0, warmup_max_lr=0.001, warmup_num_steps=1000,
                 last_batch_iteration=-1):
        self.total_num_steps = total_num_steps
        super().__init__(optimizer, warmup_min_lr, warmup_max_lr, warmup_num_steps, last_batch_iteration)
        if self.total_num_steps < self.warmup_num_steps:
            logger.warning("total_num_steps {} is less than warmup_num_steps {}".format(total_num_steps,
                                                                                        warmup_num_steps))

    def _get_gamma(self):
        if self.last_batch_iteration < self.warmup_num_steps:
            return self.inverse_log_warm_up * math.log(self.last_batch_iteration + 1)
        return max(0.0, float(self.total_num_steps - self.last_batch_iteration) /
                   float(max(1.0, self.total_num_steps - self.warmup_num_steps)))
