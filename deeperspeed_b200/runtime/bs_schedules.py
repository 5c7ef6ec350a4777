"""Batch-size scheduler (DeeperSpeed addition; reference deepspeed/runtime/bs_schedules.py:5-68).

Constructor and behaviour of the reference: the batch size climbs in `num_intervals` equally spaced steps from
ceil(`min_batch_size_multiplier` x `final_batch_size`) to `final_batch_size` over `warmup_num_steps` optimizer steps,
then stays there; `step()` advances one iteration and leaves the value in `current_batch_size`.

Two keyword-only alternatives kept from this package's first version: `start_batch_size` (an absolute starting value
instead of a fraction) and `steps_per_interval` (interval length instead of the total warm-up); with either of them
`min_batch_size_multiplier` >= 1 is the granularity every value is rounded to. An engine passed as `deepspeed=` gets
`set_train_batch_size()` whenever the value changes; `step()` also returns the current value."""
import math


def _evenly_spaced_ints(first, last, count):
    """`count` integers from first to last inclusive, truncated like numpy.linspace(..., dtype=int)."""
    if count <= 1:
        return [int(first)]
    return [int(first + (last - first) * i / (count - 1)) for i in range(count)]


class BatchSizeScheduler:
    def __init__(self, final_batch_size, min_batch_size_multiplier=0.01, warmup_num_steps=1000, num_intervals=4,
                 last_batch_iteration=-1, deepspeed=None, *, start_batch_size=None, steps_per_interval=None):
        self.final_batch_size = final_batch_size
        self.min_batch_size_multiplier = min_batch_size_multiplier
        self.num_intervals = max(int(num_intervals), 1)
        self.last_batch_iteration = last_batch_iteration
        self.deepspeed = deepspeed
        self.start_batch_size = start_batch_size
        self.steps_per_interval = steps_per_interval
        if steps_per_interval is not None:
            warmup_num_steps = int(steps_per_interval) * self.num_intervals
        self.warmup_num_steps = warmup_num_steps
        self.schedule = self._build_schedule()
        self.current_batch_size = None

    def _build_schedule(self):
        """{first iteration of an interval: batch size}; intervals that would repeat the previous size are dropped."""
        absolute = self.start_batch_size is not None or self.steps_per_interval is not None
        if absolute:
            grain = max(int(self.min_batch_size_multiplier), 1)
            start = self.start_batch_size if self.start_batch_size is not None else grain
            points = self.num_intervals + 1
        else:
            grain = 1
            start = math.ceil(self.min_batch_size_multiplier * self.final_batch_size)
            points = self.num_intervals
        sizes = _evenly_spaced_ints(start, self.final_batch_size, points)
        if grain > 1:
            sizes = [min(max(int(round(b / grain)) * grain, grain), self.final_batch_size) for b in sizes]
        starts = _evenly_spaced_ints(0, self.warmup_num_steps, points)
        schedule, previous = {}, None
        for at, size in zip(starts, sizes):
            if size != previous and at not in schedule:
                schedule[at] = size
            previous = size
        return schedule

    def get_current_batch_size(self):
        reached = [at for at in self.schedule if at <= self.last_batch_iteration]
        return self.schedule[max(reached) if reached else min(self.schedule)]

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        size = self.get_current_batch_size()
        if size != self.current_batch_size:
            self.current_batch_size = size
            if self.deepspeed is not None and hasattr(self.deepspeed, "set_train_batch_size"):
                self.deepspeed.set_train_batch_size(size)
        return size

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]
