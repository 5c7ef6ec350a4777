"""Batch-size scheduler (DeeperSpeed addition; reference deepspeed/runtime/bs_schedules.py:5-68): linearly ramps the
effective batch size from `start_batch_size` to `final_batch_size` over `num_intervals` intervals of
`steps_per_interval` optimizer steps, in multiples of `min_batch_size_multiplier`."""


class BatchSizeScheduler:
    def __init__(self, final_batch_size, start_batch_size=1, num_intervals=8, steps_per_interval=1000,
                 min_batch_size_multiplier=1, last_batch_iteration=-1, deepspeed=None):
        self.final_batch_size = final_batch_size
        self.start_batch_size = start_batch_size
        self.num_intervals = max(int(num_intervals), 1)
        self.steps_per_interval = max(int(steps_per_interval), 1)
        self.min_batch_size_multiplier = max(int(min_batch_size_multiplier), 1)
        self.last_batch_iteration = last_batch_iteration
        self.deepspeed = deepspeed
        self.schedule = self._build_schedule()
        self.current_batch_size = None

    def _build_schedule(self):
        sched = {}
        span = self.final_batch_size - self.start_batch_size
        for i in range(self.num_intervals + 1):
            bs = self.start_batch_size + span * i / self.num_intervals
            bs = int(round(bs / self.min_batch_size_multiplier)) * self.min_batch_size_multiplier
            bs = min(max(bs, self.min_batch_size_multiplier), self.final_batch_size)
            sched[i * self.steps_per_interval] = bs
        return sched

    def get_current_batch_size(self):
        it = max(self.last_batch_iteration, 0)
        keys = [k for k in self.schedule if k <= it]
        return self.schedule[max(keys)] if keys else self.start_batch_size

    def step(self, last_batch_iteration=None):
        if last_batch_iteration is None:
            last_batch_iteration = self.last_batch_iteration + 1
        self.last_batch_iteration = last_batch_iteration
        bs = self.get_current_batch_size()
        if bs != self.current_batch_size:
            self.current_batch_size = bs
            if self.deepspeed is not None and hasattr(self.deepspeed, "set_train_batch_size"):
                self.deepspeed.set_train_batch_size(bs)
        return bs

    def state_dict(self):
        return {"last_batch_iteration": self.last_batch_iteration}

    def load_state_dict(self, sd):
        self.last_batch_iteration = sd["last_batch_iteration"]
