"""Timers. `SynchronizedWallClockTimer` keeps the reference API (deepspeed/utils/timer.py:19-101) but measures
on the device with CUDA events when CUDA is present, so starting/stopping a timer does not serialise the stream;
elapsed() synchronises only on the recorded end event. `ThroughputTimer` as reference timer.py:105-182.

Beyond the reference (which has no NVTX ranges and no trace export): with DSB_NVTX=1 every timer interval
is also an NVTX range (visible in nsys / ncu), and `enable_trace()` + `export_chrome_trace(path)` (or
DSB_TIMER_TRACE=<path>, written at interpreter exit) dump every interval with device timestamps as a Chrome /
Perfetto trace, one row per timer name."""
import atexit
import json
import os
import time

import torch

from .logging import log_dist

try:
    import psutil
except ImportError:  # pragma: no cover
    psutil = None


def _cuda():
    return torch.cuda.is_available()


_NVTX = os.environ.get("DSB_NVTX", "0") == "1"


class SynchronizedWallClockTimer:
    class Timer:
        def __init__(self, name, owner=None):
            self.name_ = name
            self.elapsed_ = 0.0
            self.started_ = False
            self._t0 = None
            self._pending = []  # (start_event, end_event) pairs not yet folded into elapsed_
            self._owner = owner

        def start(self):
            assert not self.started_, f"timer {self.name_} has already been started"
            if _cuda():
                if _NVTX:
                    torch.cuda.nvtx.range_push(self.name_)
                self._t0 = torch.cuda.Event(enable_timing=True)
                self._t0.record()
            else:
                self._t0 = time.time()
            self.started_ = True

        def stop(self, reset=False):
            assert self.started_, "timer is not started"
            if _cuda():
                e = torch.cuda.Event(enable_timing=True)
                e.record()
                self._pending.append((self._t0, e))
                if _NVTX:
                    torch.cuda.nvtx.range_pop()
            else:
                now = time.time()
                dt = now - self._t0
                self.elapsed_ = dt if reset else self.elapsed_ + dt
                if self._owner is not None and self._owner._trace is not None:
                    self._owner._trace.append((self.name_, (self._t0 - self._owner._trace_t0) * 1e6, dt * 1e6))
            self.started_ = False

        def _fold(self):
            own = self._owner
            for s, e in self._pending:
                e.synchronize()
                ms = s.elapsed_time(e)
                self.elapsed_ += ms / 1000.0
                if own is not None and own._trace is not None:
                    own._trace.append((self.name_, own._trace_base.elapsed_time(s) * 1e3, ms * 1e3))
            self._pending = []

        def _fold_keep(self):
            self._fold()

        def reset(self):
            if self._owner is not None and self._owner._trace is not None:
                self._fold()   # a traced interval is recorded even if its value is discarded
            self.elapsed_ = 0.0
            self._pending = []
            self.started_ = False

        def elapsed(self, reset=True):
            started = self.started_
            if started:
                self.stop()
            self._fold()
            v = self.elapsed_
            if reset:
                self.reset()
            if started:
                self.start()
            return v

    def __init__(self):
        self.timers = {}
        self._trace = None
        self._trace_base = None
        self._trace_t0 = 0.0
        path = os.environ.get("DSB_TIMER_TRACE")
        if path:
            self.enable_trace()
            atexit.register(self._export_at_exit, path)

    def __call__(self, name):
        if name not in self.timers:
            self.timers[name] = self.Timer(name, owner=self)
        return self.timers[name]

    # ---- trace export (no reference counterpart)
    def enable_trace(self):
        """Start collecting (name, start, duration) of every timer interval; timestamps are device time relative to
        this call (CUDA events) or host time on CPU."""
        self._trace = []
        self._trace_t0 = time.time()
        if _cuda():
            self._trace_base = torch.cuda.Event(enable_timing=True)
            self._trace_base.record()

    def trace_events(self):
        if self._trace is None:
            return []
        for t in self.timers.values():
            if not t.started_:
                t._fold_keep()
        return list(self._trace)

    def export_chrome_trace(self, path, rank=0):
        """Write the collected intervals as a Chrome / Perfetto trace (open in chrome://tracing or ui.perfetto.dev)."""
        names = sorted({n for n, _, _ in self.trace_events()})
        tid = {n: i for i, n in enumerate(names)}
        ev = [{"name": n, "ph": "X", "ts": ts, "dur": dur, "pid": rank, "tid": tid[n]} for n, ts, dur in self._trace or []]
        ev += [{"name": "thread_name", "ph": "M", "pid": rank, "tid": i, "args": {"name": n}} for n, i in tid.items()]
        with open(path, "w") as f:
            json.dump({"traceEvents": ev, "displayTimeUnit": "ms"}, f)
        return len(ev)

    def _export_at_exit(self, path):
        try:
            import torch.distributed as dist
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
            root, ext = os.path.splitext(path)
            self.export_chrome_trace(f"{root}.rank{rank}{ext or '.json'}", rank=rank)
        except Exception:   # never fail interpreter shutdown over a trace
            pass

    @staticmethod
    def memory_usage():
        if not _cuda():
            return ""
        gb = 1024 ** 3
        return " | mem_allocated: {:.4f} GB | max_mem_allocated: {:.4f} GB | cache_allocated: {:.4f} GB | " \
               "max_cache_allocated: {:.4f} GB".format(torch.cuda.memory_allocated() / gb,
                                                       torch.cuda.max_memory_allocated() / gb,
                                                       torch.cuda.memory_reserved() / gb,
                                                       torch.cuda.max_memory_reserved() / gb)

    def log(self, names, normalizer=1.0, reset=True, memory_breakdown=False, ranks=None, return_values=True):
        """Log a group of timers; returns {name: elapsed ms / normalizer} (reference utils/timer.py:84-102) unless
        `return_values` is false."""
        assert normalizer > 0.0
        s, values = "time (ms)", {}
        for name in names:
            if name in self.timers:
                values[name] = self.timers[name].elapsed(reset=reset) * 1000.0 / normalizer
                s += " | {}: {:.2f}".format(name, values[name])
            else:
                log_dist(f"logging failed for timer {name}", ranks=[0])
        if memory_breakdown:
            s += self.memory_usage()
        log_dist(s, ranks=ranks or [0])
        return values if return_values else None


class ThroughputTimer:
    def __init__(self, batch_size, num_workers, start_step=2, steps_per_output=50, monitor_memory=False,
                 logging_fn=None):
        self.start_time = 0
        self.end_time = 0
        self.started = False
        self.batch_size = batch_size or 1
        self.num_workers = num_workers
        self.start_step = start_step
        self.epoch_count = 0
        self.local_step_count = 0
        self.total_step_count = 0
        self.total_elapsed_time = 0
        self.steps_per_output = steps_per_output
        self.monitor_memory = monitor_memory
        self.logging = logging_fn or (lambda m: log_dist(m, ranks=[0]))
        self.initialized = False

    def update_epoch_count(self):
        self.epoch_count += 1
        self.local_step_count = 0

    def _init_timer(self):
        self.initialized = True

    def start(self):
        self._init_timer()
        self.started = True
        if self.total_step_count >= self.start_step:
            if _cuda():
                torch.cuda.synchronize()
            self.start_time = time.time()

    def stop(self, report_speed=True):
        if not self.started:
            return
        self.started = False
        self.total_step_count += 1
        self.local_step_count += 1
        if self.total_step_count > self.start_step:
            if _cuda():
                torch.cuda.synchronize()
            self.end_time = time.time()
            self.total_elapsed_time += self.end_time - self.start_time
            if report_speed and self.local_step_count % self.steps_per_output == 0:
                self.logging("{}/{}, SamplesPerSec={}".format(self.epoch_count, self.local_step_count,
                                                              self.avg_samples_per_sec()))
                if self.monitor_memory and psutil is not None:
                    vm = psutil.virtual_memory()
                    self.logging("{}/{}, vm percent: {}, swap percent: {}".format(
                        self.epoch_count, self.local_step_count, vm.percent, psutil.swap_memory().percent))

    def avg_samples_per_sec(self):
        if self.total_step_count > self.start_step and self.total_elapsed_time > 0:
            samples_per_step = self.batch_size * self.num_workers
            avg = self.total_elapsed_time / (self.total_step_count - self.start_step)
            return samples_per_step / avg
        return float("-inf")
