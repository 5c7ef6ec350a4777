"""forward / backward / step of DeepSpeedEngine (mixin; see engine.py for the class).

Call contract per reference deepspeed/runtime/engine.py:959-1460: `loss = engine(x); engine.backward(loss);
engine.step()` once per micro-batch; the optimizer runs on the micro-step that completes a global batch. What is
specific to this engine: gradients accumulate straight into the flat arena, so "reduction" is an in-place bucket
collective on arena slices (`optimizer.reduce_gradients`), overflow / norm / clip never leave the device, and wall-clock
breakdown timers are CUDA-event based (one context manager, no device synchronisation on the hot path).
"""
import contextlib

import torch
import torch.distributed as dist

from ..utils.logging import logger, log_dist
from . import constants as C
from .grad_reduce import GradientReducer, flatten_dense as _flatten, unflatten_like as _unflatten, \
    split_half_float_double_csr  # noqa: F401  (re-exported names)

MEMORY_OPT_ALLREDUCE_SIZE = 500000000

_MICRO_TIMERS = ["forward_microstep", "backward_microstep", "backward_inner_microstep", "backward_allreduce_microstep",
                 "step_microstep"]
_BATCH_TIMERS = ["forward", "backward", "backward_inner", "backward_allreduce", "step"]


def _get_global_rank(group, group_rank):
    try:
        return dist.get_global_rank(group, group_rank)
    except Exception:
        return group_rank


def is_zero_supported_optimizer(optimizer):
    from ..ops.adam import FusedAdam, DeepSpeedCPUAdam
    return isinstance(optimizer, (torch.optim.Adam, torch.optim.AdamW, FusedAdam, DeepSpeedCPUAdam))


class TrainStepMixin:
    # ------------------------------------------------------------------ small utilities
    @contextlib.contextmanager
    def _timed(self, *names):
        """wall_clock_breakdown timers around a block (started in order, stopped in reverse)."""
        on = self.wall_clock_breakdown()
        if on:
            for n in names:
                self.timers(n).start()
        try:
            yield
        finally:
            if on:
                for n in reversed(names):
                    self.timers(n).stop()

    def _tb_scalars(self, events, flush=True):
        """Write (tag, value, step) triples if TensorBoard is on and this is global rank 0."""
        writer = getattr(self, "summary_writer", None)
        if not (self.tensorboard_enabled() and self.global_rank == 0 and writer is not None):
            return
        for tag, value, step in events:
            writer.add_scalar(tag, value, step)
        if flush:
            writer.flush()

    def _profiling_this_step(self):
        return self.flops_profiler_enabled() and self.global_rank == 0 and \
            self.global_steps == self.flops_profiler_profile_step()

    # ------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        r"""Execute forward propagation."""
        profiling = self._profiling_this_step()
        if profiling:
            from ..profiling.flops_profiler.profiler import FlopsProfiler
            self.flops_profiler = FlopsProfiler(self.module)
            self.flops_profiler.start_profile(ignore_list=None)
        if self.module.training and self.progressive_layer_drop:
            kwargs.update(self.progressive_layer_drop.get_state())

        zero3 = self.zero_optimization_partition_weights()
        if zero3:     # parameters are fetched on access while a forward is running
            for m in self.module.modules():
                m._parameters._in_forward = True
        if self.training_dataloader is None:
            self.tput_timer.start()
        with self._timed("forward_microstep", "forward"):
            amp = getattr(self, "amp", None)
            if amp is not None:
                with amp.autocast():
                    loss = self.module(*inputs, **kwargs)
            else:
                loss = self.module(*inputs, **kwargs)
        if zero3:
            if not torch._C.is_grad_enabled():
                self.optimizer.end_of_pass()      # no backward will come to release the kept tail and close the trace
            for m in self.module.modules():
                m._parameters._in_forward = False

        if profiling:
            prof = self.flops_profiler
            prof.print_model_profile(profile_step=self.global_steps, module_depth=self.flops_profiler_module_depth(),
                                     top_modules=self.flops_profiler_top_modules(),
                                     detailed=self.flops_profiler_detailed())
            prof.flops, prof.params = prof.get_total_flops(), prof.get_total_params()
            prof.end_profile()
        return loss

    # ------------------------------------------------------------------ gradient reduction
    def _has_arenas(self):
        return hasattr(self.optimizer, "reduce_gradients") and hasattr(self.optimizer, "arenas")

    def allreduce_gradients(self, bucket_size=MEMORY_OPT_ALLREDUCE_SIZE):
        if self.zero_optimization_partition_weights():
            self.optimizer.overlapping_partition_gradients_reduce_epilogue()
        elif self.is_gradient_accumulation_boundary():
            if self._has_arenas():
                self.optimizer.reduce_gradients()      # in-place bucket collectives on the gradient arena
            else:
                self.buffered_allreduce_fallback(elements_per_buffer=bucket_size)

    # ------------------------------------------------------------------ backward
    def backward(self, loss, allreduce_gradients=True, release_loss=False):
        r"""Execute backward pass on the loss."""
        if not allreduce_gradients:
            logger.warning("Argument `allreduce_gradients` is deprecated, ignored, and will soon be removed")
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        gas = self.gradient_accumulation_steps()
        if gas > 1:
            loss = self._scale_loss(loss.float())
        boundary = self.is_gradient_accumulation_boundary()
        if boundary:
            self.summary_events = [("Train/Samples/train_loss", None, self.global_samples)]
            if self.tensorboard_enabled() and self.global_rank == 0 and getattr(self, "summary_writer", None) is not None:
                self.summary_events[0] = ("Train/Samples/train_loss", loss.mean().item() * gas, self.global_samples)
                self._tb_scalars(self.summary_events)

        with self._timed("backward_microstep", "backward"):
            with self._timed("backward_inner_microstep", "backward_inner"):
                if hasattr(self.optimizer, "is_gradient_accumulation_boundary"):
                    self.optimizer.is_gradient_accumulation_boundary = boundary
                if getattr(self, "amp", None) is not None:
                    self.amp.backward(loss)            # GradScaler's loss scale
                elif self.zero_optimization() or hasattr(self.optimizer, "backward"):
                    self.optimizer.backward(loss)      # applies the loss scale
                else:
                    loss.backward()
            with self._timed("backward_allreduce_microstep", "backward_allreduce"):
                if self.enable_backward_allreduce:
                    self.allreduce_gradients()
        return loss

    def is_gradient_accumulation_boundary(self):
        """True when the current micro-step completes a global batch (an optimizer step will follow)."""
        return (self.micro_steps + 1) % self.gradient_accumulation_steps() == 0

    def zero_grad(self):
        for p in self.module.parameters():
            p.grad = None

    def clip_fp32_gradients(self):
        from . import utils as ds_utils
        ds_utils.clip_grad_norm_(parameters=self.module.parameters(), max_norm=self.gradient_clipping(), mpu=self.mpu)

    # ------------------------------------------------------------------ step
    def _take_model_step(self, lr_kwargs):
        """One optimizer step plus its bookkeeping (skipped steps, LR schedule, counters)."""
        wrapped = self.zero_optimization() or (self.fp16_enabled() and self.precision() != torch.float32)
        amp = getattr(self, "amp", None)
        if self.gradient_clipping() > 0.0 and not wrapped and amp is None:
            self.clip_fp32_gradients()      # the fp16 / ZeRO wrappers clip inside their own step
        if self.store_gradients:
            keep = (lambda g: g.clone().cpu()) if self.store_gradients_cpu else (lambda g: g.clone())
            self.stored_gradients = [keep(p.grad) for p in self.module.parameters() if p.grad is not None]
        if amp is not None:                 # unscale -> clip -> step (skipped on overflow) -> new scale
            amp.step(self.optimizer, [p for p in self.module.parameters() if p.grad is not None],
                     self.gradient_clipping())
        else:
            self.optimizer.step()
        if not wrapped:
            self.zero_grad()                # a client optimizer's own zero_grad is not relied upon
        elif not hasattr(self.optimizer, "arenas"):
            self.optimizer.zero_grad()      # arena optimizers clear their gradient arena inside step()

        if getattr(self.optimizer, "overflow", False) or (amp is not None and amp.skipped):
            self.skipped_steps += 1
        elif self.lr_scheduler is not None:
            self.lr_scheduler.step(**(lr_kwargs or {}))
        if self.global_rank in (0, None) and (self.global_steps + 1) % self.steps_per_print() == 0:
            self._report_progress(self.global_steps + 1)
        self.global_steps += 1
        self.global_samples += self.train_batch_size()

    def step(self, lr_kwargs=None):
        r"""Execute the weight update step after forward and backward propagation on effective_train_batch."""
        assert self.optimizer is not None, "must provide optimizer during init in order to use step"
        boundary = self.is_gradient_accumulation_boundary()
        with self._timed("step_microstep", "step"):
            if boundary:
                if self.progressive_layer_drop:
                    self.progressive_layer_drop.update_state(self.global_steps)
                self._take_model_step(lr_kwargs)
            self.tput_timer.stop(self.global_rank in (0, None))
            if boundary:
                self.summary_events = [("Train/Samples/lr", self.get_lr()[0], self.global_samples)]
                if self.fp16_enabled() and hasattr(self.optimizer, "cur_scale"):
                    self.summary_events.append(("Train/Samples/loss_scale", self.optimizer.cur_scale, self.global_samples))
                self._tb_scalars(self.summary_events)
        if self.wall_clock_breakdown():
            self.timers.log(names=_MICRO_TIMERS, memory_breakdown=self.memory_breakdown())
            if boundary:
                self._tb_scalars([(f"Train/Samples/elapsed_time_ms_{n}", self.timers(n).elapsed(reset=False) * 1000.0,
                                   self.global_samples) for n in _BATCH_TIMERS])
                self.timers.log(_BATCH_TIMERS)
        self.micro_steps += 1

    # ------------------------------------------------------------------ optimizer queries
    def _get_optimizer_param(self, param_name):
        if not self.optimizer:
            return []
        return [group.get(param_name, 0.0) for group in self.optimizer.param_groups]

    def get_lr(self):
        return self._get_optimizer_param("lr")

    def get_type(self):
        return self._get_optimizer_param("type")

    def get_mom(self):
        key = "momentum" if self.optimizer_name() in ["SGD", "RMSprop"] else "betas"
        return self._get_optimizer_param(key)

    def get_pld_theta(self):
        return self.progressive_layer_drop.get_theta() if self.progressive_layer_drop else None

    def get_global_grad_norm(self):
        return getattr(self.optimizer, "_global_grad_norm", None)

    def _report_progress(self, step):
        log_dist(f"step={step}, skipped={self.skipped_steps}, lr={self.get_lr()}, mom={self.get_mom()}", ranks=[0])

    # ------------------------------------------------------------------ reduction without an arena (delegated)
    def _reducer(self):
        r = getattr(self, "_grad_reducer", None)
        if r is None:
            r = self._grad_reducer = GradientReducer(self)
        return r

    def allreduce_bucket(self, bucket):
        return self._reducer().reduce_flat(bucket)

    def allreduce_and_copy(self, small_bucket):
        self._reducer().reduce_in_place(small_bucket)

    def allreduce_no_retain(self, bucket, numel_per_bucket=500000000):
        self._reducer().reduce_dense(bucket, numel_per_bucket)

    def buffered_allreduce_fallback(self, grads=None, elements_per_buffer=500000000):
        self._reducer().reduce_all(elements_per_buffer)

    def csr_allreduce_no_retain(self, bucket):
        self._reducer().reduce_sparse(bucket)

    def csr_allreduce_bucket(self, bucket):
        return [self._reducer().reduce_csr(csr) for csr in bucket]

    def csr_allreduce(self, csr):
        return self._reducer().reduce_csr(csr)

    def csr_all_gather(self, value):
        return self._reducer().gather_rows(value)

    def all_gather_scalar(self, value):
        out = [torch.zeros_like(value) for _ in range(self.dp_world_size)]
        dist.all_gather(out, value, group=self.data_parallel_group)
        return out
