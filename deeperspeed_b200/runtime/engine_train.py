"""forward / backward / step / gradient reduction of DeepSpeedEngine (mixin; see engine.py for the class).
Reference call stack: deepspeed/runtime/engine.py:959-1460."""
import torch
import torch.distributed as dist

from ..utils.logging import logger, log_dist
from . import constants as C
from .csr_tensor import CSRTensor

MEMORY_OPT_ALLREDUCE_SIZE = 500000000


def _get_global_rank(group, group_rank):
    try:
        return dist.get_global_rank(group, group_rank)
    except Exception:
        return group_rank


def is_zero_supported_optimizer(optimizer):
    from ..ops.adam import FusedAdam, DeepSpeedCPUAdam
    ok = (torch.optim.Adam, torch.optim.AdamW, FusedAdam, DeepSpeedCPUAdam)
    return isinstance(optimizer, ok)


def _flatten(tensors):
    return torch.cat([t.contiguous().view(-1) for t in tensors])


def _unflatten(flat, tensors):
    outs, off = [], 0
    for t in tensors:
        outs.append(flat.narrow(0, off, t.numel()).view_as(t))
        off += t.numel()
    return outs


class TrainStepMixin:
    # ------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        r"""Execute forward propagation."""
        if self.flops_profiler_enabled() and self.global_steps == self.flops_profiler_profile_step() and self.global_rank == 0:
            from ..profiling.flops_profiler.profiler import FlopsProfiler
            self.flops_profiler = FlopsProfiler(self.module)
            self.flops_profiler.start_profile(ignore_list=None)

        if self.module.training and self.progressive_layer_drop:
            kwargs.update(self.progressive_layer_drop.get_state())

        if self.zero_optimization_partition_weights():
            for module in self.module.modules():
                module._parameters._in_forward = True

        if self.wall_clock_breakdown():
            self.timers("forward_microstep").start()
            self.timers("forward").start()
        if self.training_dataloader is None:
            self.tput_timer.start()
        loss = self.module(*inputs, **kwargs)

        if self.zero_optimization_partition_weights():
            if not torch._C.is_grad_enabled():
                self.optimizer.param_coordinator.reset_step()
            for module in self.module.modules():
                module._parameters._in_forward = False

        if self.wall_clock_breakdown():
            self.timers("forward").stop()
            self.timers("forward_microstep").stop()

        if self.flops_profiler_enabled() and self.global_steps == self.flops_profiler_profile_step() and self.global_rank == 0:
            self.flops_profiler.print_model_profile(profile_step=self.global_steps,
                                                    module_depth=self.flops_profiler_module_depth(),
                                                    top_modules=self.flops_profiler_top_modules(),
                                                    detailed=self.flops_profiler_detailed())
            self.flops_profiler.flops = self.flops_profiler.get_total_flops()
            self.flops_profiler.params = self.flops_profiler.get_total_params()
            self.flops_profiler.end_profile()
        return loss

    # ------------------------------------------------------------------ gradient reduction
    def _has_arenas(self):
        return hasattr(self.optimizer, "reduce_gradients") and hasattr(self.optimizer, "arenas")

    def allreduce_gradients(self, bucket_size=MEMORY_OPT_ALLREDUCE_SIZE):
        if self.zero_optimization_partition_weights():
            self.optimizer.overlapping_partition_gradients_reduce_epilogue()
        elif self.is_gradient_accumulation_boundary():
            if self._has_arenas():
                # in-place bucket collectives on the gradient arena (reduce-scatter under ZeRO, all-reduce otherwise)
                self.optimizer.reduce_gradients()
            else:
                self.buffered_allreduce_fallback(elements_per_buffer=bucket_size)

    def backward(self, loss, allreduce_gradients=True, release_loss=False):
        r"""Execute backward pass on the loss."""
        if not allreduce_gradients:
            logger.warning("Argument `allreduce_gradients` is deprecated, ignored, and will soon be removed")
        if self.gradient_accumulation_steps() > 1:
            loss = self._scale_loss(loss.float())

        if self.tensorboard_enabled() and self.is_gradient_accumulation_boundary() and self.global_rank == 0 \
                and getattr(self, "summary_writer", None) is not None:
            self.summary_events = [("Train/Samples/train_loss", loss.mean().item() * self.gradient_accumulation_steps(),
                                    self.global_samples)]
            for event in self.summary_events:
                self.summary_writer.add_scalar(event[0], event[1], event[2])
            self.summary_writer.flush()

        if self.wall_clock_breakdown():
            self.timers("backward_microstep").start()
            self.timers("backward").start()
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        if self.wall_clock_breakdown():
            self.timers("backward_inner_microstep").start()
            self.timers("backward_inner").start()

        if hasattr(self.optimizer, "is_gradient_accumulation_boundary"):
            self.optimizer.is_gradient_accumulation_boundary = self.is_gradient_accumulation_boundary()
        if self.zero_optimization() or hasattr(self.optimizer, "backward"):
            self.optimizer.backward(loss)
        else:
            loss.backward()

        if self.wall_clock_breakdown():
            self.timers("backward_inner").stop()
            self.timers("backward_inner_microstep").stop()
            self.timers("backward_allreduce_microstep").start()
            self.timers("backward_allreduce").start()

        if self.enable_backward_allreduce:
            self.allreduce_gradients()

        if self.wall_clock_breakdown():
            self.timers("backward_allreduce").stop()
            self.timers("backward_allreduce_microstep").stop()
            self.timers("backward").stop()
            self.timers("backward_microstep").stop()
        if release_loss:
            pass
        return loss

    def is_gradient_accumulation_boundary(self):
        """True when the current micro-step completes a global batch (an optimizer step will follow)."""
        return (self.micro_steps + 1) % self.gradient_accumulation_steps() == 0

    def zero_grad(self):
        for param_name, param in self.module.named_parameters():
            param.grad = None

    def clip_fp32_gradients(self):
        from . import utils as ds_utils
        ds_utils.clip_grad_norm_(parameters=self.module.parameters(), max_norm=self.gradient_clipping(), mpu=self.mpu)

    def _take_model_step(self, lr_kwargs):
        if self.gradient_clipping() > 0.0:
            if not self.fp16_enabled() or self.precision() == torch.float32:
                if not self.zero_optimization():
                    self.clip_fp32_gradients()
        if self.store_gradients:
            self.stored_gradients = [p.grad.clone().cpu() if self.store_gradients_cpu else p.grad.clone()
                                     for p in self.module.parameters() if p.grad is not None]
        self.optimizer.step()

        # zero grad in basic optimizer could be unreliable and may not exhibit the behaviour we want
        if not self.zero_optimization() and not (self.fp16_enabled() and self.precision() != torch.float32):
            self.zero_grad()
        elif not hasattr(self.optimizer, "arenas"):
            self.optimizer.zero_grad()

        report_progress = self.global_rank == 0 if self.global_rank else True
        overflow = getattr(self.optimizer, "overflow", False)
        if overflow:
            self.skipped_steps += 1
        else:
            if self.lr_scheduler is not None:
                self.lr_scheduler.step(**(lr_kwargs or {}))
        if report_progress and (self.global_steps + 1) % self.steps_per_print() == 0:
            self._report_progress(self.global_steps + 1)
        self.global_steps += 1
        self.global_samples += self.train_batch_size()

    def step(self, lr_kwargs=None):
        r"""Execute the weight update step after forward and backward propagation on effective_train_batch."""
        if self.wall_clock_breakdown():
            self.timers("step_microstep").start()
            self.timers("step").start()
        assert self.optimizer is not None, "must provide optimizer during init in order to use step"
        report_progress = self.global_rank == 0 if self.global_rank else True

        if self.is_gradient_accumulation_boundary():
            if self.progressive_layer_drop:
                self.progressive_layer_drop.update_state(self.global_steps)
            self._take_model_step(lr_kwargs)
        self.tput_timer.stop(report_progress)

        if self.tensorboard_enabled() and self.is_gradient_accumulation_boundary() and self.global_rank == 0 \
                and getattr(self, "summary_writer", None) is not None:
            self.summary_events = [("Train/Samples/lr", self.get_lr()[0], self.global_samples)]
            if self.fp16_enabled() and hasattr(self.optimizer, "cur_scale"):
                self.summary_events.append(("Train/Samples/loss_scale", self.optimizer.cur_scale, self.global_samples))
            for event in self.summary_events:
                self.summary_writer.add_scalar(event[0], event[1], event[2])
            self.summary_writer.flush()

        if self.wall_clock_breakdown():
            self.timers("step").stop()
            self.timers("step_microstep").stop()
            timer_names = ["forward_microstep", "backward_microstep", "backward_inner_microstep",
                           "backward_allreduce_microstep", "step_microstep"]
            self.timers.log(names=timer_names, memory_breakdown=self.memory_breakdown())
            if self.is_gradient_accumulation_boundary():
                if self.tensorboard_enabled() and self.global_rank == 0 and getattr(self, "summary_writer", None):
                    for nm in ("forward", "backward", "backward_inner", "backward_allreduce", "step"):
                        self.summary_writer.add_scalar(f"Train/Samples/elapsed_time_ms_{nm}",
                                                       self.timers(nm).elapsed(reset=False) * 1000.0,
                                                       self.global_samples)
                    self.summary_writer.flush()
                self.timers.log(["forward", "backward", "backward_inner", "backward_allreduce", "step"])
        self.micro_steps += 1

    def _get_optimizer_param(self, param_name):
        result = []
        if not self.optimizer:
            return result
        for group in self.optimizer.param_groups:
            if param_name in group:
                result.append(group[param_name])
            else:
                result.append(0.0)
        return result

    def get_lr(self):
        return self._get_optimizer_param("lr")

    def get_type(self):
        return self._get_optimizer_param("type")

    def get_mom(self):
        if self.optimizer_name() in ["SGD", "RMSprop"]:
            return self._get_optimizer_param("momentum")
        return self._get_optimizer_param("betas")

    def get_pld_theta(self):
        if self.progressive_layer_drop:
            return self.progressive_layer_drop.get_theta()
        return None

    def get_global_grad_norm(self):
        return getattr(self.optimizer, "_global_grad_norm", None)

    def _report_progress(self, step):
        lr = self.get_lr()
        mom = self.get_mom()
        log_dist(f"step={step}, skipped={self.skipped_steps}, lr={lr}, mom={mom}", ranks=[0])

    # ------------------------------------------------------------------ bucketed all-reduce fallback (no arena)
    def allreduce_bucket(self, bucket):
        tensor = _flatten(bucket)
        tensor_to_allreduce = tensor
        if self.allreduce_always_fp32():
            tensor_to_allreduce = tensor.float()
        if self.postscale_gradients():
            if self.gradient_predivide_factor() != 1.0:
                tensor_to_allreduce.mul_(1.0 / self.gradient_predivide_factor())
            dist.all_reduce(tensor_to_allreduce, group=self.data_parallel_group)
            if self.gradient_average:
                if self.gradient_predivide_factor() != self.dp_world_size:
                    tensor_to_allreduce.mul_(self.gradient_predivide_factor() / self.dp_world_size)
        else:
            tensor_to_allreduce.div_(self.dp_world_size)
            dist.all_reduce(tensor_to_allreduce, group=self.data_parallel_group)
        if self.allreduce_always_fp32() and tensor is not tensor_to_allreduce:
            tensor.copy_(tensor_to_allreduce)
        return tensor

    def allreduce_and_copy(self, small_bucket):
        allreduced = self.allreduce_bucket(small_bucket)
        for buf, synced in zip(small_bucket, _unflatten(allreduced, small_bucket)):
            buf.copy_(synced)

    def allreduce_no_retain(self, bucket, numel_per_bucket=500000000):
        small_bucket = []
        numel = 0
        for tensor in bucket:
            small_bucket.append(tensor)
            numel = numel + tensor.numel()
            if numel > numel_per_bucket:
                self.allreduce_and_copy(small_bucket)
                small_bucket = []
                numel = 0
        if len(small_bucket) > 0:
            self.allreduce_and_copy(small_bucket)

    def buffered_allreduce_fallback(self, grads=None, elements_per_buffer=500000000):
        grads = []
        for param_name, param in self.module.named_parameters():
            if param.grad is None:
                # every rank must take part with the same tensor list, so missing grads become zeros
                param.grad = torch.zeros(param.size(), dtype=param.dtype, device=param.device)
                grads.append(param.grad.data)
            else:
                grad_data = param.grad.data
                if self.sparse_gradients_enabled() and param_name in self.csr_tensor_module_names:
                    csr = CSRTensor(grad_data)
                    if grad_data.is_sparse:   # the reduced gradient replaces the sparse one as a dense tensor
                        csr.orig_param = param
                    grads.append(csr)
                elif grad_data.is_sparse:
                    param.grad = grad_data.to_dense()
                    grads.append(param.grad.data)
                else:
                    grads.append(grad_data)
        split_buckets = split_half_float_double_csr(grads)
        for i, bucket_tuple in enumerate(split_buckets):
            bucket_type, bucket = bucket_tuple
            if bucket_type == CSRTensor.type():
                self.csr_allreduce_no_retain(bucket)
            else:
                self.allreduce_no_retain(bucket, numel_per_bucket=elements_per_buffer)

    def csr_allreduce_no_retain(self, bucket):
        allreduced_csrs = self.csr_allreduce_bucket(bucket)
        for csr in allreduced_csrs:
            dense_tensor = csr.to_dense()
            if csr.orig_param is not None:
                csr.orig_param.grad = dense_tensor
            else:
                csr.orig_dense_tensor.copy_(dense_tensor)

    def csr_allreduce_bucket(self, bucket):
        return [self.csr_allreduce(csr) for csr in bucket]

    def csr_allreduce(self, csr):
        if self.postscale_gradients():
            csr.values.div_(self.dp_world_size)  # pre-divide: the gather below only concatenates
        else:
            csr.values.div_(self.dp_world_size)
        indices_device_list = self.csr_all_gather(csr.indices)
        values_device_list = self.csr_all_gather(csr.values)
        csr.indices = torch.cat(indices_device_list)
        csr.values = torch.cat(values_device_list)
        return csr

    def csr_all_gather(self, value):
        my_size = torch.LongTensor([value.size()[0]]).to(self.device)
        all_sizes = self.all_gather_scalar(my_size)
        max_size = int(torch.cat(all_sizes).max())
        fill_size = max_size - int(my_size)
        assert value.dim() in [1, 2]
        if value.dim() == 1:
            if fill_size > 0:
                value = torch.cat([value, value.new_zeros(fill_size)])
            tensor_list = [value.new_zeros(max_size) for _ in range(self.dp_world_size)]
        else:
            if fill_size > 0:
                value = torch.cat([value, value.new_zeros(fill_size, value.size()[1])])
            tensor_list = [value.new_zeros(max_size, value.size()[1]) for _ in range(self.dp_world_size)]
        dist.all_gather(tensor_list, value, group=self.data_parallel_group)
        tensors = []
        for dev_idx, t in enumerate(tensor_list):
            size = int(all_sizes[dev_idx][0])
            tensors.append(t.index_select(0, torch.arange(size, dtype=torch.long, device=self.device)))
        return tensors

    def all_gather_scalar(self, value):
        tensor_list = [value.new_zeros(value.size()) for _ in range(self.dp_world_size)]
        dist.all_gather(tensor_list, value, group=self.data_parallel_group)
        return tensor_list


def split_half_float_double_csr(tensors):
    buckets = {}
    order = []
    for t in tensors:
        key = t.type()
        if key not in buckets:
            buckets[key] = []
            order.append(key)
        buckets[key].append(t)
    return [(k, buckets[k]) for k in order]
