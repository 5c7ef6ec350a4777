"""PipelineEngine: executes pipeline schedules (1F1B training, inference) over a PipelineModule.

Public behaviour per reference deepspeed/runtime/pipe/engine.py:52-1316 (train_batch / eval_batch / inference_batch,
tensor-meta handshake, tied-weight reduction, loss aggregation, per-layer checkpoints, timer_values).
B200-native differences: NCCL send/recv are stream-ordered (the host never blocks on p2p); data-parallel gradient
reduction is an in-place reduce-scatter on the gradient arena fused with the ZeRO-1 step instead of a full
all-reduce followed by a separate sharded step; host->device batch copies are non-blocking from pinned memory.
"""
from types import MethodType

import torch
import torch.distributed as dist

from ...utils.logging import logger
from ...utils.timer import ThroughputTimer
from ..engine import DeepSpeedEngine, MEMORY_OPT_ALLREDUCE_SIZE
from ..utils import PartitionedTensor
from ..dataloader import RepeatingLoader
from ..activation_checkpointing import checkpointing as ds_checkpointing
from . import p2p
from . import schedule
from .module import PipelineModule, PipelineError
from .engine_exec import PipeExecMixin, _DTYPES

TARGET_ID = -2
LOG_STAGE = -2
DATA_PARALLEL_ID = -2


def is_even(number):
    return number % 2 == 0


def _tensor_bytes(tensor):
    return tensor.numel() * tensor.element_size()


class PipelineEngine(PipeExecMixin, DeepSpeedEngine):
    """A training engine hybrid pipeline, data, and model parallel training. Created by deepspeed.initialize()
    when a PipelineModule is provided."""

    def __init__(self, *super_args, **super_kwargs):
        super().__init__(*super_args, **super_kwargs)
        assert isinstance(self.module, PipelineModule), "model must base PipelineModule"
        assert self.zero_optimization_stage() < 2, "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"

        # the pipeline schedule decides when gradients are reduced
        self.enable_backward_allreduce = False
        self.has_bool_tensors = False
        self.eval_return_logits = False
        self.outputs = None
        self.pipeline_enable_backward_allreduce = True
        assert not self.elasticity_enabled(), "Elasticity is not currently supported with pipeline parallelism."

        self.log_batch_step_id = -1
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()

        self.grid = self.module._grid
        if self.grid.get_global_rank() == 0:
            logger.info(f"CONFIG: micro_batches={self.micro_batches} micro_batch_size={self.micro_batch_size}")
        self.global_rank = self.grid.get_global_rank()
        assert self.dp_world_size == self.grid.data_parallel_size
        assert self.train_batch_size() == self.micro_batch_size * self.micro_batches * self.grid.data_parallel_size

        self.num_stages = self.grid.pipe_parallel_size
        self.stage_id = self.grid.get_stage_id()
        self.prev_stage = self.stage_id - 1
        self.next_stage = self.stage_id + 1

        self.data_iterator = None
        self.batch_fn = None
        self._force_grad_boundary = False
        self.batch_timer = ThroughputTimer(batch_size=self.micro_batch_size * self.micro_batches,
                                           num_workers=self.dp_world_size, logging_fn=self.tput_log,
                                           monitor_memory=False, steps_per_output=self.steps_per_print())
        if self.training_data:
            self._build_data_iter(self.training_data)

        self.is_pipe_parallel = self.grid.pipe_parallel_size > 1
        self.is_data_parallel = self.grid.data_parallel_size > 1
        self.is_model_parallel = self.grid.model_parallel_size > 1
        # PartitionedTensor slicing of activations across the TP group before a send
        self.is_pipe_partitioned = self.is_model_parallel
        self.is_grad_partitioned = False

        model_parameters = filter(lambda p: p.requires_grad, self.module.parameters())
        num_params = sum(p.numel() for p in model_parameters)
        unique_params = num_params
        if self.module.tied_comms:
            tied_params = 0
            for key, d in self.module.tied_comms.items():
                if self.global_rank != min(d["ranks"]):
                    tied_params += sum(p.numel() for p in d["module"].parameters())
            unique_params -= tied_params
        params_tensor = torch.LongTensor(data=[num_params, unique_params]).to(self.device)
        dist.all_reduce(params_tensor, group=self.grid.get_model_parallel_group())
        params_tensor = params_tensor.tolist()
        total_params, unique_params = params_tensor[0], params_tensor[1]
        if self.grid.data_parallel_id == 0:
            logger.info(f"RANK={self.global_rank} STAGE={self.stage_id} LAYERS={self.module._local_stop - self.module._local_start} "
                        f"[{self.module._local_start}, {self.module._local_stop}) STAGE_PARAMS={num_params} ({num_params / 1e6:0.3f}M) "
                        f"TOTAL_PARAMS={total_params} ({total_params / 1e6:0.3f}M) "
                        f"UNIQUE_PARAMS={unique_params} ({unique_params / 1e6:0.3f}M)")

        if self.is_pipe_parallel:
            p2p.init_process_groups(self.grid)

        # pipeline buffers
        self.num_pipe_buffers = 0
        self.pipe_buffers = {"inputs": [], "labels": [], "outputs": [], "output_tensors": []}
        self.pipe_recv_buf = None
        self.grad_layer = None
        self.meta_buffer = None
        self.first_output_send = True
        self.first_gradient_send = True

        self.loss = torch.tensor(0.0).to(self.device)
        self.total_loss = None
        self.agg_loss = torch.tensor(0.0, requires_grad=False).to(self.device)
        self.dp_group_loss = torch.tensor(0.0, requires_grad=False).to(self.device)

        if self._config.pipeline["activation_checkpoint_interval"] > 0:
            self.module.activation_checkpoint_interval = self._config.pipeline["activation_checkpoint_interval"]
        if self.is_last_stage():
            self.loss_model = self.module.loss_fn

        self.has_attention_mask = self.module.__class__.__name__ == "GPT2ModelPipe"
        self.timer_values = None
        self.fp32_comm = self.precision() == torch.bfloat16 and self.allreduce_always_fp32() and \
            self._config._param_dict.get("fp32_allreduce", None) is True

        # one-time neighbour handshake so the first real transfer does not pay communicator setup
        if self.is_pipe_parallel:
            if is_even(self.stage_id):
                if not self.is_last_stage():
                    p2p.send(self.loss, self.next_stage)
                if not self.is_first_stage():
                    p2p.recv(self.loss, self.prev_stage)
            else:
                if not self.is_first_stage():
                    p2p.recv(self.loss, self.prev_stage)
                if not self.is_last_stage():
                    p2p.send(self.loss, self.next_stage)

        if self.wall_clock_breakdown():
            for name in ("forward_microstep", "backward_microstep", "backward_inner_microstep",
                         "backward_allreduce_microstep", "backward_allreduce", "step_microstep"):
                self.timers(name).start()
                self.timers(name).stop()

    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def _build_data_iter(self, dataset):
        sampler = torch.utils.data.distributed.DistributedSampler(dataset, num_replicas=self.dp_world_size,
                                                                  rank=self.mpu.get_data_parallel_rank(), shuffle=False)
        pipe_dataloader = self.deepspeed_io(dataset, data_sampler=sampler)
        pipe_dataloader = RepeatingLoader(pipe_dataloader)
        self.set_dataloader(pipe_dataloader)

    # ------------------------------------------------------------------ reductions
    def _exec_reduce_tied_grads(self):
        self.module.allreduce_tied_weight_gradients()

    def _exec_reduce_grads(self):
        self._force_grad_boundary = True
        if self.pipeline_enable_backward_allreduce:
            if self._has_arenas():
                self.optimizer.reduce_gradients()
            elif self.is_data_parallel:
                self.buffered_allreduce_fallback(elements_per_buffer=MEMORY_OPT_ALLREDUCE_SIZE)
        self._force_grad_boundary = False

    def _reserve_pipe_buffers(self, num_buffers):
        if self.num_pipe_buffers >= num_buffers:
            return
        num_added = num_buffers - self.num_pipe_buffers
        for key in self.pipe_buffers:
            self.pipe_buffers[key].extend([None] * num_added)
        self.num_pipe_buffers = num_buffers

    # ------------------------------------------------------------------ public batch APIs
    def _hooked(self, layers_to_hook):
        """Context for the `layers_to_hook` argument of the batch APIs (reference pipe/engine.py:292-293,346-347)."""
        import contextlib

        @contextlib.contextmanager
        def ctx():
            if layers_to_hook is None:
                yield
                return
            if not self._hook_handles:
                self.register_forward_hook(layers_to_hook)
            self.layers_to_hook = layers_to_hook
            self.layer_outputs = {}
            try:
                yield
            finally:
                self.layers_to_hook = []
        return ctx()

    def train_batch(self, data_iter=None, layers_to_hook=None):
        """Progress the pipeline to train the next batch of data (gradient_accumulation_steps micro-batches).
        Returns the loss averaged over micro-batches and data-parallel replicas."""
        with self._hooked(layers_to_hook):
            return self._train_batch(data_iter)

    def _train_batch(self, data_iter=None):
        if not torch._C.is_grad_enabled():
            raise RuntimeError("train_batch() requires gradients enabled. Use eval_batch() instead.")
        if data_iter is not None:
            self.set_dataiterator(data_iter)
        self.module.train()
        self.total_loss = None
        self.timers("train_batch").start()
        sched = schedule.TrainSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        self._exec_schedule(sched)
        self.agg_train_loss = self._aggregate_total_loss()
        self.timers("train_batch").stop()

        if self.global_steps % self.steps_per_print() == 0:
            if self.global_rank == 0:
                elapsed = self.timers("train_batch").elapsed(reset=True)
                iter_time = elapsed / self.steps_per_print()
                tput = self.train_batch_size() / max(iter_time, 1e-9)
                print(f"steps: {self.global_steps} loss: {self.agg_train_loss:0.4f} iter time (s): {iter_time:0.3f} "
                      f"samples/sec: {tput:0.3f}")
        if self.tensorboard_enabled() and self.global_rank == 0 and getattr(self, "summary_writer", None):
            self.summary_writer.add_scalar("Train/Samples/train_loss", self.agg_train_loss.mean().item(),
                                           self.global_samples)
            if self.global_steps % self.steps_per_print() == 0:
                self.summary_writer.flush()
        if self.wall_clock_breakdown() and self.global_steps % self.steps_per_print() == 0:
            names = ["pipe_send_output", "pipe_send_grad", "pipe_recv_input", "pipe_recv_grad", "batch_input",
                     "forward_microstep", "backward_microstep", "step_microstep", "reduce_grads", "reduce_tied_grads"]
            vals = {n: self.timers(n).elapsed(reset=True) for n in names if n in self.timers.timers}
            comms = sum(vals.get(n, 0.0) for n in ("pipe_send_output", "pipe_send_grad", "pipe_recv_input",
                                                   "pipe_recv_grad", "reduce_grads", "reduce_tied_grads"))
            total = max(sum(vals.values()), 1e-9)
            self.timer_values = {
                "%comms": 100.0 * comms / total,
                "%optimizer_step": 100.0 * vals.get("step_microstep", 0.0) / total,
                "%forward": 100.0 * vals.get("forward_microstep", 0.0) / total,
                "%backward": 100.0 * vals.get("backward_microstep", 0.0) / total,
            }
            if self.global_rank == 0:
                logger.info(" | ".join(f"{k}: {v:.2f}" for k, v in self.timer_values.items()))
        return self.agg_train_loss

    def eval_batch(self, data_iter, return_logits=False, layers_to_hook=None):
        """Evaluate the pipeline on a batch of data (no gradients, no optimizer step)."""
        with self._hooked(layers_to_hook):
            return self._eval_batch(data_iter, return_logits)

    def _eval_batch(self, data_iter, return_logits=False):
        self.eval_return_logits = return_logits
        self.module.eval()
        self.total_loss = None
        train_iterator = self.data_iterator
        self.set_dataiterator(data_iter)
        sched = schedule.InferenceSchedule(micro_batches=self.micro_batches, stages=self.num_stages, stage_id=self.stage_id)
        with torch.no_grad():
            self._exec_schedule(sched)
        self.agg_eval_loss = self._aggregate_total_loss()
        if self.tensorboard_enabled() and self.global_rank == 0 and getattr(self, "summary_writer", None):
            self.summary_writer.add_scalar("Train/Samples/eval_loss", self.agg_eval_loss.mean().item(), self.global_samples)
            self.summary_writer.flush()
        self.set_dataiterator(train_iterator)
        if return_logits:
            outputs = self.outputs
            self.outputs = None
            self.eval_return_logits = False
            return self.agg_eval_loss, outputs
        return self.agg_eval_loss

    def inference_batch(self, data_iter, layers_to_hook=None):
        with self._hooked(layers_to_hook):
            return self._inference_batch(data_iter)

    def _inference_batch(self, data_iter):
        """Run one forward-only micro-batch through the pipeline and return (logits, presents) on every rank of the
        pipe group (broadcast from the last stage). Shapes may vary call to call, so the meta handshake is redone."""
        self.module.eval()
        self.total_loss = None
        train_iterator = self.data_iterator
        self.set_dataiterator(data_iter)
        # shapes may differ from training: reset the cached meta/handshake buffers
        self.first_output_send = True
        self.pipe_recv_buf = None
        sched = schedule.InferenceSchedule(micro_batches=1, stages=self.num_stages, stage_id=self.stage_id)
        self._inference_mode = True
        with torch.no_grad():
            self._exec_schedule(sched)
        self._inference_mode = False
        outputs = self.pipe_buffers["outputs"][0] if self.is_last_stage() else None
        logits, presents = None, None
        if self.is_last_stage():
            if isinstance(outputs, (tuple, list)):
                logits = outputs[0]
                presents = outputs[1] if len(outputs) > 1 else None
            else:
                logits = outputs
        if self.is_pipe_parallel:
            src = self.grid.stage_to_global(self.num_stages - 1)
            logits = self._bcast_tensor(logits, src)
            has_presents = torch.tensor([1 if presents is not None else 0], device=self.device)
            dist.broadcast(has_presents, src=src, group=self.grid.get_pipe_parallel_group())
            if int(has_presents.item()):
                presents = self._bcast_tensor(presents, src)
        self.set_dataiterator(train_iterator)
        self.first_output_send = True
        self.pipe_recv_buf = None
        return logits, presents

    def _bcast_tensor(self, tensor, src):
        group = self.grid.get_pipe_parallel_group()
        if self.global_rank == src:
            meta = torch.tensor([tensor.dim()] + list(tensor.shape) + [_DTYPES.index(tensor.dtype)], device=self.device)
            n = torch.tensor([meta.numel()], device=self.device)
        else:
            n = torch.tensor([0], device=self.device)
        dist.broadcast(n, src=src, group=group)
        if self.global_rank != src:
            meta = torch.zeros(int(n.item()), dtype=torch.long, device=self.device)
        dist.broadcast(meta, src=src, group=group)
        if self.global_rank != src:
            m = meta.tolist()
            tensor = torch.empty(m[1:1 + m[0]], dtype=_DTYPES[m[-1]], device=self.device)
        tensor = tensor.contiguous()
        dist.broadcast(tensor, src=src, group=group)
        return tensor

    def is_first_stage(self):
        return self.stage_id == 0

    def is_last_stage(self):
        return self.stage_id == self.num_stages - 1

    def _aggregate_total_loss(self):
        # scale loss, average among DP ranks, and bcast loss to the rest of my DP group
        # a module without loss_fn "returns" its last-stage output: that is a whole tensor, not a scalar
        # (reference tests/unit/test_pipe_module.py compares eval_batch() with the sequential forward)
        src_rank = self.grid.stage_to_global(self.num_stages - 1)
        flag = torch.zeros(1, dtype=torch.long, device=self.device)
        if self.is_last_stage():
            flag[0] = int(isinstance(self.total_loss, torch.Tensor) and self.total_loss.numel() > 1)
        if self.is_pipe_parallel:
            dist.broadcast(flag, src=src_rank, group=self.grid.get_pipe_parallel_group())
        if int(flag.item()):
            agg = None
            if self.is_last_stage():
                agg = self._scale_loss(self.total_loss).clone().detach()
                self.dp_group_loss = agg.clone()
                if self.is_data_parallel:
                    dist.all_reduce(agg, group=self.mpu.get_data_parallel_group())
                    agg /= self.dp_world_size
            if self.is_pipe_parallel:
                agg = self._bcast_tensor(agg, src_rank)
            return agg
        if self.is_last_stage():
            loss = self._scale_loss(self.total_loss)
            self.dp_group_loss = loss.clone().detach()
            agg_loss = self.dp_group_loss.clone().detach()
            if self.is_data_parallel:
                dist.all_reduce(agg_loss, group=self.mpu.get_data_parallel_group())
                agg_loss /= self.dp_world_size
            assert self.global_rank in self.grid.pp_group
            losses = torch.stack([self.dp_group_loss.float().reshape(()), agg_loss.float().reshape(())])
            if self.is_pipe_parallel:
                dist.broadcast(tensor=losses, src=self.global_rank, group=self.mpu.get_pipe_parallel_group())
        else:
            src_rank = self.grid.stage_to_global(self.num_stages - 1)
            assert src_rank in self.grid.pp_group
            losses = torch.zeros(2, dtype=torch.float32, device=self.device)
            dist.broadcast(tensor=losses, src=src_rank, group=self.grid.get_pipe_parallel_group())
            self.dp_group_loss = losses[0].clone().detach()
            agg_loss = losses[1].clone().detach()
        return agg_loss

    def set_dataloader(self, loader):
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = loader
            self.data_iterator = iter(self.training_dataloader)

    def set_dataiterator(self, iterator):
        if self.is_first_stage() or self.is_last_stage():
            self.training_dataloader = None
            self.data_iterator = iterator

    def set_batch_fn(self, fn):
        self.batch_fn = fn

    def is_gradient_accumulation_boundary(self):
        """The schedule, not the micro-step counter, decides when an optimizer step happens."""
        return self._force_grad_boundary

    def log_for_device(self, *msg):
        if LOG_STAGE == self.stage_id or LOG_STAGE == -1:
            if DATA_PARALLEL_ID == self.grid.data_parallel_id or DATA_PARALLEL_ID == -1:
                print(f"RANK={dist.get_rank()} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} "
                      f"MBATCH-ID={self.microbatch_id} STEP-ID={self.log_batch_step_id} ::", *msg, flush=True)

    def tput_log(self, *msg):
        if self.global_rank == 0 and self.global_steps % self.steps_per_print() == 0:
            print(*msg)

    def _next_batch(self):
        if self.is_model_parallel:
            mp_rank = self.grid.get_slice_parallel_rank()
        else:
            mp_rank = 0
        batch = None
        # only TP rank 0 pulls from the iterator; batch_fn may broadcast to the rest of the TP group
        if mp_rank == 0:
            batch = next(self.data_iterator)
        if self.batch_fn:
            batch = self.batch_fn(batch)
        return batch
