"""PipelineEngine: pipeline + data (+ tensor) parallel training over a PipelineModule.

Public behaviour per reference deepspeed/runtime/pipe/engine.py:52-1316: `train_batch` / `eval_batch` /
`inference_batch`, loss averaged over micro-batches and data-parallel replicas and broadcast to every stage, tied
weight reduction, per-layer checkpoint files, `timer_values`, and forward / backward / step disabled in favour of the
batch APIs.

The work itself is done by `executor.StageExecutor`: schedules are compiled into per-tick plans, the transfers of a tick
run on per-direction links (`p2p.StageLinks`) concurrently with each other and are only awaited where their data is
consumed, and the fill / drain bubble of 1F1B can be trimmed by feeding the first and last micro-batch in slices
(`pipeline.edge_split`). The data-parallel reduction of a stage is the in-place reduce-scatter of the ZeRO-1 arena
fused with the sharded step, not an all-reduce followed by a separate step.
"""
import contextlib

import torch
import torch.distributed as dist

from ...utils.logging import logger
from ...utils.timer import ThroughputTimer
from ..engine import DeepSpeedEngine, MEMORY_OPT_ALLREDUCE_SIZE
from ..dataloader import RepeatingLoader
from . import p2p
from . import schedule
from .executor import StageExecutor, slot_rows, _DTYPES
from .module import PipelineModule, PipelineError

TARGET_ID = -2
LOG_STAGE = -2
DATA_PARALLEL_ID = -2

_COMM_TIMERS = ("pipe_send_output", "pipe_send_grad", "pipe_recv_input", "pipe_recv_grad", "reduce_grads",
                "reduce_tied_grads")
_ALL_TIMERS = _COMM_TIMERS[:4] + ("batch_input", "forward_microstep", "backward_microstep", "step_microstep") + \
    _COMM_TIMERS[4:]


def is_even(number):
    return number % 2 == 0


def _tensor_bytes(tensor):
    return tensor.numel() * tensor.element_size()


class PipelineEngine(DeepSpeedEngine):
    """Training engine for hybrid pipeline / data / tensor parallelism; `deepspeed.initialize()` returns one when the
    model is a PipelineModule."""

    def __init__(self, *super_args, **super_kwargs):
        super().__init__(*super_args, **super_kwargs)
        assert isinstance(self.module, PipelineModule), "model must base PipelineModule"
        assert self.zero_optimization_stage() < 2, "ZeRO-2 and ZeRO-3 are incompatible with pipeline parallelism"
        assert not self.elasticity_enabled(), "Elasticity is not currently supported with pipeline parallelism."

        # gradients are reduced where the schedule says so, never from engine.backward()
        self.enable_backward_allreduce = False
        self.pipeline_enable_backward_allreduce = True
        self._force_grad_boundary = False
        self.has_bool_tensors = False
        self.eval_return_logits = False
        self.outputs = None
        self.log_batch_step_id = -1

        grid = self.grid = self.module._grid
        self.global_rank = grid.get_global_rank()
        self.micro_batch_size = self.train_micro_batch_size_per_gpu()
        self.micro_batches = self.gradient_accumulation_steps()
        assert self.dp_world_size == grid.data_parallel_size
        assert self.train_batch_size() == self.micro_batch_size * self.micro_batches * grid.data_parallel_size
        if self.global_rank == 0:
            logger.info(f"CONFIG: micro_batches={self.micro_batches} micro_batch_size={self.micro_batch_size}")

        self.num_stages = grid.pipe_parallel_size
        self.stage_id = grid.get_stage_id()
        self.prev_stage, self.next_stage = self.stage_id - 1, self.stage_id + 1
        self.is_pipe_parallel = self.num_stages > 1
        self.is_data_parallel = grid.data_parallel_size > 1
        self.is_model_parallel = grid.model_parallel_size > 1
        # with tensor parallelism the first activation crosses a stage boundary as one slice per TP rank; the first
        # gradient can travel the same way (`pipeline.partition_gradients`, off by default like the reference's flag)
        self.is_pipe_partitioned = self.is_model_parallel
        self.is_grad_partitioned = self.is_model_parallel and bool(self._config.pipeline.get("partition_gradients", False))

        self.data_iterator = None
        self.batch_fn = None
        self.batch_timer = ThroughputTimer(batch_size=self.micro_batch_size * self.micro_batches,
                                           num_workers=self.dp_world_size, logging_fn=self.tput_log,
                                           monitor_memory=False, steps_per_output=self.steps_per_print())
        if self.training_data:
            self._build_data_iter(self.training_data)

        self._report_parameters()

        if self.is_pipe_parallel:
            p2p.init_process_groups(grid)
        self._exec = StageExecutor(self)
        self.pipe_buffers = self._exec.buffers          # slot buffers ("inputs", "labels", "outputs", ...)

        split = int(self._config.pipeline.get("edge_split", 1) or 1)
        if split > 1 and (self.is_model_parallel or not self.is_pipe_parallel or self.micro_batch_size % split):
            split = 1
        self.edge_split = split

        self.loss = torch.tensor(0.0).to(self.device)
        self.total_loss = None
        self.agg_loss = torch.tensor(0.0, requires_grad=False).to(self.device)
        self.dp_group_loss = torch.tensor(0.0, requires_grad=False).to(self.device)

        if self._config.pipeline["activation_checkpoint_interval"] > 0:
            self.module.activation_checkpoint_interval = self._config.pipeline["activation_checkpoint_interval"]
        self.loss_model = self.module.loss_fn if self.is_last_stage() else None
        self.has_attention_mask = self.module.__class__.__name__ == "GPT2ModelPipe"
        self.timer_values = None
        self.fp32_comm = self.precision() == torch.bfloat16 and self.allreduce_always_fp32() and \
            self._config._param_dict.get("fp32_allreduce", None) is True

        if self.is_pipe_parallel:
            self._warm_up_links()
        if self.wall_clock_breakdown():
            for name in ("forward_microstep", "backward_microstep", "backward_inner_microstep",
                         "backward_allreduce_microstep", "backward_allreduce", "step_microstep"):
                self.timers(name).start()
                self.timers(name).stop()

    # ------------------------------------------------------------------ construction helpers
    def _report_parameters(self):
        mine = sum(p.numel() for p in self.module.parameters() if p.requires_grad)
        unique = mine
        for d in (self.module.tied_comms or {}).values():
            if self.global_rank != min(d["ranks"]):       # tied copies count once, on their lowest rank
                unique -= sum(p.numel() for p in d["module"].parameters())
        counts = torch.LongTensor(data=[mine, unique]).to(self.device)
        dist.all_reduce(counts, group=self.grid.get_model_parallel_group())
        total, unique = counts.tolist()
        if self.grid.data_parallel_id == 0:
            lo, hi = self.module._local_start, self.module._local_stop
            logger.info(f"RANK={self.global_rank} STAGE={self.stage_id} LAYERS={hi - lo} [{lo}, {hi}) "
                        f"STAGE_PARAMS={mine} ({mine / 1e6:0.3f}M) TOTAL_PARAMS={total} ({total / 1e6:0.3f}M) "
                        f"UNIQUE_PARAMS={unique} ({unique / 1e6:0.3f}M)")

    def _warm_up_links(self):
        """One scalar over every link so the first real transfer does not pay for communicator set-up. Both ends of a
        link touch its two communicators in the same order (activation direction first) and a stage serves its upstream
        pair before its downstream pair, so communicator creation cannot interlock."""
        links = self._exec.links
        for link in (links.act_in, links.grad_out, links.act_out, links.grad_in):
            if link is not None:
                link.post(torch.zeros(1, device=self.device)).wait()

    def set_has_attention_mask(self, value):
        assert isinstance(value, bool)
        self.has_attention_mask = value

    def _build_data_iter(self, dataset):
        sampler = torch.utils.data.distributed.DistributedSampler(dataset, num_replicas=self.dp_world_size,
                                                                  rank=self.mpu.get_data_parallel_rank(), shuffle=False)
        self.set_dataloader(RepeatingLoader(self.deepspeed_io(dataset, data_sampler=sampler)))

    # ------------------------------------------------------------------ hooks the executor calls
    def _module_forward(self, inputs):
        return DeepSpeedEngine.forward(self, inputs)

    def _loss_backward(self, loss):
        DeepSpeedEngine.backward(self, loss)

    def _reduce_stage_gradients(self):
        self._force_grad_boundary = True
        if self.pipeline_enable_backward_allreduce:
            if self._has_arenas():
                self.optimizer.reduce_gradients()
            elif self.is_data_parallel:
                self.buffered_allreduce_fallback(elements_per_buffer=MEMORY_OPT_ALLREDUCE_SIZE)
        self._force_grad_boundary = False

    def _pipeline_optimizer_step(self, lr_kwargs=None):
        if self.wall_clock_breakdown():
            self.timers("step_microstep").start()
            self.timers("step").start()
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self._force_grad_boundary = False
        writer = getattr(self, "summary_writer", None)
        if self.tensorboard_enabled() and self.global_rank == 0 and writer:
            writer.add_scalar("Train/Samples/lr", self.get_lr()[0], self.global_samples)
            if self.fp16_enabled() and hasattr(self.optimizer, "cur_scale"):
                writer.add_scalar("Train/Samples/loss_scale", self.optimizer.cur_scale, self.global_samples)
        if self.wall_clock_breakdown():
            self.timers("step_microstep").stop()
            self.timers("step").stop()

    # aliases of the reference's instruction handlers (some integrations call them directly)
    _exec_reduce_grads = _reduce_stage_gradients
    _exec_optimizer_step = _pipeline_optimizer_step

    def _exec_reduce_tied_grads(self):
        self.module.allreduce_tied_weight_gradients()

    def _exec_schedule(self, pipe_schedule, rows=None):
        self._exec.run(pipe_schedule, rows)

    def _reserve_pipe_buffers(self, num_buffers):
        self._exec.reserve(num_buffers)

    @property
    def num_pipe_buffers(self):
        return len(self.pipe_buffers["inputs"])

    # ------------------------------------------------------------------ batch APIs
    @contextlib.contextmanager
    def _hooked(self, layers_to_hook):
        """`layers_to_hook` of the batch APIs: capture the outputs of the named layers for the duration of the call."""
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

    def train_batch(self, data_iter=None, layers_to_hook=None):
        """Train on the next batch (gradient_accumulation_steps micro-batches through the 1F1B schedule, then the
        optimizer step). Returns the loss averaged over micro-batches and data-parallel replicas, on every stage."""
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
        rows = slot_rows(self.micro_batches, self.micro_batch_size, self.edge_split)
        sched = schedule.TrainSchedule(micro_batches=len(rows), stages=self.num_stages, stage_id=self.stage_id)
        self._exec.run(sched, rows)
        self.agg_train_loss = self._aggregate_total_loss()
        self.timers("train_batch").stop()
        self._report_train_step()
        return self.agg_train_loss

    def _report_train_step(self):
        printing = self.global_steps % self.steps_per_print() == 0
        if printing and self.global_rank == 0:
            iter_time = self.timers("train_batch").elapsed(reset=True) / self.steps_per_print()
            print(f"steps: {self.global_steps} loss: {self.agg_train_loss:0.4f} iter time (s): {iter_time:0.3f} "
                  f"samples/sec: {self.train_batch_size() / max(iter_time, 1e-9):0.3f}")
        writer = getattr(self, "summary_writer", None)
        if self.tensorboard_enabled() and self.global_rank == 0 and writer:
            writer.add_scalar("Train/Samples/train_loss", self.agg_train_loss.mean().item(), self.global_samples)
            if printing:
                writer.flush()
        if self.wall_clock_breakdown() and printing:
            t = {n: self.timers(n).elapsed(reset=True) for n in _ALL_TIMERS if n in self.timers.timers}
            total = max(sum(t.values()), 1e-9)
            share = lambda *names: 100.0 * sum(t.get(n, 0.0) for n in names) / total
            self.timer_values = {"%comms": share(*_COMM_TIMERS), "%optimizer_step": share("step_microstep"),
                                 "%forward": share("forward_microstep"), "%backward": share("backward_microstep")}
            if self.global_rank == 0:
                logger.info(" | ".join(f"{k}: {v:.2f}" for k, v in self.timer_values.items()))

    @contextlib.contextmanager
    def _borrowed_iterator(self, data_iter):
        """Evaluate from `data_iter` without losing the training iterator."""
        keep = self.data_iterator
        self.set_dataiterator(data_iter)
        try:
            yield
        finally:
            self.set_dataiterator(keep)

    def eval_batch(self, data_iter, return_logits=False, layers_to_hook=None):
        """Forward-only pass over gradient_accumulation_steps micro-batches; returns the averaged loss (and the last
        stage's logits of the final micro-batch if asked)."""
        with self._hooked(layers_to_hook):
            return self._eval_batch(data_iter, return_logits)

    def _eval_batch(self, data_iter, return_logits=False):
        self.eval_return_logits = return_logits
        self.module.eval()
        self.total_loss = None
        sched = schedule.InferenceSchedule(micro_batches=self.micro_batches, stages=self.num_stages,
                                           stage_id=self.stage_id)
        with self._borrowed_iterator(data_iter), torch.no_grad():
            self._exec.run(sched)
        self.agg_eval_loss = self._aggregate_total_loss()
        writer = getattr(self, "summary_writer", None)
        if self.tensorboard_enabled() and self.global_rank == 0 and writer:
            writer.add_scalar("Train/Samples/eval_loss", self.agg_eval_loss.mean().item(), self.global_samples)
            writer.flush()
        if not return_logits:
            return self.agg_eval_loss
        outputs, self.outputs, self.eval_return_logits = self.outputs, None, False
        return self.agg_eval_loss, outputs

    def inference_batch(self, data_iter, layers_to_hook=None):
        with self._hooked(layers_to_hook):
            return self._inference_batch(data_iter)

    def _inference_batch(self, data_iter):
        """One forward-only micro-batch; (logits, presents) come back on every rank of the pipe group. Shapes may
        differ from call to call, so the activation shapes are renegotiated before and after."""
        self.module.eval()
        self.total_loss = None
        self._exec.reset_wire()
        sched = schedule.InferenceSchedule(micro_batches=1, stages=self.num_stages, stage_id=self.stage_id)
        self._inference_mode = True
        try:
            with self._borrowed_iterator(data_iter), torch.no_grad():
                self._exec.run(sched)
        finally:
            self._inference_mode = False
        logits = presents = None
        if self.is_last_stage():
            out = self.pipe_buffers["outputs"][0]
            if isinstance(out, (tuple, list)):
                logits, presents = out[0], (out[1] if len(out) > 1 else None)
            else:
                logits = out
        if self.is_pipe_parallel:
            src = self.grid.stage_to_global(self.num_stages - 1)
            logits = self._bcast_tensor(logits, src)
            flag = torch.tensor([int(presents is not None)], device=self.device)
            dist.broadcast(flag, src=src, group=self.grid.get_pipe_parallel_group())
            if int(flag.item()):
                presents = self._bcast_tensor(presents, src)
        self._exec.reset_wire()
        return logits, presents

    def _bcast_tensor(self, tensor, src):
        """Broadcast a tensor of unknown shape / dtype from `src` over the pipe group (shape header first)."""
        group = self.grid.get_pipe_parallel_group()
        header = torch.zeros(16, dtype=torch.long, device=self.device)
        if self.global_rank == src:
            desc = [tensor.dim(), _DTYPES.index(tensor.dtype)] + list(tensor.shape)
            header[:len(desc)] = torch.tensor(desc, device=self.device)
        dist.broadcast(header, src=src, group=group)
        if self.global_rank != src:
            h = header.tolist()
            tensor = torch.empty(h[2:2 + h[0]], dtype=_DTYPES[h[1]], device=self.device)
        tensor = tensor.contiguous()
        dist.broadcast(tensor, src=src, group=group)
        return tensor

    def is_first_stage(self):
        return self.stage_id == 0

    def is_last_stage(self):
        return self.stage_id == self.num_stages - 1

    # ------------------------------------------------------------------ loss
    def _aggregate_total_loss(self):
        """Mean over micro-batches (engine loss scaling) and data-parallel replicas on the last stage, then shared
        with the other stages. A module without loss_fn "returns" its last-stage output, which is a whole tensor and
        is broadcast with its shape (reference tests compare eval_batch() with the sequential forward)."""
        last = self.grid.stage_to_global(self.num_stages - 1)
        pipe_group = self.grid.get_pipe_parallel_group()
        agg = None
        if self.is_last_stage():
            mine = self._scale_loss(self.total_loss)
            self.dp_group_loss = mine.clone().detach()
            agg = self.dp_group_loss.clone()
            if self.is_data_parallel:
                dist.all_reduce(agg, group=self.mpu.get_data_parallel_group())
                agg /= self.dp_world_size
        if not self.is_pipe_parallel:
            return agg
        wide = torch.zeros(1, dtype=torch.long, device=self.device)
        if self.is_last_stage():
            wide[0] = int(agg.numel() > 1)
        dist.broadcast(wide, src=last, group=pipe_group)
        if int(wide.item()):
            return self._bcast_tensor(agg, last)
        pair = torch.zeros(2, dtype=torch.float32, device=self.device)
        if self.is_last_stage():
            pair = torch.stack([self.dp_group_loss.float().reshape(()), agg.float().reshape(())])
        dist.broadcast(pair, src=last, group=pipe_group)
        if not self.is_last_stage():
            self.dp_group_loss = pair[0].clone().detach()
        return pair[1].clone().detach()

    # ------------------------------------------------------------------ data plumbing
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

    def _next_batch(self):
        """Only tensor-parallel rank 0 pulls from the iterator; `batch_fn` may broadcast to the rest of its group."""
        tp_rank = self.grid.get_slice_parallel_rank() if self.is_model_parallel else 0
        batch = next(self.data_iterator) if tp_rank == 0 else None
        return self.batch_fn(batch) if self.batch_fn else batch

    def is_gradient_accumulation_boundary(self):
        """The schedule, not the micro-step counter, decides when an optimizer step happens."""
        return self._force_grad_boundary

    # ------------------------------------------------------------------ logging
    def log_for_device(self, *msg):
        if LOG_STAGE in (self.stage_id, -1) and DATA_PARALLEL_ID in (self.grid.data_parallel_id, -1):
            print(f"RANK={dist.get_rank()} PIPE-ID={self.stage_id} DATA-ID={self.grid.data_parallel_id} "
                  f"MBATCH-ID={self.microbatch_id} STEP-ID={self.log_batch_step_id} ::", *msg, flush=True)

    def tput_log(self, *msg):
        if self.global_rank == 0 and self.global_steps % self.steps_per_print() == 0:
            print(*msg)

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        return

    # ------------------------------------------------------------------ only the batch APIs are public
    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    # ------------------------------------------------------------------ checkpoints
    def module_state_dict(self):
        """One file per layer (written by data-parallel rank 0 of each stage); the model-states file keeps None."""
        assert isinstance(self.module, torch.nn.Module)
        assert self._curr_ckpt_path is not None, \
            "PipelineEngine expects module_state_dict() to be called from save_checkpoint()"
        self.module.save_state_dict(self._curr_ckpt_path)
        return None

    def load_module_state_dict(self, state_dict, strict=True):
        if state_dict is not None and not isinstance(state_dict, str):
            super().load_module_state_dict(state_dict, strict)
            return
        self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict)
