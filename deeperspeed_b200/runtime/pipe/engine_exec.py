"""Instruction interpreter of the PipelineEngine (mixin): one method per pipe instruction.
Reference: deepspeed/runtime/pipe/engine.py:654-1308.

Wire protocol: on the first SendActivation a stage ships ONE fixed-size int64 meta vector
[kind, n, (dtype_id, ndim, *shape) * n] (kind 0 = tensor, 2 = tuple); the receiver allocates its buffers once and
reuses them for every micro-batch. Gradients of the received activations travel back with the mirrored shapes.
(The reference sends type / count / ndims / shape as separate tiny messages, pipe/engine.py:826-937.)"""
import torch
import torch.distributed as dist

from ..utils import PartitionedTensor
from . import p2p
from . import schedule
from .module import PipelineError

_DTYPES = [torch.float16, torch.bfloat16, torch.float32, torch.float64, torch.int64, torch.int32, torch.int16,
           torch.int8, torch.uint8, torch.bool]
_META_LEN = 128


class PipeExecMixin:
    # ------------------------------------------------------------------ forward
    def _exec_forward_pass(self, buffer_id):
        self.tput_timer.start()
        self.mem_status("BEFORE FWD", reset_max=True)

        if isinstance(self.pipe_buffers["inputs"][buffer_id], tuple):
            inputs = tuple(t.clone() if torch.is_tensor(t) else t for t in self.pipe_buffers["inputs"][buffer_id])
        else:
            inputs = self.pipe_buffers["inputs"][buffer_id].clone()

        # rebuild activations that were scattered over the TP group for the send
        if self.is_pipe_partitioned and not self.is_first_stage():
            part_input = PartitionedTensor.from_meta(meta=inputs[0], local_part=inputs[1],
                                                     group=self.grid.get_slice_parallel_group())
            inputs = tuple([part_input.full(), *inputs[2:]]) if len(inputs) > 2 else part_input.full()
            if torch.is_tensor(inputs):
                inputs.requires_grad = True
            else:
                inputs[0].requires_grad = True
            self.pipe_buffers["inputs"][buffer_id] = inputs

        self._zero_grads(inputs)
        outputs = super(PipeExecMixin, self).forward(inputs)

        if self.is_pipe_partitioned and not self.is_last_stage():
            first = outputs[0] if isinstance(outputs, tuple) else outputs
            part = PartitionedTensor(tensor=first, group=self.grid.get_slice_parallel_group())
            # keep the full tensor alive for backward, send only meta + local slice
            self.pipe_buffers["output_tensors"][buffer_id] = first
            rest = outputs[1:] if isinstance(outputs, tuple) else ()
            outputs = tuple([part.to_meta(), part.data(), *rest])
            part = None

        self.pipe_buffers["outputs"][buffer_id] = outputs

        if self.is_last_stage():
            if getattr(self, "_inference_mode", False):
                return
            if self.loss_model is not None:
                labels = self.pipe_buffers["labels"][buffer_id]
                self.loss = self.loss_model(outputs, labels)
            else:
                self.loss = outputs
            if self.eval_return_logits:
                self.outputs = outputs
            if isinstance(self.loss, torch.Tensor):
                if self.total_loss is None:
                    self.total_loss = torch.zeros_like(self.loss)
                self.total_loss += self.loss.detach()
            else:
                if self.total_loss is None:
                    self.total_loss = [torch.zeros_like(l) for l in self.loss]
                for idx, l in enumerate(self.loss):
                    self.total_loss[idx] += l.detach()

    # ------------------------------------------------------------------ backward
    def _exec_backward_pass(self, buffer_id):
        assert self.optimizer is not None, "must provide optimizer during init in order to use backward"
        self.mem_status("BEFORE BWD", reset_max=True)

        if self.is_last_stage():
            # the last stage owns the loss: go through the engine so loss scaling applies
            super(PipeExecMixin, self).backward(self.loss)
            self.mem_status("AFTER BWD")
            return

        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.wall_clock_breakdown():
            self.timers("backward_microstep").start()
            self.timers("backward").start()
            self.timers("backward_inner_microstep").start()
            self.timers("backward_inner").start()

        if self.is_pipe_partitioned:
            # outputs = (meta, local part, ...): backward runs through the kept full tensor
            full = self.pipe_buffers["output_tensors"][buffer_id]
            outputs = tuple([full, *outputs[2:]]) if len(outputs) > 2 else full

        grad_tensors = self.grad_layer
        if isinstance(outputs, tuple):
            out_tensors = [t for t in outputs if torch.is_tensor(t) and t.is_floating_point() and t.requires_grad]
            assert len(out_tensors) == len(grad_tensors)
            torch.autograd.backward(tensors=out_tensors, grad_tensors=list(grad_tensors))
        else:
            torch.autograd.backward(tensors=(outputs,), grad_tensors=(grad_tensors[0],))

        self.pipe_buffers["output_tensors"][buffer_id] = None
        self.pipe_buffers["outputs"][buffer_id] = None
        if self.wall_clock_breakdown():
            self.timers("backward_inner").stop()
            self.timers("backward_inner_microstep").stop()
            self.timers("backward").stop()
            self.timers("backward_microstep").stop()
        self.mem_status("AFTER BWD")

    # ------------------------------------------------------------------ data
    def _to_device(self, x):
        if torch.is_tensor(x):
            return x.to(self.device, non_blocking=True)
        if isinstance(x, (tuple, list)):
            return type(x)(self._to_device(v) for v in x)
        return x

    def _exec_load_micro_batch(self, buffer_id):
        if self.wall_clock_breakdown():
            self.timers("batch_input").start()
        batch = self._next_batch()

        if self.is_first_stage():
            loaded = None
            if torch.is_tensor(batch[0]):
                loaded = batch[0].clone().to(self.device, non_blocking=True).detach()
                loaded.requires_grad = loaded.is_floating_point()
            else:
                assert isinstance(batch[0], (tuple, list))
                loaded = []
                for x in batch[0]:
                    assert torch.is_tensor(x)
                    mine = x.clone().detach().to(self.device, non_blocking=True)
                    mine.requires_grad = mine.is_floating_point()
                    loaded.append(mine)
                loaded = tuple(loaded)
            self.pipe_buffers["inputs"][buffer_id] = loaded

        if self.is_last_stage():
            loaded = batch[1]
            if torch.is_tensor(batch[1]):
                loaded = batch[1].to(self.device, non_blocking=True)
            elif isinstance(batch[1], (tuple, list)):
                loaded = tuple(x.to(self.device, non_blocking=True).detach() if torch.is_tensor(x) else x
                               for x in batch[1])
            self.pipe_buffers["labels"][buffer_id] = loaded
        if self.wall_clock_breakdown():
            self.timers("batch_input").stop()

    # ------------------------------------------------------------------ meta handshake
    def _send_tensor_meta(self, buffer, recv_stage):
        meta = [0] * _META_LEN
        tensors = [buffer] if torch.is_tensor(buffer) else list(buffer)
        meta[0] = 0 if torch.is_tensor(buffer) else 2
        meta[1] = len(tensors)
        pos = 2
        for t in tensors:
            assert torch.is_tensor(t), "pipeline stages exchange tensors or tuples of tensors"
            entry = [_DTYPES.index(t.dtype), t.dim()] + list(t.size())
            assert pos + len(entry) <= _META_LEN, "activation tuple too large for the meta message"
            meta[pos:pos + len(entry)] = entry
            pos += len(entry)
        p2p.send(torch.tensor(meta, dtype=torch.long, device=self.device), recv_stage)

    def _recv_tensor_meta(self, send_stage):
        meta_t = torch.zeros(_META_LEN, dtype=torch.long, device=self.device)
        p2p.recv(meta_t, send_stage)
        meta = meta_t.tolist()
        kind, count = meta[0], meta[1]
        pos = 2
        bufs = []
        for _ in range(count):
            dtype = _DTYPES[meta[pos]]
            ndim = meta[pos + 1]
            shape = meta[pos + 2:pos + 2 + ndim]
            pos += 2 + ndim
            if dtype == torch.bool:
                self.has_bool_tensors = True
            bufs.append(torch.zeros(shape, dtype=dtype, device=self.device))
        if kind == 0:
            return bufs[0]
        return tuple(bufs)

    # ------------------------------------------------------------------ activations
    def _exec_send_activations(self, buffer_id):
        if self.wall_clock_breakdown():
            self.timers("pipe_send_output").start()
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.first_output_send:
            self.first_output_send = False
            self._send_tensor_meta(outputs, self.next_stage)
        if isinstance(outputs, torch.Tensor):
            p2p.send(outputs, self.next_stage, fp32_comm=self.fp32_comm)
        elif isinstance(outputs, tuple):
            for buffer in outputs:
                if buffer.dtype == torch.bool:
                    # NCCL has no bool: ship as uint8
                    p2p.send(buffer.to(torch.uint8), self.next_stage)
                else:
                    p2p.send(buffer, self.next_stage, fp32_comm=self.fp32_comm)
        else:
            raise NotImplementedError(f"Could not send output of type {type(outputs)}")
        if self.wall_clock_breakdown():
            self.timers("pipe_send_output").stop()

    def _exec_recv_activations(self, buffer_id):
        if self.wall_clock_breakdown():
            self.timers("pipe_recv_input").start()
        recvd = None
        if self.pipe_recv_buf is None:
            self.pipe_recv_buf = self._recv_tensor_meta(self.prev_stage)
        if isinstance(self.pipe_recv_buf, torch.Tensor):
            p2p.recv(self.pipe_recv_buf, self.prev_stage, fp32_comm=self.fp32_comm)
            recvd = self.pipe_recv_buf.clone().detach()
            recvd.requires_grad = recvd.is_floating_point()
        else:
            assert isinstance(self.pipe_recv_buf, tuple)
            recvd = [None] * len(self.pipe_recv_buf)
            for idx, buffer in enumerate(self.pipe_recv_buf):
                assert torch.is_tensor(buffer)
                if buffer.dtype == torch.bool:
                    tmp = torch.empty(buffer.shape, dtype=torch.uint8, device=self.device)
                    p2p.recv(tmp, self.prev_stage)
                    recvd[idx] = tmp.bool()
                else:
                    p2p.recv(buffer, self.prev_stage, fp32_comm=self.fp32_comm)
                    recvd[idx] = buffer.clone().detach()
            recvd = tuple(recvd)
            for buffer in recvd:
                buffer.requires_grad = buffer.is_floating_point()
        self.pipe_buffers["inputs"][buffer_id] = recvd
        if self.wall_clock_breakdown():
            self.timers("pipe_recv_input").stop()

    # ------------------------------------------------------------------ gradients
    def _exec_send_grads(self, buffer_id):
        if self.wall_clock_breakdown():
            self.timers("pipe_send_grad").start()
        inputs = self.pipe_buffers["inputs"][buffer_id]
        if self.is_grad_partitioned:
            raise NotImplementedError("gradient partitioning over the TP group is not enabled")
        if isinstance(inputs, torch.Tensor):
            assert inputs.grad is not None
            p2p.send(inputs.grad, self.prev_stage, fp32_comm=self.fp32_comm)
        else:
            for idx, buffer in enumerate(inputs):
                # integer / bool tensors (ids, masks) carry no gradient
                if not buffer.is_floating_point():
                    assert buffer.grad is None
                    continue
                if self.is_pipe_partitioned and idx < 2:
                    continue
                assert buffer.grad is not None
                p2p.send(buffer.grad, self.prev_stage, fp32_comm=self.fp32_comm)
            if self.is_pipe_partitioned:
                # the rebuilt full activation is inputs[0] after forward re-assembly
                pass
        # the micro-batch is done on this stage: release its input buffer
        self.pipe_buffers["inputs"][buffer_id] = None
        if self.wall_clock_breakdown():
            self.timers("pipe_send_grad").stop()

    def _exec_recv_grads(self, buffer_id):
        if self.wall_clock_breakdown():
            self.timers("pipe_recv_grad").start()
        outputs = self.pipe_buffers["outputs"][buffer_id]
        if self.is_pipe_partitioned:
            full = self.pipe_buffers["output_tensors"][buffer_id]
            outputs = tuple([full, *outputs[2:]]) if len(outputs) > 2 else full
        # gradient buffers mirror the floating-point outputs; allocated once
        if self.grad_layer is None:
            if isinstance(outputs, torch.Tensor):
                self.grad_layer = [torch.zeros(outputs.size(), dtype=outputs.dtype, device=self.device)]
            else:
                self.grad_layer = [torch.zeros(t.size(), dtype=t.dtype, device=self.device)
                                   for t in outputs if torch.is_tensor(t) and t.is_floating_point() and t.requires_grad]
        for buffer in self.grad_layer:
            p2p.recv(buffer, self.next_stage, fp32_comm=self.fp32_comm)
        if self.wall_clock_breakdown():
            self.timers("pipe_recv_grad").stop()

    # ------------------------------------------------------------------ optimizer
    def _exec_optimizer_step(self, lr_kwargs=None):
        if self.wall_clock_breakdown():
            self.timers("step_microstep").start()
            self.timers("step").start()
        self.mem_status("BEFORE STEP", reset_max=True)
        self._force_grad_boundary = True
        self._take_model_step(lr_kwargs)
        self._force_grad_boundary = False
        self.mem_status("AFTER STEP")
        if self.tensorboard_enabled() and self.global_rank == 0 and getattr(self, "summary_writer", None):
            self.summary_writer.add_scalar("Train/Samples/lr", self.get_lr()[0], self.global_samples)
            if self.fp16_enabled() and hasattr(self.optimizer, "cur_scale"):
                self.summary_writer.add_scalar("Train/Samples/loss_scale", self.optimizer.cur_scale, self.global_samples)
        if self.wall_clock_breakdown():
            self.timers("step_microstep").stop()
            self.timers("step").stop()

    def _zero_grads(self, inputs):
        if isinstance(inputs, torch.Tensor):
            if inputs.grad is not None:
                inputs.grad.data.zero_()
        else:
            for t in inputs:
                if torch.is_tensor(t) and t.grad is not None:
                    t.grad.data.zero_()

    # ------------------------------------------------------------------ API overrides
    def forward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def backward(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def step(self, *args, **kwargs):
        raise PipelineError("Only train_batch() is accessible in pipeline mode.")

    def mem_status(self, msg, print_rank=-1, reset_max=False):
        return

    # ------------------------------------------------------------------ checkpoints
    def module_state_dict(self):
        """One file per layer (written by data-parallel rank 0 of each stage); the model-states file keeps None."""
        assert isinstance(self.module, torch.nn.Module)
        assert self._curr_ckpt_path is not None, "PipelineEngine expects module_state_dict() to be called from save_checkpoint()"
        self.module.save_state_dict(self._curr_ckpt_path)
        return None

    def load_module_state_dict(self, state_dict, strict=True):
        if (state_dict is not None) and (not isinstance(state_dict, str)):
            super(PipeExecMixin, self).load_module_state_dict(state_dict, strict)
            return
        self.module.load_state_dir(load_dir=self._curr_ckpt_path, strict=strict)

    # ------------------------------------------------------------------ interpreter
    _INSTRUCTION_MAP = {
        schedule.OptimizerStep: _exec_optimizer_step,
        schedule.ReduceGrads: lambda self: self._exec_reduce_grads(),
        schedule.ReduceTiedGrads: lambda self: self._exec_reduce_tied_grads(),
        schedule.LoadMicroBatch: _exec_load_micro_batch,
        schedule.ForwardPass: _exec_forward_pass,
        schedule.BackwardPass: _exec_backward_pass,
        schedule.SendActivation: _exec_send_activations,
        schedule.RecvActivation: _exec_recv_activations,
        schedule.SendGrad: _exec_send_grads,
        schedule.RecvGrad: _exec_recv_grads,
    }

    def _exec_schedule(self, pipe_schedule):
        self._reserve_pipe_buffers(pipe_schedule.num_pipe_buffers())
        for step_cmds in pipe_schedule:
            for cmd in step_cmds:
                fn = self._INSTRUCTION_MAP.get(type(cmd))
                if fn is None:
                    raise RuntimeError(f"{self.__class__.__name__} does not understand instruction {repr(cmd)}")
                if isinstance(cmd, schedule.ReduceGrads) and self.wall_clock_breakdown():
                    self.timers("reduce_grads").start()
                    fn(self, **cmd.kwargs)
                    self.timers("reduce_grads").stop()
                elif isinstance(cmd, schedule.ReduceTiedGrads) and self.wall_clock_breakdown():
                    self.timers("reduce_tied_grads").start()
                    fn(self, **cmd.kwargs)
                    self.timers("reduce_tied_grads").stop()
                else:
                    fn(self, **cmd.kwargs)

