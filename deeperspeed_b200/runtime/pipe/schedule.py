"""Pipeline instruction set and schedules (1F1B training, inference, pure data-parallel).

Instruction streams are identical to reference deepspeed/runtime/pipe/schedule.py:129-314 (golden table in
SURVEY C.4); they are generated from closed-form (step, stage) -> micro-batch maps rather than by the reference's
even/odd helper cascade."""
from abc import ABC, abstractmethod

from ..utils import call_to_str


class PipeInstruction:
    """Base class of everything the pipeline engine executes; kwargs become attributes."""

    def __init__(self, **kwargs):
        self.name = self.__class__.__name__
        self.kwargs = kwargs
        for key, val in kwargs.items():
            setattr(self, key, val)

    def __repr__(self):
        return call_to_str(self.name, **self.kwargs)

    def __eq__(self, other):
        return type(self) is type(other) and self.kwargs == other.kwargs

    def __hash__(self):
        return hash((self.name, tuple(sorted(self.kwargs.items()))))


class OptimizerStep(PipeInstruction):
    """Step the optimizer and zero gradients (issued after ReduceGrads / ReduceTiedGrads)."""


class ReduceGrads(PipeInstruction):
    """Reduce gradients among the data-parallel replicas of a stage."""


class ReduceTiedGrads(PipeInstruction):
    """Reduce gradients of tied modules over the stages that share them."""


class BufferOpInstruction(PipeInstruction):
    """An instruction operating on pipeline buffer `buffer_id`."""

    def __init__(self, buffer_id, **kwargs):
        super().__init__(buffer_id=buffer_id, **kwargs)


class LoadMicroBatch(BufferOpInstruction):
    """buffers['inputs'][buffer_id] = next(data_iter)"""


class ForwardPass(BufferOpInstruction):
    """buffers['outputs'][buffer_id] = forward(buffers['inputs'][buffer_id])"""


class BackwardPass(BufferOpInstruction):
    """backward through buffers['outputs'][buffer_id] with buffers['gradients'][buffer_id]"""


class SendActivation(BufferOpInstruction):
    """send(buffers['outputs'][buffer_id]) to the next stage"""


class RecvActivation(BufferOpInstruction):
    """buffers['inputs'][buffer_id] = recv() from the previous stage"""


class SendGrad(BufferOpInstruction):
    """send(buffers['inputs'][buffer_id].grad) to the previous stage"""


class RecvGrad(BufferOpInstruction):
    """buffers['gradients'][buffer_id] = recv() from the next stage"""


class PipeSchedule(ABC):
    """Generator of per-step instruction lists for one stage. Every step is barrier-safe: a send on one stage is
    matched by the receive on its neighbour in the same step."""

    def __init__(self, micro_batches, stages, stage_id):
        super().__init__()
        self.micro_batches = micro_batches
        self.stages = stages
        self.stage_id = stage_id
        self.prev_stage = self.stage_id - 1
        self.next_stage = self.stage_id + 1

    @abstractmethod
    def steps(self):
        pass

    def num_pipe_buffers(self):
        return self.micro_batches

    def _valid_micro_batch(self, micro_batch_id):
        return 0 <= micro_batch_id < self.micro_batches

    def _valid_stage(self, stage_id):
        return 0 <= stage_id < self.stages

    @property
    def stage(self):
        return self.stage_id

    @property
    def num_stages(self):
        return self.stages

    @property
    def num_micro_batches(self):
        return self.micro_batches

    @property
    def is_first_stage(self):
        return self.stage_id == 0

    @property
    def is_last_stage(self):
        return self.stage_id == self.stages - 1

    def _buffer_idx(self, micro_batch_id):
        assert self._valid_micro_batch(micro_batch_id)
        return micro_batch_id % self.num_pipe_buffers()

    def __iter__(self):
        self.it = None
        return self

    def __next__(self):
        if self.it is None:
            self.it = self.steps()
        return next(self.it)


class InferenceSchedule(PipeSchedule):
    """Forward-only pipelining with two alternating buffers."""

    def steps(self):
        total_steps = self.micro_batches + self.stages - 1
        for step_id in range(total_steps):
            cmds = []
            micro_batch_id = step_id - self.stage_id
            # alternate send/recv buffers; even stages use (step % 2) to receive, odd stages the other one
            if self.stage_id % 2 == 0:
                recv_buf, send_buf = step_id % 2, (step_id + 1) % 2
            else:
                recv_buf, send_buf = (step_id + 1) % 2, step_id % 2

            if (self.is_first_stage or self.is_last_stage) and self._valid_micro_batch(micro_batch_id):
                cmds.append(LoadMicroBatch(recv_buf))

            send = SendActivation(send_buf) if (self._valid_stage(self.next_stage)
                                                and self._valid_micro_batch(micro_batch_id - 1)) else None
            recv = RecvActivation(recv_buf) if (self._valid_stage(self.prev_stage)
                                                and self._valid_micro_batch(micro_batch_id)) else None
            # even stages send first, odd stages receive first (pairs up blocking p2p without deadlock)
            order = (send, recv) if self.stage_id % 2 == 0 else (recv, send)
            cmds.extend(c for c in order if c is not None)

            if self._valid_micro_batch(micro_batch_id):
                cmds.append(ForwardPass(recv_buf))
            yield cmds

    def num_pipe_buffers(self):
        return 2


class TrainSchedule(PipeSchedule):
    """1F1B: each stage alternates forward and backward steps once the pipeline is full, which bounds the number
    of live activations by the pipeline depth instead of the number of micro-batches."""

    def _step_to_micro_batch(self, step_id):
        """(micro_batch_id, is_forward) for this stage at `step_id` (ids may be out of range = idle slot)."""
        same_parity = (step_id % 2) == (self.stage_id % 2)
        if same_parity:
            # forward slot
            base = step_id // 2 if step_id % 2 == 0 else (step_id - 1) // 2
            return base - self.stage_id // 2, True
        if step_id % 2 == 0:  # even step on an odd stage
            return step_id // 2 - self.stages + (self.stage_id + 1) // 2, False
        return (step_id - 1) // 2 - self.stages + 1 + self.stage_id // 2, False

    def steps(self):
        prev_micro_batch_id = -1
        total_steps = 2 * (self.micro_batches + self.stages - 1)
        for step_id in range(total_steps):
            micro_batch_id, is_forward = self._step_to_micro_batch(step_id)
            prev_buffer = self._buffer_idx(prev_micro_batch_id) if self._valid_micro_batch(prev_micro_batch_id) else None
            curr_buffer = self._buffer_idx(micro_batch_id) if self._valid_micro_batch(micro_batch_id) else None
            cmds = []

            # exchange activations / gradients produced by the previous step with the neighbours
            if is_forward:
                if self._valid_micro_batch(micro_batch_id) and self._valid_stage(self.prev_stage):
                    cmds.append(RecvActivation(curr_buffer))
                if self._valid_micro_batch(prev_micro_batch_id) and self._valid_stage(self.prev_stage):
                    cmds.append(SendGrad(prev_buffer))
            else:
                if self._valid_micro_batch(prev_micro_batch_id) and self._valid_stage(self.next_stage):
                    cmds.append(SendActivation(prev_buffer))
                if self._valid_micro_batch(micro_batch_id) and self._valid_stage(self.next_stage):
                    cmds.append(RecvGrad(curr_buffer))

            if (self.stage_id == 0 or self.stage_id == self.stages - 1) and is_forward \
                    and self._valid_micro_batch(micro_batch_id):
                cmds.append(LoadMicroBatch(curr_buffer))

            if self._valid_micro_batch(micro_batch_id):
                cmds.append(ForwardPass(curr_buffer) if is_forward else BackwardPass(curr_buffer))

            if step_id == total_steps - 1:
                cmds.append(ReduceTiedGrads())
                cmds.append(ReduceGrads())
                cmds.append(OptimizerStep())

            prev_micro_batch_id = micro_batch_id
            yield cmds

    def num_pipe_buffers(self):
        """In-flight micro-batches a stage must hold: deeper for early stages, never fewer than 2."""
        buffers = min(self.stages - self.stage_id + 1, self.micro_batches)
        return max(2, buffers)


class DataParallelSchedule(PipeSchedule):
    """Plain gradient accumulation expressed in the pipeline ISA (single stage)."""

    def steps(self):
        for step_id in range(self.micro_batches):
            cmds = [LoadMicroBatch(buffer_id=0), ForwardPass(buffer_id=0), BackwardPass(buffer_id=0)]
            if step_id == self.micro_batches - 1:
                cmds.extend([ReduceGrads(), OptimizerStep()])
            yield cmds

    def num_pipe_buffers(self):
        return 1


def _is_even(x):
    return x % 2 == 0


def _is_odd(x):
    return x % 2 != 0
