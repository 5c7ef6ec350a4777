"""Pipeline instruction set and schedules (1F1B training, forward-only inference, plain gradient accumulation).

The instruction names, their `buffer_id` argument and the per-step streams are the contract of reference
deepspeed/runtime/pipe/schedule.py:129-474 (golden table in SURVEY C.4; checked in tests/test_pure_logic.py). They
are produced here from a clock model instead of the reference's even/odd helper cascade:

  tick `t`, stage `p` of `P`:  (t - p) even -> forward slot of micro-batch (t - p) / 2
                               (t - p) odd  -> backward slot of micro-batch (t + p + 1) / 2 - P

A slot whose micro-batch index falls outside [0, M) is idle (pipeline fill / drain). Activations produced in a
forward slot travel at the next tick, gradients produced in a backward slot likewise, so every send has its
matching receive in the same tick of the neighbour stage. The engine does not interpret these lists one by one:
`executor.compile_plan` fuses the communication of a tick into one bidirectional exchange.
"""
from abc import ABC, abstractmethod
from collections import namedtuple

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


# one slot of the clock model: micro-batch index (possibly out of range = idle) and its direction
Slot = namedtuple("Slot", ["micro_batch", "forward"])


class PipeSchedule(ABC):
    """Generator of per-tick instruction lists for one stage."""

    def __init__(self, micro_batches, stages, stage_id):
        super().__init__()
        self.micro_batches = micro_batches
        self.stages = stages
        self.stage_id = stage_id
        self.prev_stage = stage_id - 1
        self.next_stage = stage_id + 1

    @abstractmethod
    def steps(self):
        """Yield one list of instructions per tick."""

    def num_pipe_buffers(self):
        return self.micro_batches

    # ---- predicates / aliases the reference exposes
    def _valid_micro_batch(self, micro_batch_id):
        return 0 <= micro_batch_id < self.micro_batches

    def _valid_stage(self, stage_id):
        return 0 <= stage_id < self.stages

    stage = property(lambda self: self.stage_id)
    num_stages = property(lambda self: self.stages)
    num_micro_batches = property(lambda self: self.micro_batches)
    is_first_stage = property(lambda self: self.stage_id == 0)
    is_last_stage = property(lambda self: self.stage_id == self.stages - 1)

    def _buffer_idx(self, micro_batch_id):
        assert self._valid_micro_batch(micro_batch_id)
        return micro_batch_id % self.num_pipe_buffers()

    def _buf_or_none(self, micro_batch_id):
        return self._buffer_idx(micro_batch_id) if self._valid_micro_batch(micro_batch_id) else None

    def __iter__(self):
        self.it = None
        return self

    def __next__(self):
        if self.it is None:
            self.it = self.steps()
        return next(self.it)


class InferenceSchedule(PipeSchedule):
    """Forward-only pipelining over two alternating buffers: stage p works on micro-batch t - p at tick t, ships the
    previous tick's output and receives this tick's input. Even stages send before they receive, odd stages the
    other way round, so blocking transports pair up as well."""

    def steps(self):
        has_prev, has_next = self.stage_id > 0, self.stage_id < self.stages - 1
        flip = self.stage_id & 1
        for tick in range(self.micro_batches + self.stages - 1):
            mb = tick - self.stage_id
            work = self._valid_micro_batch(mb)
            in_buf = (tick + flip) & 1
            out_buf = in_buf ^ 1
            cmds = []
            if work and (self.is_first_stage or self.is_last_stage):
                cmds.append(LoadMicroBatch(in_buf))
            tx = SendActivation(out_buf) if has_next and self._valid_micro_batch(mb - 1) else None
            rx = RecvActivation(in_buf) if has_prev and work else None
            cmds += [c for c in ((rx, tx) if flip else (tx, rx)) if c is not None]
            if work:
                cmds.append(ForwardPass(in_buf))
            yield cmds

    def num_pipe_buffers(self):
        return 2


class TrainSchedule(PipeSchedule):
    """1F1B: once the pipeline is full every stage alternates one forward and one backward slot, so the number of live
    activations is bounded by the pipeline depth, not by the number of micro-batches."""

    def slot(self, tick):
        """The clock model of the module docstring."""
        d = tick - self.stage_id
        if d % 2 == 0:
            return Slot(d // 2, True)
        return Slot((tick + self.stage_id + 1) // 2 - self.stages, False)

    def _step_to_micro_batch(self, step_id):
        s = self.slot(step_id)
        return s.micro_batch, s.forward

    def steps(self):
        has_prev, has_next = self.stage_id > 0, self.stage_id < self.stages - 1
        edge_stage = not (has_prev and has_next)
        n_ticks = 2 * (self.micro_batches + self.stages - 1)
        before = None            # buffer worked on in the previous tick (None: idle)
        for tick in range(n_ticks):
            slot = self.slot(tick)
            now = self._buf_or_none(slot.micro_batch)
            cmds = []
            if slot.forward:
                # input for this forward comes from upstream; the gradient finished last tick goes back upstream
                if has_prev and now is not None:
                    cmds.append(RecvActivation(now))
                if has_prev and before is not None:
                    cmds.append(SendGrad(before))
                if edge_stage and now is not None:
                    cmds.append(LoadMicroBatch(now))
                if now is not None:
                    cmds.append(ForwardPass(now))
            else:
                # the activation finished last tick goes downstream; the gradient for this backward comes from there
                if has_next and before is not None:
                    cmds.append(SendActivation(before))
                if has_next and now is not None:
                    cmds.append(RecvGrad(now))
                if now is not None:
                    cmds.append(BackwardPass(now))
            if tick == n_ticks - 1:
                cmds += [ReduceTiedGrads(), ReduceGrads(), OptimizerStep()]
            before = now
            yield cmds

    def num_pipe_buffers(self):
        """In-flight micro-batches a stage must hold: deeper for early stages, never fewer than 2."""
        return max(2, min(self.stages - self.stage_id + 1, self.micro_batches))


class DataParallelSchedule(PipeSchedule):
    """Plain gradient accumulation expressed in the pipeline ISA (single stage)."""

    def steps(self):
        for mb in range(self.micro_batches):
            cmds = [LoadMicroBatch(buffer_id=0), ForwardPass(buffer_id=0), BackwardPass(buffer_id=0)]
            if mb == self.micro_batches - 1:
                cmds += [ReduceGrads(), OptimizerStep()]
            yield cmds

    def num_pipe_buffers(self):
        return 1


def _is_even(x):
    return x % 2 == 0


def _is_odd(x):
    return x % 2 != 0
