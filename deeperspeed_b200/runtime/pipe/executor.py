"""Stage executor of the PipelineEngine.

What the reference does by interpreting one instruction at a time with blocking transfers (reference
deepspeed/runtime/pipe/engine.py:654-1316) is organised here as a compiled plan:

  * `compile_plan` folds every tick of a schedule into one record: the transfers of the tick (both directions), the
    micro-batch to load, the forward or backward to run, and the end-of-batch tail. A plan is built once per
    (schedule kind, slot count) and replayed.
  * All transfers of a tick are posted together on the per-direction links of `p2p.StageLinks` (receives first, then
    sends) and only the receives are awaited before the compute that needs them; sends are collected one tick
    later. An activation going down and a gradient coming up therefore overlap instead of queueing behind each other.
  * Receive buffers belong to a pipeline slot and are handed to autograd as detached views, so nothing is cloned.
  * Wire format: the first activation transfer is preceded by ONE int64 meta message
    [n, (dtype, ndim, *shape) * n]; afterwards both ends know every shape. Gradients mirror the floating-point
    activations. With tensor parallelism the first activation (and, if enabled, the first gradient) crosses the
    boundary as (partition meta, local slice) and is re-assembled over the slice group on the other side.
  * Bubble trimming (`pipeline: {"edge_split": k}`): the first and the last micro-batch of a batch are fed as k
    slices of 1/k of the rows each. The 1F1B fill and drain then cost (stages - 1) / k micro-batch times instead of
    (stages - 1); gradients are unchanged because every slice's loss is weighted by its share of the micro-batch.
"""
from collections import namedtuple

import torch
import torch.distributed as dist

from ..utils import PartitionedTensor, partition_uniform
from . import p2p
from . import schedule as isa

Tick = namedtuple("Tick", ["recvs", "sends", "load", "compute", "tail"])

_DTYPES = [torch.float16, torch.bfloat16, torch.float32, torch.float64, torch.int64, torch.int32, torch.int16,
           torch.int8, torch.uint8, torch.bool]
_META_LEN = 128

_KIND = {isa.RecvActivation: ("recv", "act"), isa.RecvGrad: ("recv", "grad"),
         isa.SendActivation: ("send", "act"), isa.SendGrad: ("send", "grad")}
_TAIL = {isa.ReduceTiedGrads: "reduce_tied", isa.ReduceGrads: "reduce", isa.OptimizerStep: "step"}


def compile_plan(sched):
    """Per-tick records of a schedule (see module docstring). Unknown instructions are rejected here, once."""
    plan = []
    for cmds in sched:
        recvs, sends, load, compute, tail = [], [], None, None, []
        for c in cmds:
            t = type(c)
            if t in _KIND:
                (recvs if _KIND[t][0] == "recv" else sends).append((_KIND[t][1], c.buffer_id))
            elif t is isa.LoadMicroBatch:
                load = c.buffer_id
            elif t is isa.ForwardPass:
                compute = ("fwd", c.buffer_id)
            elif t is isa.BackwardPass:
                compute = ("bwd", c.buffer_id)
            elif t in _TAIL:
                tail.append(_TAIL[t])
            else:
                raise RuntimeError(f"pipeline executor does not understand instruction {c!r}")
        plan.append(Tick(tuple(recvs), tuple(sends), load, compute, tuple(tail)))
    return plan


def slot_rows(micro_batches, micro_batch_size, edge_split):
    """Rows of every schedule slot: the first and last micro-batch cut into `edge_split` slices, the rest whole."""
    k = edge_split
    if k <= 1 or micro_batches < 2 or micro_batch_size % k:
        return [micro_batch_size] * micro_batches
    part = micro_batch_size // k
    return [part] * k + [micro_batch_size] * (micro_batches - 2) + [part] * k


def _split_rows(x, k):
    """k row slices of a tensor or (nested) tuple of tensors; non-tensors are repeated."""
    if torch.is_tensor(x):
        return list(torch.chunk(x, k, dim=0))
    if isinstance(x, (tuple, list)):
        cols = [_split_rows(v, k) for v in x]
        return [type(x)(c[i] for c in cols) for i in range(k)]
    return [x] * k


class SlotFeed:
    """Hands the data iterator's micro-batches to the schedule slots, slicing the edge micro-batches."""

    def __init__(self, rows, micro_batch_size):
        self.rows = rows
        self.mbs = micro_batch_size
        self.queue = []
        self.slot = 0

    def reset(self):
        self.queue.clear()
        self.slot = 0

    def next(self, fetch):
        if not self.queue:
            batch = fetch()
            k = self.mbs // self.rows[self.slot]
            self.queue = _split_rows(batch, k) if (k > 1 and batch is not None) else [batch] * k
        self.slot += 1
        return self.queue.pop(0)


class StageExecutor:
    """Runs compiled plans for one pipeline stage. `engine` is the PipelineEngine (for the module, the optimizer step,
    the data iterator, the grid and the timers); the executor owns the slot buffers and the wire state."""

    def __init__(self, engine):
        self.e = engine
        self.links = p2p.links() if engine.is_pipe_parallel else None
        self.plans = {}
        self.buffers = {"inputs": [], "labels": [], "outputs": [], "output_tensors": [], "wire": [], "grads": []}
        self.pending = []          # sends in flight: (work, tensor kept alive)
        self.rx = {}               # (kind, slot buffer, rows) -> raw receive tensors, allocated once
        self.rows_of = {}          # slot buffer -> rows of the micro-batch (slice) it currently holds
        self.reset_wire()

    # ------------------------------------------------------------------ plan / buffers
    def reset_wire(self):
        """Forget the negotiated activation shapes (a call with different shapes renegotiates them)."""
        self.announced = False
        self.act_spec = None       # (is_tuple, [(dtype, shape), ...]) of what the previous stage sends
        self.rx.clear()

    def reserve(self, n):
        for lst in self.buffers.values():
            if len(lst) < n:
                lst.extend([None] * (n - len(lst)))

    def plan_for(self, sched):
        key = (type(sched), sched.micro_batches, sched.stages, sched.stage_id)
        plan = self.plans.get(key)
        if plan is None:
            plan = self.plans[key] = compile_plan(sched)
        return plan

    def run(self, sched, rows=None):
        e = self.e
        self.reserve(sched.num_pipe_buffers())
        self.rows = rows if rows is not None else [e.micro_batch_size] * sched.micro_batches
        self.feed = SlotFeed(self.rows, e.micro_batch_size)
        self.fwd_count = self.bwd_count = 0
        for tick in self.plan_for(sched):
            if tick.recvs or tick.sends:
                self.exchange(tick)
            if tick.load is not None:
                self.load(tick.load)
            if tick.compute is not None:
                kind, buf = tick.compute
                if kind == "fwd":
                    self.forward(buf)
                    self.fwd_count += 1
                else:
                    self.backward(buf)
                    self.bwd_count += 1
            for name in tick.tail:
                self.tail(name)
        self.collect_sends()

    # ------------------------------------------------------------------ helpers
    def _timer(self, name, start):
        if self.e.wall_clock_breakdown():
            (self.e.timers(name).start if start else self.e.timers(name).stop)()

    def _rows_now(self, kind):
        i = self.fwd_count if kind == "act" else self.bwd_count
        return self.rows[i] if i < len(self.rows) else self.e.micro_batch_size

    def _scaled(self, shape, rows):
        mbs = self.e.micro_batch_size
        if rows == mbs or not shape:
            return list(shape)
        assert shape[0] % mbs == 0, "edge_split needs the micro-batch rows in dim 0 of every tensor crossing a stage"
        return [shape[0] * rows // mbs] + list(shape[1:])

    def _wire_dtype(self, dtype):
        if dtype == torch.bool:
            return torch.uint8
        if self.e.fp32_comm and dtype in (torch.float16, torch.bfloat16):
            return torch.float32
        return dtype

    def _to_wire(self, t):
        wd = self._wire_dtype(t.dtype)
        return (t if wd == t.dtype else t.to(wd)).contiguous()

    def _from_wire(self, raw, dtype):
        t = raw if raw.dtype == dtype else raw.to(dtype)
        t = t.detach()
        if t.is_floating_point():
            t.requires_grad_()
        return t

    def _rx_bufs(self, kind, buf, rows, specs):
        key = (kind, buf, rows)
        bufs = self.rx.get(key)
        if bufs is None:
            bufs = self.rx[key] = [torch.empty(self._scaled(shape, rows) if scale else list(shape),
                                               dtype=self._wire_dtype(dtype), device=self.e.device)
                                   for dtype, shape, scale in specs]
        return bufs

    def collect_sends(self):
        for work, _keep in self.pending:
            work.wait()
        self.pending.clear()

    def _post_send(self, link, t):
        w = self._to_wire(t)
        self.pending.append((link.post(w), w))

    # ------------------------------------------------------------------ transfers of one tick
    def exchange(self, tick):
        self.collect_sends()                      # last tick's sends were matched one compute step ago
        waits, arrived = [], []
        for kind, buf in tick.recvs:
            name = "pipe_recv_input" if kind == "act" else "pipe_recv_grad"
            self._timer(name, True)
            raws = self._post_recv_act(buf) if kind == "act" else self._post_recv_grad(buf)
            waits += raws[0]
            arrived.append((kind, buf, raws[1]))
            self._timer(name, False)
        for kind, buf in tick.sends:
            name = "pipe_send_output" if kind == "act" else "pipe_send_grad"
            self._timer(name, True)
            (self._send_act if kind == "act" else self._send_grad)(buf)
            self._timer(name, False)
        for w in waits:
            w.wait()
        for kind, buf, raws in arrived:
            (self._land_act if kind == "act" else self._land_grad)(buf, raws)

    # ---- activations
    def _send_act(self, buf):
        link = self.links.act_out
        wire = self.buffers["wire"][buf]
        if not self.announced:
            self.announced = True
            meta = [0] * _META_LEN
            meta[0] = len(wire)
            pos = 1
            for t in wire:
                assert torch.is_tensor(t), "pipeline stages exchange tensors or tuples of tensors"
                entry = [_DTYPES.index(t.dtype), t.dim()] + list(t.size())
                assert pos + len(entry) <= _META_LEN, "activation tuple too large for the meta message"
                meta[pos:pos + len(entry)] = entry
                pos += len(entry)
            meta[-1] = int(self.buffers["outputs"][buf] is not None and isinstance(self.buffers["outputs"][buf], tuple))
            # the announced shapes are those of a WHOLE micro-batch
            rows, mbs = self._rows_of_send(buf), self.e.micro_batch_size
            if rows != mbs:
                pos = 1
                for t in wire:
                    if t.dim() > 0:
                        meta[pos + 2] = t.size(0) * mbs // rows
                    pos += 2 + t.dim()
            self._post_send(link, torch.tensor(meta, dtype=torch.long, device=self.e.device))
        for t in wire:
            self._post_send(link, t)

    def _rows_of_send(self, buf):
        return self.rows_of.get(buf, self.e.micro_batch_size)   # remembered per slot by `forward`

    def _post_recv_act(self, buf):
        link = self.links.act_in
        if self.act_spec is None:
            meta_t = torch.zeros(_META_LEN, dtype=torch.long, device=self.e.device)
            link.post(meta_t).wait()
            meta = meta_t.tolist()
            specs, pos = [], 1
            for _ in range(meta[0]):
                dtype, ndim = _DTYPES[meta[pos]], meta[pos + 1]
                specs.append((dtype, meta[pos + 2:pos + 2 + ndim]))
                pos += 2 + ndim
                if dtype == torch.bool:
                    self.e.has_bool_tensors = True
            self.act_spec = (bool(meta[-1]), specs)
        partitioned = self.e.is_pipe_partitioned
        specs = [(d, s, not (partitioned and i < 2)) for i, (d, s) in enumerate(self.act_spec[1])]
        raws = self._rx_bufs("act", buf, self._rows_now("act"), specs)
        return [link.post(t) for t in raws], raws

    def _land_act(self, buf, raws):
        is_tuple, specs = self.act_spec
        ts = tuple(self._from_wire(raw, dtype) for raw, (dtype, _) in zip(raws, specs))
        if self.e.is_pipe_partitioned and len(ts) > 1 and ts[1].is_floating_point():
            ts[1].requires_grad_(False)       # the local slice; the re-assembled tensor is the autograd leaf
        self.buffers["inputs"][buf] = ts if (is_tuple or len(ts) > 1) else ts[0]

    # ---- gradients
    def _grad_targets(self, buf):
        """The tensors of this slot's forward output that receive a gradient from downstream."""
        outs = self.buffers["outputs"][buf]
        outs = outs if isinstance(outs, tuple) else (outs,)
        return [t for t in outs if torch.is_tensor(t) and t.is_floating_point() and t.requires_grad]

    def _send_grad(self, buf):
        link = self.links.grad_out
        inputs = self.buffers["inputs"][buf]
        inputs = inputs if isinstance(inputs, tuple) else (inputs,)
        grads = []
        for t in inputs:
            if not (torch.is_tensor(t) and t.is_floating_point() and t.requires_grad):
                continue                      # ids, masks and partition metadata carry no gradient
            assert t.grad is not None, "a floating-point activation received no gradient"
            grads.append(t.grad)
        if self.e.is_grad_partitioned and grads:
            part = PartitionedTensor(tensor=grads[0], group=self.e.grid.get_slice_parallel_group())
            grads = [part.to_meta(), part.data()] + grads[1:]
        for g in grads:
            self._post_send(link, g)
        self.buffers["inputs"][buf] = None    # the micro-batch is finished on this stage

    def _post_recv_grad(self, buf):
        link = self.links.grad_in
        targets = self._grad_targets(buf)
        specs = [(t.dtype, list(t.size()), False) for t in targets]
        if self.e.is_grad_partitioned and targets:
            group = self.e.grid.get_slice_parallel_group()
            parts, me = dist.get_world_size(group=group), dist.get_rank(group=group)
            first = targets[0]
            cuts = partition_uniform(num_items=first.numel(), num_parts=parts)
            specs = [(torch.long, [first.dim() + parts + 4], False),
                     (first.dtype, [cuts[me + 1] - cuts[me]], False)] + specs[1:]
        raws = self._rx_bufs("grad", buf, self._rows_now("grad"), specs)
        return [link.post(t) for t in raws], raws

    def _land_grad(self, buf, raws):
        targets = self._grad_targets(buf)
        if self.e.is_grad_partitioned and targets:
            part = PartitionedTensor.from_meta(meta=raws[0], local_part=raws[1].to(targets[0].dtype),
                                               group=self.e.grid.get_slice_parallel_group())
            raws = [part.full()] + list(raws[2:])
        grads = [r if r.dtype == t.dtype else r.to(t.dtype) for r, t in zip(raws, targets)]
        self.buffers["grads"][buf] = grads

    # ------------------------------------------------------------------ data
    def _to_device(self, x, want_grad):
        if torch.is_tensor(x):
            t = x.to(self.e.device, non_blocking=True)
            if want_grad and t.is_floating_point():
                t = t.detach().requires_grad_()
            return t
        if isinstance(x, (tuple, list)):
            return tuple(self._to_device(v, want_grad) for v in x)
        return x

    def load(self, buf):
        e = self.e
        self._timer("batch_input", True)
        batch = self.feed.next(e._next_batch)
        if e.is_first_stage():
            assert torch.is_tensor(batch[0]) or isinstance(batch[0], (tuple, list))
            self.buffers["inputs"][buf] = self._to_device(batch[0], True)
        if e.is_last_stage():
            self.buffers["labels"][buf] = self._to_device(batch[1], False)
        self._timer("batch_input", False)

    # ------------------------------------------------------------------ compute
    def forward(self, buf):
        e = self.e
        e.tput_timer.start()
        rows = self._rows_now("act")
        self.rows_of[buf] = rows
        inputs = self.buffers["inputs"][buf]
        slice_group = e.grid.get_slice_parallel_group() if e.is_pipe_partitioned else None

        if e.is_pipe_partitioned and not e.is_first_stage():
            # (partition meta, local slice, rest...) -> full activation, which becomes the autograd leaf
            full = PartitionedTensor.from_meta(meta=inputs[0], local_part=inputs[1], group=slice_group).full()
            full.requires_grad_(full.is_floating_point())
            inputs = (full,) + tuple(inputs[2:]) if len(inputs) > 2 else full
            self.buffers["inputs"][buf] = inputs

        outputs = e._module_forward(inputs)
        self.buffers["outputs"][buf] = outputs

        if e.is_last_stage():
            self._finish_last_stage(buf, outputs, rows)
            return
        if e.is_pipe_partitioned:
            first = outputs[0] if isinstance(outputs, tuple) else outputs
            part = PartitionedTensor(tensor=first, group=slice_group)
            rest = tuple(outputs[1:]) if isinstance(outputs, tuple) else ()
            self.buffers["output_tensors"][buf] = first
            self.buffers["wire"][buf] = (part.to_meta(), part.data()) + rest
        else:
            self.buffers["wire"][buf] = outputs if isinstance(outputs, tuple) else (outputs,)

    def _finish_last_stage(self, buf, outputs, rows):
        e = self.e
        if getattr(e, "_inference_mode", False):
            return
        loss = e.loss_model(outputs, self.buffers["labels"][buf]) if e.loss_model is not None else outputs
        weight = rows / e.micro_batch_size
        if weight != 1.0:      # a slice of a micro-batch counts for its share of the rows
            loss = loss * weight if torch.is_tensor(loss) else [l * weight for l in loss]
        e.loss = loss
        if e.eval_return_logits:
            e.outputs = outputs
        if torch.is_tensor(loss):
            e.total_loss = loss.detach().clone() if e.total_loss is None else e.total_loss + loss.detach()
        else:
            if e.total_loss is None:
                e.total_loss = [torch.zeros_like(l) for l in loss]
            for i, l in enumerate(loss):
                e.total_loss[i] += l.detach()

    def backward(self, buf):
        e = self.e
        assert e.optimizer is not None, "must provide optimizer during init in order to use backward"
        if e.is_last_stage():
            e._loss_backward(e.loss)          # through the engine: loss scaling and 1 / gradient_accumulation_steps
        else:
            for name in ("backward_microstep", "backward", "backward_inner_microstep", "backward_inner"):
                self._timer(name, True)
            targets = self._grad_targets(buf)
            grads = self.buffers["grads"][buf]
            assert len(targets) == len(grads)
            torch.autograd.backward(tensors=targets, grad_tensors=grads)
            for name in ("backward_inner", "backward_inner_microstep", "backward", "backward_microstep"):
                self._timer(name, False)
        for key in ("outputs", "output_tensors", "wire", "grads", "labels"):
            self.buffers[key][buf] = None
        if e.is_first_stage():
            self.buffers["inputs"][buf] = None

    # ------------------------------------------------------------------ end of batch
    def tail(self, name):
        e = self.e
        if name == "reduce_tied":
            self._timer("reduce_tied_grads", True)
            e.module.allreduce_tied_weight_gradients()
            self._timer("reduce_tied_grads", False)
        elif name == "reduce":
            self._timer("reduce_grads", True)
            e._reduce_stage_gradients()
            self._timer("reduce_grads", False)
        elif name == "step":
            e._pipeline_optimizer_step()
