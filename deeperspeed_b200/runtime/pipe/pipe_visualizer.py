"""Text visualiser of pipeline schedules (DeeperSpeed addition, reference runtime/pipe/pipe_visualizer.py:46-68):
one row per stage, one column per schedule step, F<mb>/B<mb> for forward/backward of micro-batch mb."""
from . import schedule


def schedule_grid(num_stages, num_micro_batches, schedule_cls=schedule.TrainSchedule):
    rows = []
    for stage in range(num_stages):
        sched = schedule_cls(micro_batches=num_micro_batches, stages=num_stages, stage_id=stage)
        cells = []
        fwd = bwd = 0
        for cmds in sched:
            cell = "."
            for c in cmds:
                if isinstance(c, schedule.ForwardPass):
                    cell = f"F{fwd}"
                    fwd += 1
                elif isinstance(c, schedule.BackwardPass):
                    cell = f"B{bwd}"
                    bwd += 1
            cells.append(cell)
        rows.append(cells)
    return rows


def pipeline_visualizer(num_stages, num_microbatches=None, include_all=False, *, num_micro_batches=None,
                        schedule_cls=None, print_fn=None):
    """The reference's call (pipe_visualizer.py:44): `pipeline_visualizer(num_stages, num_microbatches, include_all)`
    returns the schedule as a markdown table plus idle / busy counts. With `schedule_cls` or `print_fn` it prints this
    package's compact grid (any schedule class) and returns its rows."""
    if num_microbatches is None:
        num_microbatches = num_micro_batches
    if schedule_cls is None and print_fn is None:
        return pipeline_table(num_stages, num_microbatches, include_all)
    return print_schedule_grid(num_stages, num_microbatches, schedule_cls or schedule.TrainSchedule, print_fn or print)


def print_schedule_grid(num_stages, num_micro_batches, schedule_cls=schedule.TrainSchedule, print_fn=print):
    rows = schedule_grid(num_stages, num_micro_batches, schedule_cls)
    width = max(len(c) for r in rows for c in r) + 1
    for stage, cells in enumerate(rows):
        print_fn(f"stage {stage}: " + "".join(c.ljust(width) for c in cells))
    busy = sum(c != "." for r in rows for c in r)
    total = sum(len(r) for r in rows)
    print_fn(f"bubble fraction: {1 - busy / total:.3f}")
    return rows


# short labels of the reference's table (runtime/pipe/pipe_visualizer.py:33-44)
reprs = {schedule.ForwardPass: "fwd", schedule.BackwardPass: "bwd", schedule.RecvActivation: "recv_act",
         schedule.SendActivation: "send_act", schedule.RecvGrad: "recv_grad", schedule.SendGrad: "send_grad",
         schedule.LoadMicroBatch: "load_batch", schedule.ReduceGrads: "reduce_grads",
         schedule.ReduceTiedGrads: "reduce_tied_grads", schedule.OptimizerStep: "step"}


def expand(steps, include_all=False):
    """Instruction lists of one stage -> one label per step ("fwd_3 / send_act_3", None for an idle step). Without
    `include_all` only forward / backward passes are shown. Buffer ids are printed 1-based, as in the reference."""
    shown = tuple(reprs) if include_all else (schedule.ForwardPass, schedule.BackwardPass)
    out = []
    for cmds in steps:
        labels = []
        for cmd in cmds or ():
            if isinstance(cmd, shown):
                label = reprs[type(cmd)]
                if hasattr(cmd, "buffer_id"):
                    label += f"_{cmd.buffer_id + 1}"
                labels.append(label)
        out.append(" / ".join(labels) if labels else None)
    return out


def pipeline_table(num_stages, num_microbatches, include_all=False):
    """Markdown table of the 1F1B schedule (one row per stage) followed by the idle / busy step counts: the string the
    reference's `pipeline_visualizer` returns, built without pytablewriter."""
    rows = [expand(list(schedule.TrainSchedule(micro_batches=num_microbatches, stages=num_stages, stage_id=s).steps()),
                   include_all) for s in range(num_stages)]
    steps = max(len(r) for r in rows)
    header = ["GPU ID"] + [str(i) for i in range(steps)]
    lines = ["# Pipe Schedule", "", "|" + "|".join(header) + "|", "|" + "|".join("---" for _ in header) + "|"]
    for s, row in enumerate(rows):
        cells = [f"GPU {s}"] + [c or "" for c in row + [None] * (steps - len(row))]
        lines.append("|" + "|".join(cells) + "|")
    idle = sum(c is None for r in rows for c in r)
    busy = sum(c is not None for r in rows for c in r)
    lines += ["", f"Num Devices: {num_stages}", f"Num Microbatches: {num_microbatches} ", f"Idle Time: {idle}",
              f"Non Idle Time: {busy}"]
    return "\n".join(lines)

