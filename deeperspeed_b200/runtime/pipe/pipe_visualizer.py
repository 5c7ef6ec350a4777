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


def pipeline_visualizer(num_stages, num_micro_batches, schedule_cls=schedule.TrainSchedule, print_fn=print):
    rows = schedule_grid(num_stages, num_micro_batches, schedule_cls)
    width = max(len(c) for r in rows for c in r) + 1
    for stage, cells in enumerate(rows):
        print_fn(f"stage {stage}: " + "".join(c.ljust(width) for c in cells))
    busy = sum(c != "." for r in rows for c in r)
    total = sum(len(r) for r in rows)
    print_fn(f"bubble fraction: {1 - busy / total:.3f}")
    return rows
