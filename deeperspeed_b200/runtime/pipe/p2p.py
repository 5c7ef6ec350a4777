"""Pipeline point-to-point transport.

NCCL send/recv between adjacent stage ranks, stream-ordered (asynchronous to the host), so the next kernels are
queued while a transfer is in flight; on CPU/gloo the same calls are blocking. The reference emulates p2p with a
2-rank broadcast and forces it synchronous (reference deepspeed/runtime/pipe/p2p.py:31-61); the pair groups are
still built for API parity (`_get_send_recv_group`)."""
import torch
import torch.distributed as dist

_groups = None
_grid = None


def init_process_groups(grid):
    """Create the 2-rank neighbour groups (collective: every rank walks the same list)."""
    global _groups, _grid
    _grid = grid
    assert _grid.pipe_parallel_size > 1, "There is no pipeline parallelism"
    _groups = [dist.new_group(ranks=group) for group in _grid.p2p_groups]


def _is_valid_send_recv(src_stage, dest_stage):
    """Only ring neighbours talk to each other (the last stage wraps around to the first)."""
    n = _grid.pipe_parallel_size
    assert (src_stage - dest_stage) % n in (1, n - 1), \
        "Functionality currently limited to send and receive between adjacent ranks only"


def _wire(tensor, fp32_comm):
    return tensor.float() if (fp32_comm and tensor.is_floating_point() and tensor.dtype != torch.float32) else tensor


def send(tensor, dest_stage, async_op=False, fp32_comm=False):
    src_stage = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, dest_stage)
    dst = _grid.stage_to_global(stage_id=dest_stage)
    t = _wire(tensor, fp32_comm).contiguous()
    if async_op:
        return dist.isend(t, dst=dst)
    dist.send(t, dst=dst)
    return None


def recv(tensor, src_stage, async_op=False, fp32_comm=False):
    dest_stage = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, dest_stage)
    src = _grid.stage_to_global(stage_id=src_stage)
    if fp32_comm and tensor.is_floating_point() and tensor.dtype != torch.float32:
        tmp = torch.empty_like(tensor, dtype=torch.float32)
        dist.recv(tmp, src=src)
        tensor.copy_(tmp)
        return None
    if async_op:
        return dist.irecv(tensor, src=src)
    dist.recv(tensor, src=src)
    return None


def barrier(stage_id):
    """Barrier with the next-stage neighbour."""
    group_id = _grid.stage_to_global(stage_id=stage_id)
    dist.barrier(group=_groups[group_id])


def _get_send_recv_group(src_stage, dest_stage):
    """The 2-rank group of a neighbour pair: groups are listed by the pair's lower stage, the wrap-around pair
    (last, first) by the last stage."""
    lo, hi = sorted((src_stage, dest_stage))
    wraps = lo == 0 and hi == _grid.pipe_parallel_size - 1
    owner = hi if wraps else lo
    return _groups[_grid.stage_to_global(stage_id=owner)]
