"""Pipeline point-to-point transport: one-directional links between neighbouring stages.

Every neighbour pair gets TWO 2-rank process groups, one per direction (activations travel down, gradients travel
up). On NCCL each group is its own communicator with its own stream, so
  * an activation send and a gradient receive posted in the same tick run concurrently (full-duplex NVLink), and
  * nothing is enqueued on the default group, whose collectives (loss broadcast, ZeRO reduce-scatter) would
    otherwise serialise with the unbatched point-to-point traffic.
Sends are non-blocking (`Link.post`); the executor waits for a send only when the next exchange starts, i.e. one
compute step later. Receives are awaited right before the compute that consumes them. On gloo (CPU tests) the same
calls are asynchronous in the backend's own threads.

The reference emulates p2p with blocking 2-rank broadcasts (reference deepspeed/runtime/pipe/p2p.py:31-61); its
module-level functions (`init_process_groups`, `send`, `recv`, `barrier`) are kept on top of the links.
"""
import torch
import torch.distributed as dist

from ...utils import comm_stub

_grid = None
_links = None


class Link:
    """One direction of one neighbour pair."""

    def __init__(self, group, peer, sends):
        self.group = group
        self.peer = peer          # global rank of the other end
        self.sends = sends        # True: this rank is the sender on this link

    def post(self, tensor):
        """Start the transfer of `tensor` (send) or into `tensor` (receive); returns the work handle."""
        if comm_stub.enabled():
            return comm_stub.DONE
        # issued as a one-element batch: NCCL then runs it as a grouped point-to-point operation of this link's own
        # communicator instead of an "independent collective" of an eagerly initialised group
        op = dist.P2POp(dist.isend if self.sends else dist.irecv, tensor, self.peer, self.group)
        return dist.batch_isend_irecv([op])[-1]


class StageLinks:
    """The up to four links of this rank: activations in / out, gradients in / out (None at the pipeline ends)."""

    def __init__(self, grid):
        self.act_in = self.act_out = self.grad_in = self.grad_out = None
        self.pair_groups = {}
        me = grid.get_global_rank()
        stages = grid.pipe_parallel_size
        # collective construction: every rank creates the groups of every adjacent pair, in the same order
        for ranks in grid.pipe_groups if stages > 1 else []:
            for s in range(stages - 1):
                up, down = ranks[s], ranks[s + 1]
                g_act = dist.new_group(ranks=[up, down])
                g_grad = dist.new_group(ranks=[up, down])
                self.pair_groups[(up, down)] = g_act
                if me == up:
                    self.act_out = Link(g_act, down, sends=True)
                    self.grad_in = Link(g_grad, down, sends=False)
                elif me == down:
                    self.act_in = Link(g_act, up, sends=False)
                    self.grad_out = Link(g_grad, up, sends=True)

    def toward(self, dest_stage, my_stage, sending):
        """The link used for a transfer between my stage and an adjacent one."""
        down = dest_stage > my_stage
        if sending:
            return self.act_out if down else self.grad_out
        return self.grad_in if down else self.act_in


def init_process_groups(grid):
    """Build the links of this rank (collective: every rank must call it)."""
    global _grid, _links
    _grid = grid
    assert grid.pipe_parallel_size > 1, "There is no pipeline parallelism"
    _links = StageLinks(grid)
    return _links


def links():
    return _links


def _is_valid_send_recv(src_stage, dest_stage):
    assert abs(src_stage - dest_stage) == 1, \
        "Functionality currently limited to send and receive between adjacent ranks only"


def _wire(tensor, fp32_comm):
    if fp32_comm and tensor.is_floating_point() and tensor.dtype != torch.float32:
        return tensor.float()
    return tensor


def send(tensor, dest_stage, async_op=False, fp32_comm=False):
    """Send to an adjacent stage (activations downstream, gradients upstream)."""
    me = _grid.get_stage_id()
    _is_valid_send_recv(me, dest_stage)
    work = _links.toward(dest_stage, me, sending=True).post(_wire(tensor, fp32_comm).contiguous())
    if async_op:
        return work
    work.wait()
    return None


def recv(tensor, src_stage, async_op=False, fp32_comm=False):
    """Receive from an adjacent stage into `tensor`."""
    me = _grid.get_stage_id()
    _is_valid_send_recv(src_stage, me)
    link = _links.toward(src_stage, me, sending=False)
    if fp32_comm and tensor.is_floating_point() and tensor.dtype != torch.float32:
        wide = torch.empty_like(tensor, dtype=torch.float32)
        link.post(wide).wait()
        tensor.copy_(wide)
        return None
    work = link.post(tensor)
    if async_op:
        return work
    work.wait()
    return None


def barrier(stage_id):
    """Barrier between `stage_id` and its downstream neighbour."""
    dist.barrier(group=_get_send_recv_group(stage_id, stage_id + 1))


def _get_send_recv_group(src_stage, dest_stage):
    """The (activation-direction) group of a neighbour pair."""
    lo, hi = sorted((src_stage, dest_stage))
    return _links.pair_groups[(_grid.stage_to_global(stage_id=lo), _grid.stage_to_global(stage_id=hi))]
