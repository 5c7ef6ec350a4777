"""Cross-GPU flag / mailbox protocol of the peer-memory kernels under skew (2 GPUs).

compute-sanitizer is closed on the development pool (profiles/sanitizer_status.txt), so the hazards it would look for
are provoked directly: the ranks enter the kernels at very different times (one of them is held back by a spinning
kernel, alternating), hundreds of epochs run back to back without any host synchronisation in between, and every
result is compared with what the exchange must produce. A lost flag, a mailbox overwritten before it was read, or a
barrier that lets a rank through early shows up as a wrong value or a hang (pytest-timeout)."""
import pytest
import torch
import torch.distributed as dist

from common import run_distributed

pytestmark = pytest.mark.gpu


def _fused_zero_epochs():
    from deeperspeed_b200.ops.adam import FusedAdam
    from deeperspeed_b200.runtime.zero.stage1 import FP16_DeepSpeedZeroOptimizer_Stage1
    rank, world = dist.get_rank(), dist.get_world_size()
    n = 8 * 4096 * world + 8 * 40            # not a multiple of the CTA tile: ragged tails in every phase
    p = torch.nn.Parameter(torch.zeros(n, device="cuda", dtype=torch.float16))
    lr = 2.0 ** -7
    opt = FP16_DeepSpeedZeroOptimizer_Stage1(FusedAdam([p], lr=lr, betas=(0.0, 0.0), eps=0.0, weight_decay=0.0,
                                                       bias_correction=False),
                                             static_loss_scale=1.0, dp_process_group=dist.group.WORLD,
                                             max_elements_per_comm=int(2e9), fused_nvlink=True, clip_grad=0.0)
    assert opt.fused is not None
    grad = opt.arenas[0].grad
    steps = 300
    for it in range(steps):
        # with beta1 = beta2 = eps = 0 Adam moves every element by exactly lr * sign(mean gradient)
        sign = 1.0 if it % 3 else -1.0
        grad.fill_(sign * (rank + 1) * 2.0 ** -10)
        if it % 2 == rank:
            torch.cuda._sleep(3_000_000)      # ~1.5 ms: this rank arrives late, the other waits inside the kernel
        opt.step()
        # the kernel must have cleared the whole local gradient arena (peers' slices only after they were read)
        if it % 50 == 0:
            assert float(grad.abs().max()) == 0.0
    torch.cuda.synchronize()
    expect = sum(-lr if it % 3 else lr for it in range(steps))
    got = opt.arenas[0].data[:n].float()
    assert torch.equal(got, torch.full_like(got, expect)), (got.min().item(), got.max().item(), expect)
    # both ranks hold the same parameters: every slice reached every rank
    other = got.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(other, got)
    assert opt.fused.launches["whole_step"] == steps


def test_fused_zero_step_protocol_under_skew():
    run_distributed(_fused_zero_epochs, world_size=2, use_cuda=True)


def _onebit_epochs():
    from deeperspeed_b200.runtime.comm.nccl import NcclBackend
    rank, world = dist.get_rank(), dist.get_world_size()
    backend = NcclBackend()
    n = 32 * world * 64
    werr, serr = torch.zeros(n, device="cuda"), torch.zeros(n // world, device="cuda")
    for it in range(200):
        # every rank contributes +-c_r with the SAME sign pattern: one magnitude per rank, so the quantisation is
        # exact, error feedback stays zero and the result must be the mean magnitude with that sign pattern
        pattern = torch.where((torch.arange(n, device="cuda") + it) % 3 == 0, -1.0, 1.0)
        x = pattern * (rank + 1.0)
        if it % 2 == rank:
            torch.cuda._sleep(2_000_000)
        out = backend.compressed_allreduce(x.clone(), werr, serr, rank)
        want = pattern * (sum(range(1, world + 1)) / world)
        assert torch.allclose(out, want, rtol=1e-6, atol=1e-6), (it, (out - want).abs().max().item())
    torch.cuda.synchronize()
    assert float(werr.abs().max()) < 1e-5 and float(serr.abs().max()) < 1e-5


def test_onebit_allreduce_protocol_under_skew():
    run_distributed(_onebit_epochs, world_size=2, use_cuda=True)
