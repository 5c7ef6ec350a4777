"""Compressed (1-bit, error-feedback) all-reduce: torch restatement vs an independent simulation on CPU/gloo, and the
fused sm_100a kernel vs the restatement on 2 GPUs (oracle in the spirit of reference tests/unit/test_onebit.py:862-884)."""
import pytest
import torch
import torch.distributed as dist

from common import distributed_test, run_distributed


def torch_sim(a, world, rank, group=None):
    """All-gather everything and compute the two-phase result densely."""
    n = a.numel()
    c = n // world
    scale = a.norm() / (n ** 0.5)
    sgn = (a >= 0).float() * 2 - 1
    ws = [torch.zeros_like(a) for _ in range(world)]
    dist.all_gather(ws, scale * sgn, group=group)
    avg = torch.stack(ws).mean(0)
    out = []
    for k in range(world):
        chunk = avg[k * c:(k + 1) * c]
        sc = chunk.norm() / (c ** 0.5)
        out.append(sc * ((chunk >= 0).float() * 2 - 1))
    return torch.cat(out)


def _cpu_worker():
    from deeperspeed_b200.runtime.comm.nccl import NcclBackend
    world, rank = dist.get_world_size(), dist.get_rank()
    backend = NcclBackend()
    n = 64 * world * 4
    torch.manual_seed(10 + rank)
    a = (torch.rand(n) - 0.5) + 0.01 * rank
    werr, serr = torch.zeros(n), torch.zeros(n // world)
    want = torch_sim(a.clone(), world, rank)
    got = backend.compressed_allreduce(a.clone(), werr, serr, rank)
    assert torch.allclose(got, want, atol=1e-6), (got - want).abs().max()
    # error feedback buffers are non-trivial and bounded
    assert werr.abs().sum() > 0 and serr.abs().sum() > 0
    # 2-D input keeps its shape; padded input (numel < error size) is cut back
    b = torch.rand(8, n // 8) - 0.5
    assert backend.compressed_allreduce(b, torch.zeros(n), torch.zeros(n // world), rank).shape == b.shape
    s = torch.rand(n - 5) - 0.5
    assert backend.compressed_allreduce(s, torch.zeros(n), torch.zeros(n // world), rank).numel() == n - 5


def test_compressed_allreduce_cpu():
    run_distributed(_cpu_worker, world_size=2, use_cuda=False)


def _gpu_worker():
    from deeperspeed_b200.runtime.comm.nccl import NcclBackend
    world, rank = dist.get_world_size(), dist.get_rank()
    fused, plain = NcclBackend(use_fused=True), NcclBackend(use_fused=False)
    n = 32 * world * 4096
    torch.manual_seed(10 + rank)
    we_f, se_f = torch.zeros(n, device="cuda"), torch.zeros(n // world, device="cuda")
    we_p, se_p = we_f.clone(), se_f.clone()
    for it in range(3):   # several rounds so the error-feedback state is exercised
        a = ((torch.rand(n, device="cuda") - 0.5) + 0.01 * rank)
        got = fused.compressed_allreduce(a.clone(), we_f, se_f, rank)
        want = plain.compressed_allreduce(a.clone(), we_p, se_p, rank)
        torch.cuda.synchronize()
        bad = (got - want).abs() > 1e-5
        # a sign may legitimately differ where the server value is ~0 (summation order)
        assert bad.float().mean().item() < 1e-3, (it, bad.float().mean().item())
        assert torch.allclose(we_f, we_p, atol=1e-5)
        assert (se_f - se_p).abs().mean().item() < 1e-4


@pytest.mark.gpu
def test_compressed_allreduce_fused_2gpu():
    run_distributed(_gpu_worker, world_size=2, use_cuda=True)


def _onebit_opt_worker(kind):
    import deeperspeed_b200 as deepspeed
    from simple_model import SimpleModel, random_dataset
    cfg = {"train_batch_size": 4, "steps_per_print": 100,
           "optimizer": {"type": kind, "params": {"lr": 0.01, "freeze_step": 3, "weight_decay": 0.0}},
           "fp16": {"enabled": True, "type": "bfloat16"}}
    torch.manual_seed(1)
    model = SimpleModel(64)
    engine, opt, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=cfg)
    it = iter(torch.utils.data.DataLoader(random_dataset(256, 64, dtype=torch.bfloat16, seed=dist.get_rank()), batch_size=2))
    losses = []
    for step in range(8):
        x, y = next(it)
        loss = engine(x, y); engine.backward(loss); engine.step()
        losses.append(loss.item())
    inner = opt.optimizer
    assert (inner.adam_freeze_key if kind == "OneBitAdam" else inner.lamb_freeze_key), "compression phase not entered"
    assert engine.enable_backward_allreduce is False
    assert all(l == l for l in losses)
    # parameters stay identical across ranks: the compressed all-reduce returns the same momentum everywhere
    flat = torch.cat([p.detach().float().view(-1) for p in model.parameters()])
    ref = flat.clone()
    dist.broadcast(ref, src=0)
    assert torch.allclose(flat, ref, atol=1e-3), (flat - ref).abs().max()
    # checkpoint: loading a compression-phase state re-enters compression and resets the error buffers
    sd = inner.state_dict()
    inner.load_state_dict(sd)
    assert (inner.adam_freeze_key if kind == "OneBitAdam" else inner.lamb_freeze_key)


@pytest.mark.parametrize("kind", ["OneBitAdam", "OneBitLamb"])
def test_onebit_optimizers_cpu(kind):
    run_distributed(_onebit_opt_worker, world_size=2, use_cuda=False, args=(kind,))


def _fused_adam_step_worker():
    """1-bit Adam compression-stage step: momentum + compressed exchange + parameter update in ONE kernel against the
    same step done as separate passes around the torch.distributed restatement. The flat tensor length is NOT a
    multiple of the sign-word padding, so the padded tail of the error buffers is exercised."""
    from deeperspeed_b200.runtime.fp16.onebit.adam import OnebitAdam
    world, rank = dist.get_world_size(), dist.get_rank()
    n = 32 * world * 1024 + 8 * 5
    torch.manual_seed(3)
    init = torch.randn(n, device="cuda") * 0.1
    results = []
    for fused in (True, False):
        p = torch.nn.Parameter(init.clone())
        opt = OnebitAdam([p], None, lr=1e-2, freeze_step=2, weight_decay=0.01)
        opt.comm_backend_handle.use_fused = fused
        g = torch.Generator(device="cuda").manual_seed(100 + rank)
        for it in range(6):
            grad = torch.randn(n, device="cuda", generator=g) * 0.05 + 0.01 * rank
            if it < 2:      # warm-up stage: the engine would average the gradients
                dist.all_reduce(grad)
                grad /= world
            p.grad = grad
            opt.step()
        assert opt.adam_freeze_key
        torch.cuda.synchronize()
        st = opt.state[p]
        results.append((p.detach().clone(), st["exp_avg"].clone(), st["worker_error"].clone()))
    (p_f, m_f, e_f), (p_u, m_u, e_u) = results
    # a sign may legitimately differ where a compensated value is ~0 (summation order of the norms)
    assert ((m_f - m_u).abs() > 1e-5).float().mean().item() < 2e-3
    assert ((p_f - p_u).abs() > 1e-4).float().mean().item() < 2e-3, (p_f - p_u).abs().max().item()
    assert (e_f - e_u).abs().mean().item() < 1e-4
    # every rank ends with the same parameters
    other = p_f.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(other, p_f)


@pytest.mark.gpu
def test_onebit_adam_fused_step_2gpu():
    run_distributed(_fused_adam_step_worker, world_size=2, use_cuda=True)


class _ThreadComm:
    """In-process stand-in for an MPI communicator: `size` ranks run as threads and meet at a barrier. Accepts numpy
    arrays (host-staged transport) and torch tensors (what a CUDA-aware MPI is handed)."""

    class _Shared:
        def __init__(self, size):
            import threading
            self.size = size
            self.barrier = threading.Barrier(size)
            self.slots = [None] * size

    def __init__(self, shared, rank):
        self.shared, self.rank = shared, rank
        self.calls = []

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.shared.size

    @staticmethod
    def _view(buf):
        return buf if isinstance(buf, torch.Tensor) else torch.from_numpy(buf)

    def _meet(self, send):
        sh = self.shared
        sh.slots[self.rank] = self._view(send).clone()
        sh.barrier.wait()
        everyone = list(sh.slots)
        sh.barrier.wait()
        return everyone

    def Alltoall(self, send, recv):
        self.calls.append(("Alltoall", type(send).__name__))
        everyone = self._meet(send)
        n = everyone[0].numel() // self.shared.size
        out = self._view(recv)
        for r, src in enumerate(everyone):
            out[r * n:(r + 1) * n] = src[self.rank * n:(self.rank + 1) * n]

    def Allgather(self, send, recv):
        self.calls.append(("Allgather", type(send).__name__))
        everyone = self._meet(send)
        self._view(recv).copy_(torch.cat([e.view(-1) for e in everyone]))


@pytest.mark.parametrize("cuda_aware", [False, True])
def test_mpi_backend_both_transports(cuda_aware):
    """MpiBackend (reference runtime/comm/mpi.py): device-buffer and host-staged transports give the dense two-phase
    result; the CUDA-aware one hands the tensors themselves to MPI, the other numpy views of pinned staging buffers."""
    import threading
    from deeperspeed_b200.runtime.comm.mpi import MpiBackend
    world = 2
    n = 64 * world * 4
    inputs = []
    for rank in range(world):
        torch.manual_seed(10 + rank)
        inputs.append((torch.rand(n) - 0.5) + 0.01 * rank)
    # dense oracle
    quant = [a.norm() / n ** 0.5 * ((a >= 0).float() * 2 - 1) for a in inputs]
    avg = torch.stack(quant).mean(0)
    c = n // world
    want = torch.cat([avg[k * c:(k + 1) * c].norm() / c ** 0.5 * ((avg[k * c:(k + 1) * c] >= 0).float() * 2 - 1)
                      for k in range(world)])
    shared = _ThreadComm._Shared(world)
    results, comms, errors = [None] * world, [None] * world, []

    def run(rank):
        try:
            comms[rank] = _ThreadComm(shared, rank)
            backend = MpiBackend(cuda_aware, comm=comms[rank])
            werr, serr = torch.zeros(n), torch.zeros(n // world)
            results[rank] = backend.compressed_allreduce(inputs[rank].clone(), werr, serr, rank)
            assert werr.abs().sum() > 0 and serr.abs().sum() > 0
            short = torch.rand(n - 5) - 0.5          # padded input is cut back to its own length
            assert backend.compressed_allreduce(short, torch.zeros(n), torch.zeros(n // world), rank).numel() == n - 5
        except Exception as e:  # pragma: no cover
            errors.append(e)
            shared.barrier.abort()

    threads = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in threads]
    [t.join(60) for t in threads]
    assert not errors, errors
    for got in results:
        assert torch.allclose(got, want, atol=1e-6)
    kinds = {kind for _, kind in comms[0].calls}
    assert kinds == ({"Tensor"} if cuda_aware else {"ndarray"})
