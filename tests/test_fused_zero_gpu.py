"""Fused NVLink ZeRO step (peer-memory reduce-scatter + Adam + all-gather kernels) vs the NCCL path, 2 GPUs."""
import os

import pytest
import torch
import torch.distributed as dist

import deeperspeed_b200 as deepspeed
from common import run_distributed

pytestmark = pytest.mark.gpu


class Net(torch.nn.Module):
    def __init__(self, h=256):
        super().__init__()
        self.a = torch.nn.Linear(h, 4 * h)
        self.b = torch.nn.Linear(4 * h, h)

    def forward(self, x, y):
        return torch.nn.functional.cross_entropy(self.b(torch.nn.functional.gelu(self.a(x))).float(), y)


def _run(fused, steps=4):
    torch.manual_seed(5)
    model = Net()
    cfg = {"train_batch_size": 16, "train_micro_batch_size_per_gpu": 4, "steps_per_print": 100,
           "optimizer": {"type": "Adam", "params": {"lr": 1e-3, "weight_decay": 0.01}},
           "fp16": {"enabled": True, "initial_scale_power": 8}, "gradient_clipping": 1.0,
           "zero_optimization": {"stage": 1, "fused_nvlink": fused}}
    engine, opt, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=cfg)
    assert (opt.fused is not None) == fused
    g = torch.Generator().manual_seed(100 + dist.get_rank())
    for _ in range(steps * engine.gradient_accumulation_steps()):
        x = torch.randn(4, 256, generator=g).half().cuda()
        y = torch.randint(0, 256, (4,), generator=g).cuda()
        loss = engine(x, y)
        engine.backward(loss)
        engine.step()
    return [p.detach().float().clone() for p in model.parameters()], opt


def _fused_vs_nccl_worker():
    global _ref_opt
    ref, _ref_opt = _run(False)
    got, opt = _run(True)
    # Adam moves an element by ~lr per step whatever the gradient magnitude, so fp16-vs-fp32 reduction rounding of
    # near-zero gradients may flip a sign: bound the drift by 2 * lr * steps and require the bulk to agree tightly
    for a, b in zip(got, ref):
        d = (a - b).abs()
        assert d.max().item() <= 2 * 1e-3 * 4 + 1e-3, d.max().item()
        assert (d > 2e-3).float().mean().item() < 0.02
    # every rank sees identical parameters after the peer-store all-gather
    flat = torch.cat([p.view(-1) for p in got])
    other = flat.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(flat, other)
    # exact check with injected gradients: both paths must produce the same update from the same (rank-dependent) grads
    torch.manual_seed(77)
    g = torch.randn(opt.arenas[0].numel, device="cuda") * 0.1 * (1 + dist.get_rank())
    masters = []
    for o in (opt, _ref_opt):
        o.zero_grad()
        o.arenas[0].grad.copy_(g.half())
        before = o.fp32_masters[0].detach().clone()
        o.step()
        masters.append((o.fp32_masters[0].detach() - before))
    torch.cuda.synchronize()
    assert (masters[0] - masters[1]).abs().max().item() < 2e-5, (masters[0] - masters[1]).abs().max().item()
    # overflow is skipped consistently through the device-side flag
    opt.arenas[0].grad[7] = float("inf")
    opt.step()
    assert opt.overflow


def test_fused_matches_nccl_path():
    run_distributed(_fused_vs_nccl_worker, world_size=2, use_cuda=True)


def _lamb_reference(w, g, lr, betas, eps, wd, max_coeff, min_coeff, step, m, v):
    """Plain fp32 LAMB on one whole tensor (reference csrc/lamb/fused_lamb_cuda_kernel.cu:186-310 math)."""
    b1, b2 = betas
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    u = m / (v.sqrt() + eps) + wd * w
    wn, un = w.norm(), u.norm()
    coeff = 1.0
    if wn > 0 and un > 0:
        coeff = float(min(max(wn / un, min_coeff), max_coeff))
    step_size = lr * (1 - b2 ** step) ** 0.5 / (1 - b1 ** step)
    return w - step_size * coeff * u, coeff


def _fused_lamb_worker():
    torch.manual_seed(5)
    model = Net()
    hp = {"lr": 2e-3, "weight_decay": 0.01, "max_coeff": 10.0, "min_coeff": 0.01}
    cfg = {"train_batch_size": 8, "train_micro_batch_size_per_gpu": 4, "steps_per_print": 100,
           "optimizer": {"type": "Lamb", "params": dict(hp)},
           "fp16": {"enabled": True, "initial_scale_power": 6}, "zero_allow_untested_optimizer": True,
           # small buckets: several comm intervals, so tensors straddle both ranks AND bucket boundaries
           "zero_optimization": {"stage": 1, "fused_nvlink": True, "reduce_bucket_size": 100000}}
    engine, opt, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=cfg)
    assert opt.fused is not None and opt.fused.is_lamb
    arena = opt.arenas[0]
    assert opt.num_buckets[0] > 1
    rank, world = dist.get_rank(), dist.get_world_size()
    ws = [s.param.detach().float().clone() for s in arena.segments]
    ms = [torch.zeros_like(w) for w in ws]
    vs = [torch.zeros_like(w) for w in ws]
    for step in (1, 2):
        gens = [torch.Generator(device="cuda").manual_seed(1000 * step + r) for r in range(world)]
        per_rank = [(torch.randn(arena.numel, device="cuda", generator=gens[r]) * 0.05 * (1 + r)).half()
                    for r in range(world)]
        opt.zero_grad()
        scale = float(opt.loss_scale)
        for s in arena.segments:   # write through the parameter views so arena padding stays zero
            s.param.grad.copy_(per_rank[rank].narrow(0, s.offset, s.numel).view(s.shape))
        opt.step()
        torch.cuda.synchronize()
        assert not opt.overflow
        coeffs = opt.optimizer.get_lamb_coeffs()
        for ti, s in enumerate(arena.segments):
            g = sum(pr.narrow(0, s.offset, s.numel).float() for pr in per_rank).view(s.shape) / world / scale
            ws[ti], c = _lamb_reference(ws[ti], g, hp["lr"], (0.9, 0.999), 1e-8, hp["weight_decay"], hp["max_coeff"],
                                        hp["min_coeff"], step, ms[ti], vs[ti])
            got = s.param.detach().float()
            assert (got - ws[ti]).abs().max().item() < 2e-3 * max(1.0, ws[ti].abs().max().item()), (ti, step)
            # the ratio is per whole tensor although every tensor here is split over ranks and buckets
            _, mine = opt.fused.owned_segments(arena, opt.sub_size[0], opt.num_buckets[0], world, rank)
            if ti in mine:
                assert abs(coeffs[ti] - c) < 1e-3 * max(1.0, c), (ti, coeffs[ti], c)
    flat = arena.data.clone()
    other = flat.clone()
    dist.broadcast(other, src=0)
    assert torch.equal(flat, other)
    # overflow: skipped on every rank, moments and step count untouched
    before = opt.fp32_masters[0].detach().clone()
    opt.arenas[0].grad[3] = float("inf")
    opt.step()
    assert opt.overflow
    assert torch.equal(before, opt.fp32_masters[0].detach())
    assert opt.optimizer.state[opt.fp32_masters[0]]["step"] == 2


def test_fused_lamb_per_tensor_trust_ratio_2gpu():
    run_distributed(_fused_lamb_worker, world_size=2, use_cuda=True)


def _gemm_rs_worker():
    from deeperspeed_b200.parallel.ag_gemm import GemmReduceScatter
    rank, world = dist.get_rank(), dist.get_world_size()
    torch.cuda.set_device(rank)
    rs = GemmReduceScatter(1024 * 2048, dtype=torch.float16)
    for it, (tokens, out_f, in_f) in enumerate([(2048, 1024, 2048), (2048, 1024, 2048), (1024, 512, 520), (4096, 256, 1024)]):
        g = torch.Generator(device="cuda").manual_seed(10 * it + rank)
        go = (torch.randn(tokens, out_f, device="cuda", generator=g) * 0.1).half()
        x = (torch.randn(tokens, in_f, device="cuda", generator=g) * 0.1).half()
        got = rs(go, x, trans_a=True, trans_b=False)          # weight gradient rows owned by this rank, averaged
        ref = go.float().t() @ x.float()
        dist.all_reduce(ref)
        ref = ref[rank * (out_f // world):(rank + 1) * (out_f // world)] / world
        torch.cuda.synchronize()
        assert got.shape == ref.shape
        err = (got.float() - ref).abs().max().item()
        assert err < 2e-2 * max(1.0, ref.abs().max().item()), (it, err)
    # forward-layout operands (NT) as well
    a = (torch.randn(512, 256, device="cuda") * 0.1).half()
    b = (torch.randn(384, 256, device="cuda") * 0.1).half()
    got = rs(a, b, average=False)
    ref = a.float() @ b.float().t()
    dist.all_reduce(ref)
    ref = ref[rank * (512 // world):(rank + 1) * (512 // world)]
    assert (got.float() - ref).abs().max().item() < 2e-2 * max(1.0, ref.abs().max().item())


def test_fused_gemm_reduce_scatter_2gpu():
    run_distributed(_gemm_rs_worker, world_size=2, use_cuda=True)


def _ag_gemm_worker():
    from deeperspeed_b200.parallel.ag_gemm import ShardedWeight, ag_linear
    torch.manual_seed(3)   # same full weight on every rank; each keeps only its rows
    Nn, Kk, M = 2048, 1024, 512
    w = (torch.randn(Nn, Kk, device="cuda") * 0.05).half()
    b = torch.randn(Nn, device="cuda").half()
    sw = ShardedWeight(w, group=None)
    assert sw.supported()
    torch.manual_seed(10 + dist.get_rank())
    x = torch.randn(M, Kk, device="cuda").half()
    ref = x.float() @ w.float().t() + b.float()
    # overlapped in-kernel all-gather (default, run twice: the panel flags are re-armed per launch) and the direct
    # remote-TMA variant
    for kwargs in ({}, {}, {"direct": True}):
        y = ag_linear(x, sw, b, **kwargs)
        torch.cuda.synchronize()
        rel = ((y.float() - ref).abs().max() / ref.abs().max()).item()
        assert rel < 3e-3, (kwargs, rel)
        dist.barrier()
    # input gradient of the same sharded weight: the shards split the reduction dimension
    from deeperspeed_b200.parallel.ag_gemm import ag_linear_dgrad
    go = (torch.randn(384, Nn, device="cuda") * 0.1).half()
    ref_gi = go.float() @ w.float()
    for _ in range(2):
        gi = ag_linear_dgrad(go, sw)
        torch.cuda.synchronize()
        rel = ((gi.float() - ref_gi).abs().max() / ref_gi.abs().max()).item()
        assert gi.shape == (384, Kk) and rel < 3e-3, rel
        dist.barrier()
    # long reduction (2048 rows per shard): the two-pass form (own range, then remote ranges accumulated on top)
    torch.manual_seed(4)
    w_long = (torch.randn(4096, 512, device="cuda") * 0.05).half()
    sw_long = ShardedWeight(w_long, group=None)
    torch.manual_seed(20 + dist.get_rank())
    go = (torch.randn(640, 4096, device="cuda") * 0.1).half()
    ref_gi = go.float() @ w_long.float()
    gi = ag_linear_dgrad(go, sw_long)
    torch.cuda.synchronize()
    assert ((gi.float() - ref_gi).abs().max() / ref_gi.abs().max()).item() < 3e-3
    dist.barrier()
    # a taller problem so that several waves of tiles wait on remote panels
    x2 = torch.randn(4096, Kk, device="cuda").half()
    y2 = ag_linear(x2, sw, None)
    ref2 = x2.float() @ w.float().t()
    assert ((y2.float() - ref2).abs().max() / ref2.abs().max()).item() < 3e-3
    dist.barrier()


def test_fused_allgather_gemm_2gpu():
    run_distributed(_ag_gemm_worker, world_size=2, use_cuda=True)


def _zero3_fused_ag_worker(tmp):
    """ZeRO-3 training with Linear forwards served by the fused all-gather + GEMM must match the regular path."""
    import deeperspeed_b200 as deepspeed
    from deeperspeed_b200.parallel import ag_gemm
    calls = [0]
    dgrad_calls = [0]
    rs_calls = [0]
    orig = ag_gemm.ag_linear
    orig_dgrad = ag_gemm.ag_linear_dgrad
    orig_rs = ag_gemm.GemmReduceScatter.__call__

    def counting(*a, **k):
        calls[0] += 1
        return orig(*a, **k)

    def counting_dgrad(*a, **k):
        dgrad_calls[0] += 1
        return orig_dgrad(*a, **k)

    def counting_rs(self, *a, **k):
        rs_calls[0] += 1
        return orig_rs(self, *a, **k)
    ag_gemm.ag_linear = counting
    ag_gemm.ag_linear_dgrad = counting_dgrad
    ag_gemm.GemmReduceScatter.__call__ = counting_rs

    class MLP(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.l1 = torch.nn.Linear(256, 1024)
            self.l2 = torch.nn.Linear(1024, 512)
            self.head = torch.nn.Linear(512, 16)      # 16 rows: not eligible, regular gather

        def forward(self, x, y):
            h = torch.nn.functional.gelu(self.l1(x))
            return torch.nn.functional.cross_entropy(self.head(self.l2(h)).float(), y)

    # 64 rows per rank: the weight-gradient GEMM + reduce-scatter kernel applies as well
    cfg = {"train_batch_size": 128, "train_micro_batch_size_per_gpu": 64, "steps_per_print": 100,
           "optimizer": {"type": "Adam", "params": {"lr": 1e-3}},
           "fp16": {"enabled": True, "loss_scale": 128},
           "zero_optimization": {"stage": 3, "stage3_param_persistence_threshold": 10}}
    results = []
    for fused in (False, True):
        torch.manual_seed(7)
        with deepspeed.zero.Init(fused_allgather_gemm=fused):
            model = MLP()
        assert getattr(model.l1.weight, "ds_fused_ag", False) == fused and not model.head.weight.ds_fused_ag
        engine, _, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=cfg)
        g = torch.Generator().manual_seed(100 + dist.get_rank())
        losses = []
        for _ in range(4):
            x = torch.randn(64, 256, generator=g).cuda().half()
            y = torch.randint(0, 16, (64,), generator=g).cuda()
            loss = engine(x, y)
            engine.backward(loss)
            engine.step()
            losses.append(loss.item())
        with deepspeed.zero.GatheredParameters(list(model.parameters())):
            results.append((losses, [p.detach().float().clone() for p in model.parameters()]))
    (l0, p0), (l1, p1) = results
    assert calls[0] == 2 * 4, calls   # l1 and l2 forwards of the fused run, nothing else
    # backward of the fused run: l2's input gradient through the K-sharded all-gather + GEMM (l1's input needs no
    # gradient), both weight gradients through GEMM + reduce-scatter; no NCCL gather / reduce-scatter for them
    assert dgrad_calls[0] == 1 * 4, dgrad_calls
    assert rs_calls[0] == 2 * 4, rs_calls
    assert all(abs(a - b) < 2e-2 for a, b in zip(l0, l1)), (l0, l1)
    for a, b in zip(p0, p1):
        assert (a - b).abs().max().item() < 5e-3
    dist.barrier()


def test_zero3_fused_allgather_gemm_2gpu(tmp_path):
    run_distributed(_zero3_fused_ag_worker, world_size=2, use_cuda=True, args=(str(tmp_path),))


def _zero2_sharded_worker():
    """Stage 2 on 2 GPUs: sharded gradient memory (side-stream reduce-scatter per bucket, pooled staging) must give
    the same parameters as the full gradient arena on the NCCL path."""
    results = []
    for mode, overlap in (("full", False), ("sharded", True), ("sharded", False)):
        torch.manual_seed(5)
        model = Net()
        cfg = {"train_batch_size": 16, "train_micro_batch_size_per_gpu": 4, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-3, "weight_decay": 0.01}},
               "fp16": {"enabled": True, "initial_scale_power": 8}, "gradient_clipping": 1.0,
               "zero_optimization": {"stage": 2, "fused_nvlink": False, "gradient_memory": mode,
                                     "overlap_comm": overlap, "reduce_bucket_size": 200000}}
        engine, opt, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=cfg)
        assert opt.shard_gradients == (mode == "sharded")
        g = torch.Generator().manual_seed(100 + dist.get_rank())
        for _ in range(4 * engine.gradient_accumulation_steps()):
            x = torch.randn(4, 256, generator=g).half().cuda()
            y = torch.randint(0, 256, (4,), generator=g).cuda()
            engine.backward(engine(x, y))
            engine.step()
        torch.cuda.synchronize()
        results.append(torch.cat([p.detach().float().view(-1) for p in model.parameters()]))
        if mode == "sharded":
            own, staged = opt.gradient_memory_bytes()
            total = sum(a.numel * 2 for a in opt.arenas)
            assert sum(opt.num_buckets) > 1 and own == total // 2 and staged < total
    for got in results[1:]:
        d = (got - results[0]).abs()
        # per-micro-batch reduction rounds differently from one reduction of the accumulated arena (fp16)
        assert d.max().item() <= 2 * 1e-3 * 4 + 1e-3 and (d > 2e-3).float().mean().item() < 0.02, d.max().item()


def test_zero2_sharded_gradient_memory_2gpu():
    run_distributed(_zero2_sharded_worker, world_size=2, use_cuda=True)


def _zero2_fused_overlap_worker():
    """ZeRO-2 with overlap_comm on the fused path: the reduce phase of every bucket is launched from the bucket hooks
    on the reduction stream while backward is still running (small grid), the step then runs exchange + update only.
    Must match the NCCL path."""
    results = []
    for fused in (False, True):
        torch.manual_seed(5)
        model = Net()
        cfg = {"train_batch_size": 8, "train_micro_batch_size_per_gpu": 4, "steps_per_print": 100,
               "optimizer": {"type": "Adam", "params": {"lr": 1e-3, "weight_decay": 0.01}},
               "fp16": {"enabled": True, "initial_scale_power": 8}, "gradient_clipping": 1.0,
               "zero_optimization": {"stage": 2, "fused_nvlink": fused, "overlap_comm": True,
                                     "reduce_bucket_size": 200000}}
        engine, opt, _, _ = deepspeed.initialize(model=model, model_parameters=model.parameters(), config_params=cfg)
        assert (opt.fused is not None) == fused and sum(opt.num_buckets) > 1
        g = torch.Generator().manual_seed(100 + dist.get_rank())
        for _ in range(5):
            x = torch.randn(4, 256, generator=g).half().cuda()
            y = torch.randint(0, 256, (4,), generator=g).cuda()
            engine.backward(engine(x, y))
            engine.step()
        torch.cuda.synchronize()
        results.append(torch.cat([p.detach().float().view(-1) for p in model.parameters()]))
        if fused:
            n = opt.fused.launches
            assert n["reduce_early"] >= 5 * (sum(opt.num_buckets) - 1), n      # buckets were reduced under backward
            assert n["reduce"] <= 5, n
    d = (results[1] - results[0]).abs()
    assert d.max().item() <= 2 * 1e-3 * 5 + 1e-3 and (d > 2e-3).float().mean().item() < 0.02, d.max().item()


def test_zero2_fused_reduce_overlaps_backward_2gpu():
    run_distributed(_zero2_fused_overlap_worker, world_size=2, use_cuda=True)
