"""Host-side native library (libdsb200_host.so) and the components built on it: CPU Adam, async I/O, flatten,
swap buffers / async swapper, contiguous allocator, op builders, environment report. CPU only.
Modelled on the reference's tests/unit/test_cpu_adam.py, test_aio.py and the swap_tensor users in stage3."""
import os

import pytest
import torch

import deeperspeed_b200 as deepspeed


# ----------------------------------------------------------------------------------------------------- CPU Adam
@pytest.mark.parametrize("adamw", [True, False])
@pytest.mark.parametrize("numel", [22, 1024, 1 << 16 | 37])
def test_cpu_adam_matches_torch(adamw, numel):
    from deeperspeed_b200.ops.adam import DeepSpeedCPUAdam
    torch.manual_seed(0)
    p_ref = torch.nn.Parameter(torch.randn(numel))
    p_ds = torch.nn.Parameter(p_ref.detach().clone())
    cls = torch.optim.AdamW if adamw else torch.optim.Adam
    ref = cls([p_ref], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.05)
    ds = DeepSpeedCPUAdam([p_ds], lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.05, adamw_mode=adamw)
    for step in range(5):
        g = torch.randn(numel)
        p_ref.grad = g.clone()
        p_ds.grad = g.clone()
        ref.step()
        ds.step()
    assert torch.allclose(p_ds, p_ref, rtol=2e-5, atol=2e-6)
    st = ds.state[p_ds]
    assert st["step"] == 5 and torch.allclose(st["exp_avg"], ref.state[p_ref]["exp_avg"], rtol=1e-5, atol=1e-7)


def test_cpu_adam_state_dict_roundtrip():
    from deeperspeed_b200.ops.adam import DeepSpeedCPUAdam
    p = torch.nn.Parameter(torch.randn(300))
    opt = DeepSpeedCPUAdam([p], lr=1e-3)
    p.grad = torch.randn(300)
    opt.step()
    import copy
    sd = copy.deepcopy(opt.state_dict())   # load_state_dict keeps references to same-dtype tensors
    q = torch.nn.Parameter(p.detach().clone())
    opt2 = DeepSpeedCPUAdam([q], lr=1e-3)
    opt2.load_state_dict(sd)
    g = torch.randn(300)
    p.grad, q.grad = g.clone(), g.clone()
    opt.step()
    opt2.step()
    assert torch.equal(p, q)


# ----------------------------------------------------------------------------------------------------- async I/O
@pytest.mark.parametrize("single_submit,overlap_events", [(False, True), (True, False)])
@pytest.mark.parametrize("num_threads", [1, 3])
def test_aio_handle_roundtrip(tmp_path, single_submit, overlap_events, num_threads):
    from deeperspeed_b200.ops.aio import aio_handle
    h = aio_handle(block_size=64 * 1024, queue_depth=4, single_submit=single_submit, overlap_events=overlap_events,
                   num_threads=num_threads)
    assert h.get_block_size() == 64 * 1024 and h.get_queue_depth() == 4 and h.get_thread_count() == num_threads
    assert h.get_single_submit() == single_submit and h.get_overlap_events() == overlap_events
    src = torch.randint(0, 255, (3 * 1024 * 1024 + 512,), dtype=torch.uint8)
    f = str(tmp_path / "blob.bin")
    assert h.sync_pwrite(src, f) == 1
    assert os.path.getsize(f) == src.numel()
    dst = torch.zeros_like(src)
    assert h.sync_pread(dst, f) == 1
    assert torch.equal(src, dst)
    # asynchronous pair: several requests in flight, wait() returns how many completed
    a, b = torch.randn(100000), torch.randn(70001)
    h.async_pwrite(a, str(tmp_path / "a.bin"))
    h.async_pwrite(b, str(tmp_path / "b.bin"))
    assert h.wait() == 2
    ra, rb = torch.empty_like(a), torch.empty_like(b)
    h.async_pread(ra, str(tmp_path / "a.bin"))
    h.async_pread(rb, str(tmp_path / "b.bin"))
    assert h.wait() == 2
    assert torch.equal(a, ra) and torch.equal(b, rb)


def test_aio_free_functions_and_memcpy(tmp_path):
    from deeperspeed_b200.ops import aio
    t = torch.arange(50000, dtype=torch.float32)
    f = str(tmp_path / "t.bin")
    aio.aio_write(t, f, block_size=8192, queue_depth=2)
    r = torch.empty_like(t)
    aio.aio_read(r, f, block_size=8192, queue_depth=2, validate=True)
    assert torch.equal(t, r)
    d = torch.empty_like(t)
    aio.deepspeed_memcpy(d, t)
    assert torch.equal(d, t)
    assert aio.AsyncIOBuilder().is_compatible()


# ----------------------------------------------------------------------------------------------------- swap tensors
def test_swap_buffers_and_async_swapper(tmp_path):
    from deeperspeed_b200.ops.aio import aio_handle
    from deeperspeed_b200.runtime.swap_tensor.async_swapper import AsyncTensorSwapper
    from deeperspeed_b200.runtime.swap_tensor.utils import SwapBufferManager, SwapBufferPool, swap_in_tensors
    h = aio_handle(block_size=4096, queue_depth=4, num_threads=1)
    mgr = SwapBufferManager(num_elems=4096, count=3, dtype=torch.float32)
    bufs = mgr.allocate(num_elems=4096, count=2, dtype=torch.float32)
    assert len(bufs) == 2 and all(b.is_pinned() or not torch.cuda.is_available() for b in bufs)
    pool = SwapBufferPool(bufs)
    t1, t2 = torch.randn(1000), torch.randn(3000)
    p1, p2 = str(tmp_path / "t1.swp"), str(tmp_path / "t2.swp")
    assert pool.has_space(1024)
    pool.insert_tensor(t1, p1, 1024)
    pool.insert_tensor(t2, p2, 3072)
    pool.swap_out(h)
    assert os.path.getsize(p1) == 1024 * 4 and os.path.getsize(p2) == 3072 * 4
    r1, r2 = torch.zeros(1024), torch.zeros(3072)
    swap_in_tensors(h, [r1, r2], [p1, p2])
    assert h.wait() == 2
    assert torch.equal(r1[:1000], t1) and torch.equal(r2[:3000], t2)
    mgr.free(bufs)
    # the asynchronous swapper batches tensors through its buffers and flushes on release
    sw = AsyncTensorSwapper(h, numel_alignment=1024, timers=None)
    sw.add_buffers(mgr.allocate(num_elems=4096, count=2, dtype=torch.float32))
    ts = [torch.randn(1500) for _ in range(5)]
    paths = [str(tmp_path / f"s{i}.swp") for i in range(5)]
    sw.swap_out_tensors(ts, paths)
    sw.release_buffers()
    for t, p in zip(ts, paths):
        r = torch.zeros(2048)
        assert h.sync_pread(r, p) == 1
        assert torch.equal(r[:1500], t)


# ----------------------------------------------------------------------------------------------------- allocator / flatten
def test_contiguous_memory_allocator_defragments():
    from deeperspeed_b200.runtime.zero.contiguous_memory_allocator import ContiguousMemoryAllocator
    mem = ContiguousMemoryAllocator(1024, torch.float32, "cpu")
    a = mem.allocate_tensor(256)
    b = mem.allocate_tensor(256)
    c = mem.allocate_tensor(256)
    a.fill_(1.0); b.fill_(2.0); c.fill_(3.0)
    mem.release_tensor(b)
    # 512 free but not contiguous (256 hole + 256 tail): the allocator must compact and keep a/c contents
    p = torch.nn.Parameter(torch.zeros(1))
    mem.assign_to_param(c, p, 256, (16, 16))
    d = mem.allocate_tensor(512)
    d.fill_(4.0)
    assert mem.max_allocated() == 1024
    assert torch.all(p.data == 3.0) and p.data.shape == (16, 16)
    assert float(a.sum()) == 256.0 and float(d.sum()) == 2048.0
    mem.release_tensor(d)
    with pytest.raises(AssertionError):
        mem.allocate_tensor(2048)


def test_native_flatten_unflatten():
    from deeperspeed_b200.ops.op_builder import UtilsBuilder
    util = UtilsBuilder().load()
    ts = [torch.randn(3, 5), torch.randn(7), torch.randn(2, 2, 2)]
    flat = util.flatten(ts)
    assert flat.numel() == 15 + 7 + 8 and torch.equal(flat[:15].view(3, 5), ts[0])
    outs = util.unflatten(flat, ts)
    assert all(torch.equal(o, t) and o.shape == t.shape for o, t in zip(outs, ts))


# ----------------------------------------------------------------------------------------------------- builders / report
def test_op_builders_and_env_report(capsys):
    from deeperspeed_b200.ops import op_builder
    assert len(op_builder.ALL_OPS) >= 7
    for name, b in op_builder.ALL_OPS.items():
        assert isinstance(b.is_compatible(), bool)
        assert name in b.absolute_name()
        assert b.sources()
    from deeperspeed_b200 import env_report
    env_report.main()
    out = capsys.readouterr().out
    assert "cpu_adam" in out and "torch version" in out.lower()
