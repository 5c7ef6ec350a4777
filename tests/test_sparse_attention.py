"""Block-sparse attention: layouts, matmul modes and softmax against dense-masked oracles
(in the spirit of reference tests/unit/test_sparse_attention.py). CPU runs the emulation path; the GPU tests check
the sm_100a kernels against the same oracle."""
import pytest
import torch

from deeperspeed_b200.ops.sparse_attention import (MatMul, Softmax, SparseSelfAttention, FixedSparsityConfig,
                                                   VariableSparsityConfig, BigBirdSparsityConfig,
                                                   BSLongformerSparsityConfig, DenseSparsityConfig,
                                                   LocalSlidingWindowSparsityConfig)
from deeperspeed_b200.ops.sparse_attention.matmul import dense_to_sparse, sparse_to_dense, get_lut


def _layout(H, nb, seed=0):
    g = torch.Generator().manual_seed(seed)
    layout = (torch.rand(H, nb, nb, generator=g) < 0.4).long()
    layout[:, torch.arange(nb), torch.arange(nb)] = 1
    return layout


def test_layouts():
    f = FixedSparsityConfig(num_heads=4, block=16, num_local_blocks=4, num_global_blocks=1).make_layout(256)
    assert f.shape == (4, 16, 16) and f[0, 0, :4].all() and f[0, 5, 3] == 1 and f[0, 0, 4] == 0
    assert (f[0] == f[3]).all()
    u = FixedSparsityConfig(num_heads=2, block=16, attention="unidirectional").make_layout(128)
    assert torch.equal(u[0], torch.tril(u[0]))
    assert DenseSparsityConfig(num_heads=2).make_layout(64).all()
    v = VariableSparsityConfig(num_heads=2, num_random_blocks=1, local_window_blocks=[2, 4], global_block_indices=[0]).make_layout(256)
    assert v[0, :, 0].all() and v[0, 0, 1] == 1
    bb = BigBirdSparsityConfig(num_heads=2, num_random_blocks=1, num_sliding_window_blocks=3, num_global_blocks=1).make_layout(256)
    assert bb[0, 0].all() and bb[0, :, 0].all() and bb[0, 5, 4] == 1 and bb[0, 5, 6] == 1
    lf = BSLongformerSparsityConfig(num_heads=2, global_block_indices=[0, 5], global_block_end_indices=[1, 7]).make_layout(256)
    assert lf[0, 5].all() and lf[0, :, 6].all()
    sw = LocalSlidingWindowSparsityConfig(num_heads=2, num_sliding_window_blocks=3).make_layout(128)
    assert sw[0, 3, 2] == 1 and sw[0, 3, 4] == 0
    with pytest.raises(ValueError):
        FixedSparsityConfig(num_heads=2).make_layout(100)


def test_native_luts_and_segment():
    layout = _layout(2, 8)
    lut = get_lut(layout, 16)
    assert lut.nnz == int(layout.sum())
    nz = layout.nonzero()
    assert torch.equal(lut.blocks[:, :3].long(), nz)
    for hr in range(16):
        cols = lut.row_idx[lut.row_ptr[hr]:lut.row_ptr[hr + 1], 0].long()
        assert torch.equal(cols, layout.view(16, 8)[hr].nonzero().flatten())
    ids, widths, counts = lut.sdd_segment(4)
    assert ((ids >= 0) == (layout > 0)).all()
    assert int((counts * torch.tensor([1, 4, 16])).sum()) == lut.nnz


def _run_matmul_case(mode, trans_a, trans_b, device, dtype, block=16):
    torch.manual_seed(0)
    Bz, H, nb, D = 2, 2, 4, 32
    S = nb * block
    layout = _layout(H, nb, seed=3)

    def dense(r, c):
        return (torch.randn(Bz, H, r, c, device=device) * 0.5).to(dtype).requires_grad_()

    if mode == "sdd":
        a = dense(D, S) if trans_a else dense(S, D)
        b = dense(S, D) if trans_b else dense(D, S)
    elif mode == "dsd":
        a = (torch.randn(Bz, int(layout.sum()), block, block, device=device) * 0.5).to(dtype).requires_grad_()
        b = dense(D, S) if trans_b else dense(S, D)
    else:
        a = dense(S, D) if trans_a else dense(D, S)
        b = (torch.randn(Bz, int(layout.sum()), block, block, device=device) * 0.5).to(dtype).requires_grad_()
    op = MatMul(layout, block, mode, trans_a=trans_a, trans_b=trans_b)
    c = op(a, b)
    g = torch.randn_like(c)
    c.backward(g)
    # dense oracle
    af = a.detach().float().cpu().requires_grad_()
    bf = b.detach().float().cpu().requires_grad_()
    A = sparse_to_dense(af, layout, block) if mode == "dsd" else af
    Bm = sparse_to_dense(bf, layout, block) if mode == "dds" else bf
    A = A.transpose(-1, -2) if trans_a else A
    Bm = Bm.transpose(-1, -2) if trans_b else Bm
    ref = A @ Bm
    if mode == "sdd":
        ref = dense_to_sparse(ref, layout, block)
    ref.backward(g.float().cpu())
    tol = 2e-2 if dtype != torch.float32 else 1e-4

    def rel(x, y):
        return ((x.float().cpu() - y).abs().max() / (y.abs().max() + 1e-6)).item()

    assert rel(c.detach(), ref.detach()) < tol
    assert rel(a.grad, af.grad) < tol and rel(b.grad, bf.grad) < tol


@pytest.mark.parametrize("mode", ["sdd", "dsd", "dds"])
@pytest.mark.parametrize("trans_a", [False, True])
@pytest.mark.parametrize("trans_b", [False, True])
def test_matmul_cpu(mode, trans_a, trans_b):
    _run_matmul_case(mode, trans_a, trans_b, "cpu", torch.float32)


@pytest.mark.gpu
@pytest.mark.parametrize("block", [16, 32, 64])
@pytest.mark.parametrize("mode,trans_a,trans_b", [("sdd", False, True), ("sdd", False, False), ("dsd", False, False),
                                                  ("dsd", True, False), ("dds", False, False), ("dds", False, True)])
def test_matmul_gpu(mode, trans_a, trans_b, block):
    _run_matmul_case(mode, trans_a, trans_b, "cuda", torch.float16, block=block)


def _run_softmax_case(device, dtype, block=16):
    torch.manual_seed(1)
    Bz, H, nb = 2, 2, 4
    S = nb * block
    layout = _layout(H, nb, seed=5)
    x = (torch.randn(Bz, int(layout.sum()), block, block, device=device)).to(dtype).requires_grad_()
    kpm = torch.zeros(Bz, S, device=device, dtype=dtype)
    kpm[:, -3:] = -10000.0
    am = torch.tril(torch.ones(S, S, device=device, dtype=dtype))
    rpe = (torch.randn(H, S, S, device=device) * 0.1).to(dtype)
    sm = Softmax(layout, block)
    y = sm(x, scale=0.5, rpe=rpe, key_padding_mask=kpm, attn_mask=am, key_padding_mask_mode="add", attn_mask_mode="mul")
    g = torch.randn_like(y)
    y.backward(g)
    xf = x.detach().float().cpu().requires_grad_()
    d = sparse_to_dense(xf, layout, block, zero=float("-inf")) * 0.5 + rpe.float().cpu() + kpm.float().cpu()[:, None, None, :]
    d = d.masked_fill(am.float().cpu()[None, None] == 0, float("-inf"))
    p = torch.softmax(d, -1)
    p = torch.where(torch.isnan(p), torch.zeros_like(p), p)
    ref = dense_to_sparse(p, layout, block)
    ref.backward(g.float().cpu())
    tol = 2e-2 if dtype != torch.float32 else 1e-5
    assert (y.detach().float().cpu() - ref.detach()).abs().max() < tol
    assert (x.grad.float().cpu() - xf.grad).abs().max() < tol


def test_softmax_cpu():
    _run_softmax_case("cpu", torch.float32)


@pytest.mark.gpu
@pytest.mark.parametrize("block", [16, 32])
def test_softmax_gpu(block):
    _run_softmax_case("cuda", torch.float16, block)


def _attn_case(device, dtype):
    torch.manual_seed(2)
    Bz, H, S, D = 2, 4, 128, 32
    cfg = FixedSparsityConfig(num_heads=H, block=16, num_local_blocks=2, num_global_blocks=1)
    attn = SparseSelfAttention(cfg, max_seq_length=S)
    SparseSelfAttention.ops.clear()
    q, k, v = ((torch.randn(Bz, H, S, D, device=device) * 0.5).to(dtype) for _ in range(3))
    out = attn(q, k, v)
    layout = cfg.make_layout(S)
    mask = sparse_to_dense(torch.ones(1, int(layout.sum()), 16, 16), layout, 16)[0].bool()
    sc = (q.float().cpu() @ k.float().cpu().transpose(-1, -2)) * D ** -0.5
    sc = sc.masked_fill(~mask[None], float("-inf"))
    ref = torch.softmax(sc, -1) @ v.float().cpu()
    assert (out.float().cpu() - ref).abs().max() < (3e-2 if dtype != torch.float32 else 1e-4)


def test_sparse_self_attention_cpu():
    _attn_case("cpu", torch.float32)


@pytest.mark.gpu
def test_sparse_self_attention_gpu():
    _attn_case("cuda", torch.float16)


# ----------------------------------------------------------------------------------------------- fused flash path
def test_fused_schedules_cover_exactly_the_layout():
    """Every visible 16 x 16 block of the layout has its bit in exactly one listed tile, forward and backward."""
    from deeperspeed_b200.ops.sparse_attention.fused import build_schedules
    torch.manual_seed(3)
    S, block, H = 512, 32, 3
    nb = S // block
    layout = (torch.rand(H, nb, nb) < 0.25).long()
    layout[:, torch.arange(nb), torch.arange(nb)] = 1
    sub = layout.repeat_interleave(block // 16, 1).repeat_interleave(block // 16, 2)     # 16-granular visibility
    fwd, bwd = build_schedules(layout, block, S)
    assert fwd.shape[:2] == (H, S // 256) and bwd.shape[:2] == (H, S // 128)
    seen_f, seen_b = torch.zeros_like(sub), torch.zeros_like(sub)
    for h in range(H):
        for p in range(S // 256):
            rec = fwd[h, p].tolist()
            for e in range(rec[0]):
                kb, words = rec[1 + 3 * e], rec[2 + 3 * e:4 + 3 * e]
                for t, w in enumerate(words):
                    for bit in range(32):
                        if (w >> bit) & 1:
                            seen_f[h, p * 16 + t * 8 + bit // 4, kb * 4 + bit % 4] += 1
        for jb in range(S // 128):
            rec = bwd[h, jb].tolist()
            for e in range(rec[0]):
                qb, w = rec[1 + 2 * e], rec[2 + 2 * e]
                for bit in range(32):
                    if (w >> bit) & 1:
                        seen_b[h, qb * 4 + bit % 4, jb * 8 + bit // 4] += 1
    assert torch.equal(seen_f, sub) and torch.equal(seen_b, sub)
    # identical layouts on every head collapse to one record set
    f1, b1 = build_schedules(layout[:1].expand(H, -1, -1).contiguous(), block, S)
    assert f1.shape[0] == 1 and b1.shape[0] == 1


def _fused_case(S, D, block, per_head, key_mask, dtype=torch.float16):
    from deeperspeed_b200.ops.sparse_attention import VariableSparsityConfig
    torch.manual_seed(4)
    Bz, H = 2, 4
    if per_head:
        cfg = VariableSparsityConfig(num_heads=H, block=block, different_layout_per_head=True, num_random_blocks=2,
                                     local_window_blocks=[2], global_block_indices=[0])
    else:
        cfg = FixedSparsityConfig(num_heads=H, block=block, num_local_blocks=4, num_global_blocks=1,
                                  attention="unidirectional")
    SparseSelfAttention.ops.clear()
    SparseSelfAttention.schedules.clear()
    attn = SparseSelfAttention(cfg, max_seq_length=S).cuda()
    q, k, v = ((torch.randn(Bz, H, S, D, device="cuda") * 0.7).to(dtype).requires_grad_() for _ in range(3))
    kpm = None
    if key_mask:
        kpm = torch.zeros(Bz, S, device="cuda", dtype=dtype)
        kpm[:, -37:] = -10000.0
    out = attn(q, k, v, key_padding_mask=kpm)
    assert "BlockSparseFlashFn" in type(out.grad_fn).__name__ or "Transpose" in type(out.grad_fn).__name__
    g = torch.randn_like(out)
    out.backward(g)
    layout = attn.get_layout(S)
    vis = sparse_to_dense(torch.ones(1, int(layout.sum()), block, block), layout, block)[0].bool().cuda()
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    sc = (qf @ kf.transpose(-1, -2)) * D ** -0.5
    if kpm is not None:
        sc = sc + kpm.float()[:, None, None, :]
    sc = sc.masked_fill(~vis[None], float("-inf"))
    ref = torch.softmax(sc, -1) @ vf
    ref.backward(g.float())
    assert (out.float() - ref).abs().max() < 3e-3
    for got, want in ((q.grad, qf.grad), (k.grad, kf.grad), (v.grad, vf.grad)):
        assert (got.float() - want).abs().max() < 2e-2 * max(1.0, want.abs().max().item())


@pytest.mark.gpu
@pytest.mark.parametrize("S,D,block", [(256, 64, 16), (512, 64, 32), (768, 128, 16), (1024, 32, 64)])
@pytest.mark.parametrize("per_head", [False, True])
@pytest.mark.parametrize("key_mask", [False, True])
def test_fused_block_sparse_attention(S, D, block, per_head, key_mask):
    """The tcgen05 block-sparse flash path (forward + backward) against dense masked fp32 attention."""
    _fused_case(S, D, block, per_head, key_mask)


@pytest.mark.gpu
def test_fused_block_sparse_attention_falls_back_for_rpe():
    """Relative position embeddings are not part of the fused kernel: the three separate operators run."""
    torch.manual_seed(5)
    cfg = FixedSparsityConfig(num_heads=2, block=16, num_local_blocks=4, num_global_blocks=1)
    attn = SparseSelfAttention(cfg, max_seq_length=256).cuda()
    q, k, v = ((torch.randn(1, 2, 256, 64, device="cuda") * 0.5).half() for _ in range(3))
    rpe = (torch.randn(2, 256, 256, device="cuda") * 0.1).half()
    a = attn(q, k, v, rpe=rpe)
    SparseSelfAttention.use_fused = False
    try:
        b = attn(q, k, v, rpe=rpe)
    finally:
        SparseSelfAttention.use_fused = True
    assert torch.equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("D,dtype,p", [(128, torch.float16, 0.1), (64, torch.bfloat16, 0.25)])
def test_flash_attention_dropout_matches_masked_reference(D, dtype, p):
    """Attention dropout inside the flash kernels: the keep mask is a hash of (seed, batch, head, query, key) that
    forward and backward both derive; `dropout_keep_mask` restates it, so the result can be checked EXACTLY against
    softmax -> mask / keep -> matmul in fp32, gradients included (a wrong or inconsistent mask fails loudly)."""
    from deeperspeed_b200.ops.sparse_attention import fused
    torch.manual_seed(0)
    B, H, S = 2, 3, 384
    seed = 0xC0FFEE
    q, k, v = (torch.randn(B, H, S, D, device="cuda", dtype=dtype).requires_grad_() for _ in range(3))
    scale = D ** -0.5
    out = fused.BlockSparseFlashFn.apply(q, k, v, None, None, None, scale, (p, seed))
    dout = torch.randn_like(out)
    dq, dk, dv = torch.autograd.grad(out, (q, k, v), dout)
    keep, inv_keep = fused.dropout_keep_mask(B, H, S, p, seed, device="cuda")
    assert abs(keep.float().mean().item() - (1 - p)) < 0.01
    assert abs(inv_keep - 1 / (1 - p)) < 1e-3
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    probs = torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1) * keep * inv_keep
    ref = probs @ vf
    rq, rk, rv = torch.autograd.grad(ref, (qf, kf, vf), dout.float())
    tol = 2e-2 if dtype == torch.float16 else 6e-2
    for got, want, name in ((out, ref, "out"), (dq, rq, "dq"), (dk, rk, "dk"), (dv, rv, "dv")):
        err = (got.float() - want).abs().max().item() / (want.abs().max().item() + 1e-6)
        assert err < tol, (name, err)
    # a different seed gives a different mask; p = 0 is the plain kernel
    other = fused.BlockSparseFlashFn.apply(q, k, v, None, None, None, scale, (p, seed + 1))
    assert (other - out).abs().max().item() > 1e-3
    plain = fused.BlockSparseFlashFn.apply(q, k, v, None, None, None, scale, None)
    ref0 = torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1) @ vf
    assert (plain.float() - ref0).abs().max().item() / ref0.abs().max().item() < tol


def test_key_order_plan_is_a_permutation_that_cuts_tile_work():
    """fused.plan_key_order: the fixed pattern's global columns (one 16-key block per window) are moved to the front,
    in index order; the result is a permutation, the permuted layout is the layout with its columns reordered, and the
    schedules need far fewer tiles. Layouts that gain nothing (or differ per head) keep their order."""
    from deeperspeed_b200.ops.sparse_attention import fused, BigBirdSparsityConfig, VariableSparsityConfig
    S, block = 4096, 16
    cfg = FixedSparsityConfig(num_heads=4, block=block, num_local_blocks=8, num_global_blocks=1, attention="unidirectional")
    layout = cfg.make_layout(S)
    order, planned = fused.plan_key_order(layout, block, S)
    assert order is not None and sorted(order.tolist()) == list(range(S))
    assert fused.plan_key_order(cfg.make_layout(1024), block, 1024)[0] is None      # too little to gain at 1k tokens
    blocks = order.view(-1, block)[:, 0] // block
    assert torch.equal(planned, layout[:, :, blocks])
    # whole blocks move, tokens inside a block keep their order
    assert torch.equal(order.view(-1, block), blocks.view(-1, 1) * block + torch.arange(block))
    work = lambda lay: fused._tile_work(*fused.build_schedules(lay, block, S))
    assert work(planned) < 0.5 * work(layout)
    # the moved columns are the widely seen ones, still ascending (a causal prefix of them stays a prefix)
    moved = blocks[:int((blocks[:-1] > blocks[1:]).nonzero()[0]) + 1]
    assert torch.equal(moved, moved.sort().values) and bool((moved % 8 == 7).all())
    bb = BigBirdSparsityConfig(num_heads=2, block=block, num_random_blocks=1, num_global_blocks=2).make_layout(S)
    assert fused.plan_key_order(bb, block, S)[0] is None          # global blocks already sit together at the front
    var = VariableSparsityConfig(num_heads=2, block=block, different_layout_per_head=True, num_random_blocks=2).make_layout(S)
    assert fused.plan_key_order(var, block, S)[0] is None         # per-head layouts: one shared order does not exist


@pytest.mark.gpu
@pytest.mark.parametrize("key_mask", [False, True])
def test_fused_block_sparse_attention_with_reordered_keys(key_mask):
    """Long fixed layout: the module visits the keys in the planned order (gathers K, V and the key mask) - same result
    as dense masked attention, gradients included, and the same as with the reordering switched off."""
    from deeperspeed_b200.ops.sparse_attention import fused
    torch.manual_seed(6)
    Bz, H, S, D, block = 1, 4, 4096, 64, 16
    cfg = FixedSparsityConfig(num_heads=H, block=block, num_local_blocks=8, num_global_blocks=1, attention="unidirectional")
    SparseSelfAttention.schedules.clear()
    attn = SparseSelfAttention(cfg, max_seq_length=S).cuda()
    assert attn.get_schedules(S, torch.device("cuda", 0))[2] is not None
    q, k, v = ((torch.randn(Bz, H, S, D, device="cuda") * 0.7).half().requires_grad_() for _ in range(3))
    kpm = None
    if key_mask:
        kpm = torch.zeros(Bz, S, device="cuda", dtype=torch.float16)
        kpm[:, 100:180] = -10000.0
    out = attn(q, k, v, key_padding_mask=kpm)
    g = torch.randn_like(out)
    grads = torch.autograd.grad(out, (q, k, v), g)
    plain = SparseSelfAttention(cfg, max_seq_length=S).cuda()
    plain.reorder_keys = False
    assert plain.get_schedules(S, torch.device("cuda", 0))[2] is None
    out0 = plain(q, k, v, key_padding_mask=kpm)
    grads0 = torch.autograd.grad(out0, (q, k, v), g)
    assert (out.float() - out0.float()).abs().max() < 2e-3
    for a, b in zip(grads, grads0):
        assert (a.float() - b.float()).abs().max() < 1e-2 * max(1.0, b.float().abs().max().item())
    layout = attn.get_layout(S)
    vis = sparse_to_dense(torch.ones(1, int(layout.sum()), block, block), layout, block)[0].bool().cuda()
    sc = (q.float() @ k.float().transpose(-1, -2)) * D ** -0.5
    if kpm is not None:
        sc = sc + kpm.float()[:, None, None, :]
    ref = torch.softmax(sc.masked_fill(~vis[None], float("-inf")), -1) @ v.float()
    assert (out.float() - ref).abs().max() < 3e-3


@pytest.mark.gpu
def test_fused_rows_that_see_nothing_in_their_first_tile():
    """Regression: a query row whose first live key tile holds nothing visible FOR IT (other rows of its warp do see
    keys there) used to pick up a NaN rescale factor (-inf - -inf) and come out as zeros."""
    from deeperspeed_b200.ops.sparse_attention import SparsityConfig

    class Crossed(SparsityConfig):
        def make_layout(self, seq_len):
            layout = self.setup_layout(seq_len)
            nb = layout.shape[1]
            for i in range(nb):
                layout[:, i, (nb // 2 + i) % nb if i % 2 == 0 else i // 2] = 1      # even rows look far ahead, odd rows back
                layout[:, i, i] = 1
            return layout

    torch.manual_seed(8)
    S, D, H, block = 512, 128, 2, 16
    SparseSelfAttention.schedules.clear()
    attn = SparseSelfAttention(Crossed(num_heads=H, block=block), max_seq_length=S).cuda()
    q, k, v = ((torch.randn(1, H, S, D, device="cuda") * 0.7).half().requires_grad_() for _ in range(3))
    out = attn(q, k, v)
    g = torch.randn_like(out)
    grads = torch.autograd.grad(out, (q, k, v), g)
    layout = attn.get_layout(S)
    vis = sparse_to_dense(torch.ones(1, int(layout.sum()), block, block), layout, block)[0].bool().cuda()
    qf, kf, vf = (t.detach().float().requires_grad_() for t in (q, k, v))
    ref = torch.softmax(((qf @ kf.transpose(-1, -2)) * D ** -0.5).masked_fill(~vis[None], float("-inf")), -1) @ vf
    refg = torch.autograd.grad(ref, (qf, kf, vf), g.float())
    assert torch.isfinite(out).all() and (out.float() - ref).abs().max() < 3e-3
    for a, b in zip(grads, refg):
        assert (a.float() - b).abs().max() < 2e-2 * max(1.0, b.abs().max().item())


def test_permute_keys_gradient_and_dropout_mask_restatement_cpu():
    """CPU checks of the fused path's Python helpers: the key permutation's backward is the inverse gather, and the
    restated dropout hash is deterministic, seed-dependent and keeps the requested fraction."""
    from deeperspeed_b200.ops.sparse_attention import fused
    torch.manual_seed(0)
    x = torch.randn(2, 3, 8, 4, requires_grad=True)
    order = torch.randperm(8)
    inverse = torch.empty_like(order)
    inverse[order] = torch.arange(8)
    y = fused.PermuteKeysFn.apply(x, order, inverse)
    assert torch.equal(y, x[:, :, order])
    w = torch.randn_like(y)
    (y * w).sum().backward()
    expect = torch.zeros_like(x)
    expect[:, :, order] = w
    assert torch.equal(x.grad, expect)
    keep, inv_keep = fused.dropout_keep_mask(2, 3, 256, 0.25, seed=7)
    again, _ = fused.dropout_keep_mask(2, 3, 256, 0.25, seed=7)
    other, _ = fused.dropout_keep_mask(2, 3, 256, 0.25, seed=8)
    assert torch.equal(keep, again) and not torch.equal(keep, other)
    assert abs(keep.float().mean().item() - 0.75) < 0.01 and abs(inv_keep - 1 / 0.75) < 1e-3
    # heads and batches get different masks
    assert not torch.equal(keep[0, 0], keep[0, 1]) and not torch.equal(keep[0, 0], keep[1, 0])
    none, one = fused.dropout_keep_mask(1, 1, 64, 0.0, seed=1)
    assert bool(none.all()) and one == 1.0
