"""Numerics of every sm_100a kernel against a plain PyTorch fp32 reference of the same op."""
import pytest
import torch

from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.ops import native as N

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).abs().max() / (b.abs().max() + 1e-6)).item()


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(128, 256, 64), (256, 512, 128), (1000, 264, 200), (4096, 2048, 2048), (384, 128, 320),
                                   (77, 64, 72)])
@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, False), (True, True)])
def test_gemm_layouts(dtype, shape, ta, tb):
    M, Nn, Kk = shape
    torch.manual_seed(0)
    a = torch.randn((Kk, M) if ta else (M, Kk), device=DEV, dtype=dtype)
    b = torch.randn((Nn, Kk) if tb else (Kk, Nn), device=DEV, dtype=dtype)
    if (M % 8 and ta) or (Nn % 8):
        pytest.skip("leading dims must be multiples of 8")
    out = K.gemm(a, b, trans_a=ta, trans_b=tb)
    ref = (a.float().t() if ta else a.float()) @ (b.float().t() if tb else b.float())
    tol = 2e-3 if dtype == torch.float16 else 1.5e-2
    assert _rel(out, ref) < tol, _where(out, ref, tol)


def _where(out, ref, tol):
    """Failure message: where the mismatching elements sit (which tile / CTA half / epilogue chunk)."""
    err = (out.float() - ref).abs()
    bad = (err > tol * ref.abs().max()).nonzero()
    if bad.numel() == 0:
        return "no element above tolerance"
    r, c = bad[:, -2], bad[:, -1]
    return (f"{bad.shape[0]} bad elements, max err {err.max().item():.4f}, rows {r.min().item()}..{r.max().item()}, "
            f"cols {c.min().item()}..{c.max().item()}, first {bad[:8].tolist()}")


@pytest.mark.parametrize("epi", ["bias", "bias_gelu", "residual", "dgelu", "accum"])
def test_gemm_epilogues(epi):
    torch.manual_seed(1)
    M, Nn, Kk = 520, 1032, 512
    dt = torch.float16
    a = torch.randn(M, Kk, device=DEV, dtype=dt) * 0.1
    b = torch.randn(Nn, Kk, device=DEV, dtype=dt) * 0.1
    bias = torch.randn(Nn, device=DEV, dtype=dt)
    aux = torch.randn(M, Nn, device=DEV, dtype=dt)
    acc = a.float() @ b.float().t()
    if epi == "bias":
        out = K.gemm(a, b, epilogue=N.EPI_BIAS, bias=bias)
        ref = acc + bias.float()
    elif epi == "bias_gelu":
        pre = torch.empty(M, Nn, device=DEV, dtype=dt)
        out = K.gemm(a, b, epilogue=N.EPI_BIAS_GELU, bias=bias, aux_out=pre)
        ref = K.gelu_tanh(acc + bias.float())
        assert _rel(pre, acc + bias.float()) < 2e-3
    elif epi == "residual":
        out = K.gemm(a, b, epilogue=N.EPI_BIAS_RESIDUAL, bias=bias, aux=aux)
        ref = acc + bias.float() + aux.float()
    elif epi == "dgelu":
        out = K.gemm(a, b, epilogue=N.EPI_DGELU, aux=aux)
        ref = acc * K.gelu_tanh_grad(aux.float())
    else:
        out = aux.clone()
        K.gemm(a, b, epilogue=N.EPI_BIAS_RESIDUAL, aux=out, out=out)
        ref = acc + aux.float()
    assert _rel(out, ref) < 3e-3


@pytest.fixture(params=[256, 128])
def pair_kernel(monkeypatch, request):
    """Force the CTA-pair kernel with 256- and 128-wide pair tiles."""
    monkeypatch.setattr(K, "GEMM_KERNEL", "pair")
    lib = N.cuda_lib()
    assert lib.dsb_gemm_pair_force_bn(request.param) == 0
    yield request.param
    lib.dsb_gemm_pair_force_bn(0)


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(256, 256, 64), (512, 768, 448), (1000, 264, 200), (4096, 2048, 2048),
                                   (8192, 2048, 1024), (77, 64, 72)])
@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, False), (True, True)])
def test_gemm_pair_layouts(pair_kernel, dtype, shape, ta, tb):
    """CTA-pair (cta_group::2) kernel, all four operand layouts incl. ragged edges."""
    M, Nn, Kk = shape
    if (M % 8 and ta) or (Nn % 8):
        pytest.skip("leading dims must be multiples of 8")
    torch.manual_seed(0)
    a = torch.randn((Kk, M) if ta else (M, Kk), device=DEV, dtype=dtype)
    b = torch.randn((Nn, Kk) if tb else (Kk, Nn), device=DEV, dtype=dtype)
    out = K.gemm(a, b, trans_a=ta, trans_b=tb)
    ref = (a.float().t() if ta else a.float()) @ (b.float().t() if tb else b.float())
    assert _rel(out, ref) < (2e-3 if dtype == torch.float16 else 1.5e-2)


@pytest.mark.parametrize("epi", ["bias_gelu", "residual", "dgelu"])
def test_gemm_pair_epilogues(pair_kernel, epi):
    torch.manual_seed(1)
    M, Nn, Kk = 520, 1032, 512
    dt = torch.float16
    a = torch.randn(M, Kk, device=DEV, dtype=dt) * 0.1
    b = torch.randn(Nn, Kk, device=DEV, dtype=dt) * 0.1
    bias = torch.randn(Nn, device=DEV, dtype=dt)
    aux = torch.randn(M, Nn, device=DEV, dtype=dt)
    acc = a.float() @ b.float().t()
    if epi == "bias_gelu":
        pre = torch.empty(M, Nn, device=DEV, dtype=dt)
        out = K.gemm(a, b, epilogue=N.EPI_BIAS_GELU, bias=bias, aux_out=pre)
        ref = K.gelu_tanh(acc + bias.float())
        assert _rel(pre, acc + bias.float()) < 2e-3
    elif epi == "residual":
        out = K.gemm(a, b, epilogue=N.EPI_BIAS_RESIDUAL, bias=bias, aux=aux)
        ref = acc + bias.float() + aux.float()
    else:
        out = K.gemm(a, b, epilogue=N.EPI_DGELU, aux=aux)
        ref = acc * K.gelu_tanh_grad(aux.float())
    assert _rel(out, ref) < 3e-3


def test_gemm_pair_batched(pair_kernel):
    torch.manual_seed(2)
    q = torch.randn(6, 512, 64, device=DEV, dtype=torch.float16)
    k = torch.randn(6, 512, 64, device=DEV, dtype=torch.float16)
    sc = K.gemm(q, k, trans_b=True, alpha=0.125)
    assert _rel(sc, torch.matmul(q.float(), k.float().transpose(1, 2)) * 0.125) < 2e-3


@pytest.mark.parametrize("split", [2, 3, 4])
@pytest.mark.parametrize("shape,ta,tb", [((4096, 2048, 2048), False, True), ((1000, 520, 4096), False, True),
                                         ((2048, 2048, 8192), True, False), ((8192, 2048, 1024), False, False)])
def test_gemm_pair_split_k(monkeypatch, split, shape, ta, tb):
    """Tail-wave K split: contributors park fp32 partials, the owner slice adds them in its epilogue."""
    monkeypatch.setattr(K, "GEMM_KERNEL", "pair")
    lib = N.cuda_lib()
    assert lib.dsb_gemm_pair_force_split(split) == 0
    try:
        M, Nn, Kk = shape
        torch.manual_seed(7)
        dt = torch.float16
        a = torch.randn((Kk, M) if ta else (M, Kk), device=DEV, dtype=dt)
        b = torch.randn((Nn, Kk) if tb else (Kk, Nn), device=DEV, dtype=dt)
        bias = torch.randn(Nn, device=DEV, dtype=dt)
        aux = torch.randn(M, Nn, device=DEV, dtype=dt)
        ref = (a.float().t() if ta else a.float()) @ (b.float().t() if tb else b.float())
        for _ in range(3):   # flags must be back to zero after every launch
            out = K.gemm(a, b, trans_a=ta, trans_b=tb, epilogue=N.EPI_BIAS_RESIDUAL, bias=bias, aux=aux, alpha=0.5)
            assert _rel(out, 0.5 * ref + bias.float() + aux.float()) < 2e-3, _where(out, 0.5 * ref + bias.float() + aux.float(), 2e-3)
    finally:
        lib.dsb_gemm_pair_force_split(-1)


def test_gemm_autotune_dispatch():
    """Native tuner records a variant per problem and dsb_gemm_auto uses it (results stay correct)."""
    from deeperspeed_b200.ops import gemm_autotune as GA
    best, ms = GA.tune(1024, 512, 2048, iters=3)
    assert best in GA.VARIANTS and len(ms) == len(GA.VARIANTS) and min(x for x in ms if x > 0) > 0
    assert GA.tuned_variant(1024, 512, 2048) == best
    assert GA.tuned_variant(1024, 512, 2056) is None
    torch.manual_seed(8)
    a = torch.randn(1024, 2048, device=DEV, dtype=torch.float16)
    b = torch.randn(512, 2048, device=DEV, dtype=torch.float16)
    out = K.gemm(a, b)
    assert _rel(out, a.float() @ b.float().t()) < 2e-3
    rep = GA.tune_module(torch.nn.Linear(256, 384, bias=False).cuda().half(), 512, iters=2)
    assert len(rep) == 3


def test_gemm_batched_alpha():
    torch.manual_seed(2)
    B, S, D = 12, 256, 64
    q = torch.randn(B, S, D, device=DEV, dtype=torch.float16)
    k = torch.randn(B, S, D, device=DEV, dtype=torch.float16)
    v = torch.randn(B, S, D, device=DEV, dtype=torch.float16)
    sc = K.gemm(q, k, trans_b=True, alpha=0.125)
    assert _rel(sc, torch.matmul(q.float(), k.float().transpose(1, 2)) * 0.125) < 2e-3
    ctx = K.gemm(sc, v, trans_b=False)
    assert _rel(ctx, torch.matmul(sc.float(), v.float())) < 2e-3


def test_linear_autograd():
    torch.manual_seed(3)
    x = torch.randn(4, 96, 256, device=DEV, dtype=torch.float16, requires_grad=True)
    w = (torch.randn(512, 256, device=DEV, dtype=torch.float16) * 0.05).requires_grad_()
    b = torch.randn(512, device=DEV, dtype=torch.float16, requires_grad=True)
    y = K.linear_gelu(x, w, b)
    g = torch.randn_like(y)
    y.backward(g)
    xf, wf, bf = (t.detach().float().requires_grad_() for t in (x, w, b))
    yr = K.gelu_tanh(torch.nn.functional.linear(xf, wf, bf))
    yr.backward(g.float())
    assert _rel(y, yr) < 3e-3
    assert _rel(x.grad, xf.grad) < 5e-3
    assert _rel(w.grad, wf.grad) < 5e-3
    assert _rel(b.grad, bf.grad) < 5e-3


def test_layernorm_passthrough_and_grad_accumulate():
    """Residual stream routed through the LN op (alias output) + gamma/beta gradients accumulated in place."""
    torch.manual_seed(6)
    rows, hidden = 130, 2048
    x = torch.randn(rows, hidden, device=DEV, dtype=torch.float16, requires_grad=True)
    g = (1 + 0.1 * torch.randn(hidden, device=DEV, dtype=torch.float16)).requires_grad_()
    b = (0.1 * torch.randn(hidden, device=DEV, dtype=torch.float16)).requires_grad_()
    g.grad = torch.ones_like(g)
    b.grad = torch.full_like(b, 2.0)
    gbuf, bbuf = g.grad, b.grad
    y, xr = K.layer_norm(x, g, b, 1e-5)
    assert xr.data_ptr() == x.data_ptr()
    gy, gs = torch.randn_like(y), torch.randn_like(y)
    ((y * gy).sum() + (xr * gs).sum()).backward()
    xf, gf, bf = (t.detach().float().requires_grad_() for t in (x, g, b))
    yf = torch.nn.functional.layer_norm(xf, (hidden,), gf, bf, 1e-5)
    ((yf * gy.float()).sum() + (xf * gs.float()).sum()).backward()
    assert _rel(y, yf) < 4e-3 and _rel(x.grad, xf.grad) < 4e-3
    assert g.grad is gbuf and b.grad is bbuf          # accumulated in place, no new tensors
    assert _rel(g.grad, gf.grad + 1.0) < 4e-3 and _rel(b.grad, bf.grad + 2.0) < 4e-3
    # only the pass-through used
    x2 = x.detach().clone().requires_grad_()
    _, xr2 = K.layer_norm(x2, g, b, 1e-5)
    (xr2 * gs).sum().backward()
    assert torch.equal(x2.grad, gs)


def test_mlp_residual_autograd():
    """Fused MLP block (gelu' in the FF2 dgrad epilogue) against fp32 autograd."""
    torch.manual_seed(5)
    x = torch.randn(3, 128, 256, device=DEV, dtype=torch.float16, requires_grad=True)
    r = torch.randn(3, 128, 256, device=DEV, dtype=torch.float16, requires_grad=True)
    w1 = (torch.randn(1024, 256, device=DEV, dtype=torch.float16) * 0.05).requires_grad_()
    b1 = (torch.randn(1024, device=DEV, dtype=torch.float16) * 0.1).requires_grad_()
    w2 = (torch.randn(256, 1024, device=DEV, dtype=torch.float16) * 0.05).requires_grad_()
    b2 = (torch.randn(256, device=DEV, dtype=torch.float16) * 0.1).requires_grad_()
    y = K.mlp_residual(x, w1, b1, w2, b2, r)
    g = torch.randn_like(y)
    y.backward(g)
    ref_in = [t.detach().float().requires_grad_() for t in (x, w1, b1, w2, b2, r)]
    xf, w1f, b1f, w2f, b2f, rf = ref_in
    yr = torch.nn.functional.linear(K.gelu_tanh(torch.nn.functional.linear(xf, w1f, b1f)), w2f, b2f) + rf
    yr.backward(g.float())
    assert _rel(y, yr) < 3e-3
    for t, tf in zip((x, w1, b1, w2, b2, r), ref_in):
        assert _rel(t.grad, tf.grad) < 6e-3


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("hidden", [160, 1024, 2048, 4096, 8192])
def test_layernorm(dtype, hidden):
    torch.manual_seed(4)
    rows = 333
    x = torch.randn(rows, hidden, device=DEV, dtype=dtype, requires_grad=True)
    r = torch.randn(rows, hidden, device=DEV, dtype=dtype, requires_grad=True)
    g = (1 + 0.1 * torch.randn(hidden, device=DEV, dtype=dtype)).requires_grad_()
    b = (0.1 * torch.randn(hidden, device=DEV, dtype=dtype)).requires_grad_()
    y, s = K.layer_norm(x, g, b, 1e-5, residual=r)
    gy, gs = torch.randn_like(y), torch.randn_like(s)
    (y * gy).sum().backward(retain_graph=True)
    (s * gs).sum().backward()
    xf, rf, gf, bf = (t.detach().float().requires_grad_() for t in (x, r, g, b))
    sf = (xf + rf)
    yf = torch.nn.functional.layer_norm(sf, (hidden,), gf, bf, 1e-5)
    ((yf * gy.float()).sum() + (sf * gs.float()).sum()).backward()
    tol = 4e-3 if dtype == torch.float16 else 3e-2
    assert _rel(y, yf) < tol and _rel(s, sf) < tol
    assert _rel(x.grad, xf.grad) < tol and _rel(r.grad, rf.grad) < tol
    assert _rel(g.grad, gf.grad) < tol and _rel(b.grad, bf.grad) < tol




def _softmax_ref(x, mask, scale, causal):
    """softmax(scale * x + key-padding mask [+ causal mask]) in fp32 with plain torch ops."""
    z = x.float() * scale
    if mask is not None:
        z = z + mask.float()[:, None, None, :]
    if causal:
        Sq, Sk = z.shape[-2], z.shape[-1]
        keep = torch.ones(Sq, Sk, dtype=torch.bool, device=z.device).tril(Sk - Sq)
        z = z.masked_fill(~keep, float("-inf"))
    return torch.softmax(z, -1)


def _softmax_bwd_ref(dy, y, scale):
    yf, dyf = y.float(), dy.float()
    return scale * yf * (dyf - (dyf * yf).sum(-1, keepdim=True))

@pytest.mark.parametrize("sk", [64, 128, 512, 2048, 8192])
@pytest.mark.parametrize("causal", [False, True])
def test_softmax(sk, causal):
    torch.manual_seed(5)
    B, H, Sq = 2, 3, min(sk, 128)
    x = torch.randn(B, H, Sq, sk, device=DEV, dtype=torch.float16)
    mask = torch.zeros(B, sk, device=DEV, dtype=torch.float16)
    mask[:, -5:] = -10000.0
    y = K.softmax_fwd(x, mask, scale=0.3, causal=causal)
    assert _rel(y, _softmax_ref(x, mask, 0.3, causal)) < 2e-3
    dy = torch.randn_like(y)
    dx = K.softmax_bwd(dy, y, scale=0.3)
    assert _rel(dx, _softmax_bwd_ref(dy, y, 0.3)) < 3e-3


@pytest.mark.parametrize("sk", [72, 256, 1024, 1536])
def test_softmax_group_kernels_ragged_rows(sk):
    """Short-row kernels (several rows per warp / CTA) with a row count that does not fill the last CTA."""
    torch.manual_seed(6)
    x = torch.randn(1, 3, 5, sk, device=DEV, dtype=torch.float16)
    y = K.softmax_fwd(x, None, scale=0.7, causal=True)
    assert _rel(y, _softmax_ref(x, None, 0.7, True)) < 2e-3
    dy = torch.randn_like(y)
    assert _rel(K.softmax_bwd(dy, y, scale=0.7), _softmax_bwd_ref(dy, y, 0.7)) < 3e-3


def test_bias_gelu_dropout_transforms():
    torch.manual_seed(6)
    x = torch.randn(300, 1024, device=DEV, dtype=torch.float16)
    bias = torch.randn(1024, device=DEV, dtype=torch.float16)
    assert _rel(K.bias_gelu_fwd(x, bias), K.gelu_tanh(x.float() + bias.float())) < 2e-3
    dy = torch.randn_like(x)
    assert _rel(K.bias_gelu_bwd(dy, x, bias), dy.float() * K.gelu_tanh_grad(x.float() + bias.float())) < 3e-3
    res = torch.randn_like(x)
    out, mask = K.dropout_fwd(x, 0.25, seed=1234, offset=7, bias=bias, residual=res)
    keep = mask.float()
    assert abs(keep.mean().item() - 0.75) < 0.01
    ref = (x.float() + bias.float()) * keep / 0.75 + res.float()
    assert _rel(out, ref) < 2e-3
    out2, mask2 = K.dropout_fwd(x, 0.25, seed=1234, offset=7, bias=bias, residual=res)
    assert torch.equal(mask, mask2)
    assert _rel(K.dropout_bwd(dy, mask, 0.25), dy.float() * keep / 0.75) < 2e-3
    B, S, H, D = 2, 50, 4, 64
    qkv = torch.randn(B, S, 3 * H * D, device=DEV, dtype=torch.float16)
    qb = torch.randn(3 * H * D, device=DEV, dtype=torch.float16)
    q, k, v = K.qkv_bias_split_transpose(qkv, qb, B, S, H, D)
    # explicit index math: [B, S, 3, H, D] + bias -> three [B, H, S, D] tensors
    t = (qkv.float() + qb.float()).view(B, S, 3, H, D).permute(2, 0, 3, 1, 4)
    assert q.shape == (B, H, S, D)
    assert _rel(q, t[0]) < 1e-3 and _rel(k, t[1]) < 1e-3 and _rel(v, t[2]) < 1e-3
    assert torch.equal(K.transpose_0213(q), q.permute(0, 2, 1, 3).contiguous())
    assert _rel(K.column_sum(x), x.float().sum(0)) < 3e-3
    assert _rel(K.fused_add(x, res, dy), x.float() + res.float() + dy.float()) < 2e-3


@pytest.mark.parametrize("vocab", [1000, 50432, 50257])
def test_cross_entropy(vocab):
    torch.manual_seed(7)
    T = 257
    ld = (vocab + 7) // 8 * 8
    buf = torch.randn(T, ld, device=DEV, dtype=torch.float16) * 2
    logits = buf[:, :vocab].detach().requires_grad_()
    tg = torch.randint(0, vocab, (T,), device=DEV)
    tg[5] = -100
    ref_in = logits.detach().float().requires_grad_()
    ref = torch.nn.functional.cross_entropy(ref_in, tg, ignore_index=-100)
    (ref * 3.0).backward()
    loss = K.cross_entropy(logits, tg)
    (loss * 3.0).backward()
    assert abs(loss.item() - ref.item()) < 2e-3 * abs(ref.item())
    assert _rel(logits.grad, ref_in.grad) < 5e-3


def test_rotary():
    torch.manual_seed(8)
    B, S, H, D = 2, 128, 4, 128
    qkv = torch.randn(B, S, 3, H, D, device=DEV, dtype=torch.float16, requires_grad=True)
    q, k = K.rotary(qkv[:, :, 0], qkv[:, :, 1], rot_dim=32)
    qr, kr = K._rotary_ref(qkv[:, :, 0].detach().cpu(), qkv[:, :, 1].detach().cpu(), 32, 10000.0, False)
    assert _rel(q, qr.to(DEV)) < 2e-3 and _rel(k, kr.to(DEV)) < 2e-3
    # gradient against autograd through the fp32 reference rotation, with non-trivial upstream gradients
    gq, gk = torch.randn_like(q), torch.randn_like(k)
    (q.float() * gq.float()).sum().backward(retain_graph=True)
    (k.float() * gk.float()).sum().backward()
    ref_in = qkv.detach().float().cpu().requires_grad_()
    rq, rk = K._rotary_ref(ref_in[:, :, 0], ref_in[:, :, 1], 32, 10000.0, False)
    ((rq * gq.float().cpu()).sum() + (rk * gk.float().cpu()).sum()).backward()
    assert _rel(qkv.grad[:, :, :2], ref_in.grad[:, :, :2].to(DEV)) < 3e-3
    assert qkv.grad[:, :, 2].abs().max().item() == 0.0


@pytest.mark.parametrize("gdt", [torch.float16, torch.bfloat16, torch.float32])
def test_adam_flat_and_sumsq(gdt):
    torch.manual_seed(9)
    n = 1_000_003
    p = torch.randn(n, device=DEV)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    g = (torch.randn(n, device=DEV) * 4).to(gdt)
    lp = torch.empty(n, device=DEV, dtype=torch.float16)
    pr, mr, vr = p.cpu().clone(), m.cpu().clone(), v.cpu().clone()
    for step in (1, 2):
        K.adam_flat(p, m, v, g, lp=lp, lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=step,
                    inv_scale=0.25)
        K.adam_flat(pr, mr, vr, g.cpu(), lr=1e-2, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=step,
                    inv_scale=0.25)
    assert _rel(p, pr.to(DEV)) < 1e-5 and _rel(lp, pr.to(DEV)) < 1e-3
    skip = torch.ones(1, dtype=torch.int32, device=DEV)
    before = p.clone()
    K.adam_flat(p, m, v, g, lr=1.0, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=3, skip_dev=skip)
    assert torch.equal(p, before)
    flag = torch.zeros(1, dtype=torch.int32, device=DEV)
    s = K.sumsq(g, flag=flag)
    assert abs(s.item() - (g.float() ** 2).sum().item()) < 1e-3 * s.item() and flag.item() == 0
    g[12345] = float("inf")
    K.sumsq(g, flag=flag)
    assert flag.item() == 1


def test_wgrad_accumulates_into_existing_grad():
    torch.manual_seed(11)
    x = torch.randn(64, 256, device=DEV, dtype=torch.float16)
    w = (torch.randn(512, 256, device=DEV, dtype=torch.float16) * 0.05).requires_grad_()
    b = torch.zeros(512, device=DEV, dtype=torch.float16, requires_grad=True)
    w.grad = torch.ones_like(w)
    b.grad = torch.ones_like(b)
    gptr = w.grad.data_ptr()
    y = K.linear(x, w, b)
    g = torch.randn_like(y)
    y.backward(g)
    assert w.grad.data_ptr() == gptr
    assert _rel(w.grad, 1.0 + g.float().t() @ x.float()) < 3e-3
    assert _rel(b.grad, 1.0 + g.float().sum(0)) < 3e-3


def test_causal_self_attention_op():
    from deeperspeed_b200.ops.attention import causal_self_attention
    torch.manual_seed(12)
    B, S, H, D = 2, 256, 4, 64
    qkv = (torch.randn(B, S, 3 * H * D, device=DEV, dtype=torch.float16) * 0.5).requires_grad_()
    out = causal_self_attention(qkv, H, D, rot_dim=D)
    g = torch.randn_like(out)
    out.backward(g)
    ref_in = qkv.detach().float().cpu().requires_grad_()
    t = ref_in.view(B, S, 3, H, D)
    q, k = K._rotary_ref(t[:, :, 0], t[:, :, 1], D, 10000.0, False)
    ctx = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), t[:, :, 2].transpose(1, 2),
                                                           is_causal=True)
    ref = ctx.transpose(1, 2).reshape(B, S, H * D)
    ref.backward(g.float().cpu())
    assert _rel(out.cpu(), ref) < 5e-3
    assert _rel(qkv.grad.cpu(), ref_in.grad) < 1e-2


def test_fused_zero_kernels_single_rank():
    """The fused ZeRO step kernels with R = 1 (this GPU is its own peer): reduce + norm, Adam + push, and the LAMB
    variant's per-tensor trust ratios, against plain fp32 PyTorch. (The multi-GPU behaviour is in
    tests/test_fused_zero_gpu.py; this one also runs where only one GPU is visible.)"""
    from deeperspeed_b200.parallel.fused_zero_single import SingleRankFusedZero
    torch.manual_seed(0)
    sizes = [4096, 8, 100000, 2048 * 64, 40]
    n = (sum(sizes) + 1024 + 7) // 8 * 8
    lr, betas, eps, wd = 1e-2, (0.9, 0.999), 1e-8, 0.01
    for mode in ("adam", "lamb"):
        z = SingleRankFusedZero(n, tensor_sizes=sizes)
        used = sum(sizes)
        w0 = torch.zeros(n, device="cuda")
        w0[:used] = torch.randn(used, device="cuda") * 0.5
        z.master.copy_(w0)
        w, m, v = w0.clone(), torch.zeros_like(w0), torch.zeros_like(w0)
        for step in (1, 2, 3):
            g16 = torch.zeros(n, device="cuda", dtype=torch.float16)
            g16[:used] = (torch.randn(used, device="cuda") * 0.1).half()
            z.grad.copy_(g16)
            z.rs_norm(scale=0.5)
            g = g16.float() * 0.5
            torch.testing.assert_close(z.red, g, rtol=0, atol=0)
            torch.testing.assert_close(z.sumsq[0], (g * g).sum(), rtol=1e-4, atol=0)
            assert z.flag.item() == 0
            m = betas[0] * m + (1 - betas[0]) * g
            v = betas[1] * v + (1 - betas[1]) * g * g
            if mode == "adam":
                z.adam_ag(lr, betas, eps, wd, True, step)
                upd = (m / (1 - betas[0] ** step)) / ((v / (1 - betas[1] ** step)).sqrt() + eps) + wd * w
                w = w - lr * upd
            else:
                z.lamb_ag(lr, betas, eps, wd, 10.0, 0.01, step)
                u = m / (v.sqrt() + eps) + wd * w
                step_size = lr * (1 - betas[1] ** step) ** 0.5 / (1 - betas[0] ** step)
                pos = 0
                for ti, s in enumerate(sizes):
                    wn, un = w[pos:pos + s].norm(), u[pos:pos + s].norm()
                    c = float(min(max(wn / un, 0.01), 10.0)) if wn > 0 and un > 0 else 1.0
                    assert abs(z.coeff[ti].item() - c) < 1e-4 * max(1.0, c), (ti, z.coeff[ti].item(), c)
                    w[pos:pos + s] -= step_size * c * u[pos:pos + s]
                    pos += s
            torch.cuda.synchronize()
            torch.testing.assert_close(z.master, w, rtol=1e-5, atol=1e-6)
            torch.testing.assert_close(z.param.float(), w.half().float(), rtol=0, atol=4e-3)
        # non-finite gradients raise the flag; a set skip flag leaves everything untouched
        z.grad[5] = float("inf")
        z.rs_norm()
        assert z.flag.item() == 1
        z.skip.fill_(1)
        before = z.master.clone()
        (z.adam_ag(lr, betas, eps, wd, True, 4) if mode == "adam" else z.lamb_ag(lr, betas, eps, wd, 10.0, 0.01, 4))
        torch.cuda.synchronize()
        assert torch.equal(before, z.master)


def test_collective_gemm_variants_single_rank():
    """The GEMM kernels that carry a collective (all-gather of a row-sharded weight in forward and dgrad layout,
    reduce-scatter of the result) with ONE rank: the GPU is its own peer, so tile order, panel flags, the k-loop
    rotation, the staging stores, the barriers and the slot sum all run; results against fp32 matmul."""
    import ctypes as C
    from deeperspeed_b200.parallel import ag_gemm  # noqa: F401  (ctypes signatures)
    lib = N.cuda_lib()
    torch.manual_seed(1)
    tokens, out_f, in_f = 384, 512, 264
    w = (torch.randn(out_f, in_f, device="cuda") * 0.1).half()
    x = (torch.randn(tokens, in_f, device="cuda") * 0.1).half()
    go = (torch.randn(tokens, out_f, device="cuda") * 0.1).half()
    flags = torch.zeros(1024, dtype=torch.int32, device="cuda")
    one = lambda t: (C.c_void_p * 1)(t.data_ptr())
    dc, st = N.dtype_code(torch.float16), N.stream_ptr()
    gathered = torch.empty_like(w)
    pflags = torch.zeros(64, dtype=torch.int32, device="cuda")
    # forward: y = x @ gather(w)^T
    y = torch.empty(tokens, out_f, device="cuda", dtype=torch.float16)
    N.check(lib.dsb_gemm_ag_overlap(N.ptr(x), one(w), 1, 0, N.ptr(gathered), N.ptr(pflags), N.ptr(y), tokens, out_f,
                                    in_f, x.stride(0), y.stride(0), dc, 0, 1.0, None, None, 0, st), "ag_overlap")
    ref = x.float() @ w.float().t()
    assert ((y.float() - ref).abs().max() / ref.abs().max()).item() < 3e-3
    # dgrad: gi = go @ gather(w), shards split the reduction dimension
    gi = torch.empty(tokens, in_f, device="cuda", dtype=torch.float16)
    N.check(lib.dsb_gemm_ag_overlap_k(N.ptr(go), one(w), 1, 0, N.ptr(gathered), N.ptr(pflags), N.ptr(gi), tokens, in_f,
                                      out_f, go.stride(0), gi.stride(0), dc, 1.0, st), "ag_overlap_k")
    ref = go.float() @ w.float()
    assert ((gi.float() - ref).abs().max() / ref.abs().max()).item() < 3e-3
    # wgrad + reduce-scatter: dW = go^T @ x, all rows owned by this rank
    stage = torch.zeros(out_f * in_f, device="cuda", dtype=torch.float16)
    sync = torch.zeros(2, dtype=torch.int32, device="cuda")
    dw = torch.empty(out_f, in_f, device="cuda", dtype=torch.float16)
    for epoch in (1, 2):
        N.check(lib.dsb_gemm_reduce_scatter(N.ptr(go), N.ptr(x), N.ptr(dw), one(stage), one(flags), 1, 0, N.ptr(sync),
                                            out_f, in_f, tokens, go.stride(0), x.stride(0), 1, 1, dc, 1.0, 0.5, epoch,
                                            st), "gemm_reduce_scatter")
        ref = 0.5 * (go.float().t() @ x.float())
        torch.cuda.synchronize()
        assert ((dw.float() - ref).abs().max() / ref.abs().max()).item() < 3e-3


@pytest.mark.parametrize("variant", [2, 64, 128])
@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_flash_attention_forward(causal, dtype, variant):
    """tcgen05 flash-attention forward (all three kernels) vs fp32 softmax(QK^T)V, on strided views of a fused QKV
    buffer."""
    from deeperspeed_b200.ops.attention import flash_attn_fwd, set_flash_fwd_variant
    prev = set_flash_fwd_variant(variant)
    try:
        _check_flash_forward(causal, dtype, flash_attn_fwd)
    finally:
        set_flash_fwd_variant(prev)


def _check_flash_forward(causal, dtype, flash_attn_fwd, B=2, S=768, H=3):
    torch.manual_seed(0)
    D = 128
    qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=dtype)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    out, lse = flash_attn_fwd(q, k, v, causal=causal, return_lse=True)
    torch.cuda.synchronize()
    qf, kf, vf = (t.float().transpose(1, 2) for t in (q, k, v))      # [B, H, S, D]
    scores = qf @ kf.transpose(-1, -2) / D ** 0.5
    if causal:
        scores = scores.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device="cuda"), 1), float("-inf"))
    ref = (torch.softmax(scores, -1) @ vf).transpose(1, 2)
    tol = 4e-3 if dtype == torch.float16 else 2e-2
    assert out.shape == (B, S, H, D) and out.is_contiguous()
    assert (out.float() - ref).abs().max().item() < tol, (out.float() - ref).abs().max().item()
    assert (lse - torch.logsumexp(scores, -1)).abs().max().item() < 2e-3


@pytest.mark.parametrize("S", [2048, 8192])
def test_flash_attention_long_sequences(S):
    """The sequence lengths the benchmarks run (several waves of CTAs per head, 32 / 128 key blocks per row block)."""
    from deeperspeed_b200.ops.attention import flash_attn_fwd
    _check_flash_forward(True, torch.float16, flash_attn_fwd, B=1, S=S, H=2)


def test_flash_attention_forward_odd_block_count_falls_back():
    """S % 256 == 128: the two-tile kernel does not apply, the 64-key kernel runs instead."""
    from deeperspeed_b200.ops.attention import flash_attn_fwd
    _check_flash_forward(True, torch.float16, flash_attn_fwd, B=1, S=640, H=2)


@pytest.mark.parametrize("gen,S", [(2, 512), (2, 2048), (1, 512)])
@pytest.mark.parametrize("causal", [True, False])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_flash_attention_backward(causal, dtype, gen, S):
    """tcgen05 flash-attention backward (both generations) vs autograd through fp32 softmax(QK^T)V."""
    from deeperspeed_b200.ops.attention import flash_attn_fwd, flash_attn_bwd, set_flash_bwd_variant
    prev = set_flash_bwd_variant(gen)
    try:
        _check_flash_backward(causal, dtype, S, flash_attn_fwd, flash_attn_bwd)
    finally:
        set_flash_bwd_variant(prev)


def _check_flash_backward(causal, dtype, S, flash_attn_fwd, flash_attn_bwd):
    torch.manual_seed(1)
    B, H, D = (2, 3, 128) if S <= 512 else (1, 2, 128)
    qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=dtype)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    out, lse = flash_attn_fwd(q, k, v, causal=causal, return_lse=True)
    dout = torch.randn(B, S, H, D, device="cuda", dtype=dtype)
    dqkv = torch.empty_like(qkv)
    dq, dk, dv = flash_attn_bwd(dout, q, k, v, out, lse, causal=causal, dk=dqkv[:, :, 1], dv=dqkv[:, :, 2])
    torch.cuda.synchronize()
    qf, kf, vf = (t.float().detach().requires_grad_() for t in (q, k, v))
    scores = qf.transpose(1, 2) @ kf.transpose(1, 2).transpose(-1, -2) / D ** 0.5
    if causal:
        scores = scores.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device="cuda"), 1), float("-inf"))
    ref = (torch.softmax(scores, -1) @ vf.transpose(1, 2)).transpose(1, 2)
    rq, rk, rv = torch.autograd.grad(ref, (qf, kf, vf), dout.float())
    tol = 2e-2 if dtype == torch.float16 else 8e-2
    for name, got, want in (("dq", dq, rq), ("dk", dk.float(), rk), ("dv", dv.float(), rv)):
        err = (got.float() - want).abs().max().item()
        assert err < tol * max(1.0, want.abs().max().item()), (name, err, want.abs().max().item())


@pytest.mark.parametrize("rot", [128, 32, 0])
def test_causal_self_attention_op_native_backend(rot):
    """The attention block op on the tcgen05 flash-attention kernels (rotary in, inverse rotary out, dV written
    straight into the dQKV buffer) against the fp32 reference of the same block."""
    from deeperspeed_b200.ops import attention as A
    prev = A.set_attention_backend("native")
    try:
        torch.manual_seed(13)
        B, S, H, D = 2, 384, 3, 128
        qkv = (torch.randn(B, S, 3 * H * D, device=DEV, dtype=torch.float16) * 0.5).requires_grad_()
        out = A.causal_self_attention(qkv, H, D, rot_dim=rot)
        assert out.grad_fn is not None and "Native" in type(out.grad_fn).__name__
        g = torch.randn_like(out)
        out.backward(g)
    finally:
        A.set_attention_backend(prev)
    ref_in = qkv.detach().float().cpu().requires_grad_()
    t = ref_in.view(B, S, 3, H, D)
    q, k = t[:, :, 0], t[:, :, 1]
    if rot > 0:
        q, k = K._rotary_ref(q, k, rot, 10000.0, False)
    ctx = torch.nn.functional.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), t[:, :, 2].transpose(1, 2),
                                                           is_causal=True)
    ref = ctx.transpose(1, 2).reshape(B, S, H * D)
    ref.backward(g.float().cpu())
    assert _rel(out.cpu(), ref) < 5e-3
    assert _rel(qkv.grad.cpu(), ref_in.grad) < 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_embedding_gather_and_scatter_add(dtype):
    """K.embedding: gather forward; backward adds into the table's EXISTING .grad (arena view) with packed atomics,
    or returns a fresh gradient when there is none. Oracle: torch.nn.functional.embedding in fp32, repeated ids."""
    torch.manual_seed(0)
    vocab, dim = 1000, 264
    table = (torch.randn(vocab, dim, device="cuda") * 0.5).to(dtype).requires_grad_()
    ids = torch.randint(0, 50, (4, 96), device="cuda")          # few distinct rows: many repeats per row
    dout = (torch.randn(4, 96, dim, device="cuda") * 0.1).to(dtype)
    ref_t = table.detach().float().requires_grad_()
    ref = torch.nn.functional.embedding(ids, ref_t)
    ref.backward(dout.float())
    out = K.embedding(ids, table)
    assert torch.equal(out, table.detach()[ids])
    out.backward(dout)                                           # no .grad yet: autograd receives the gradient
    assert (table.grad.float() - ref_t.grad).abs().max().item() < (2e-2 if dtype == torch.float16 else 1e-1)
    # with a gradient buffer in place the kernel accumulates into it and autograd gets None
    base = table.grad.detach().clone()
    keep_ptr = table.grad.data_ptr()
    K.embedding(ids, table).backward(dout)
    assert table.grad.data_ptr() == keep_ptr
    assert (table.grad.float() - 2 * base.float()).abs().max().item() < (4e-2 if dtype == torch.float16 else 2e-1)
    # untouched rows stay exactly zero
    assert float(table.grad[50:].abs().max()) == 0.0


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_fused_zero_one_kernel_step_single_rank(dtype):
    """dsb_fused_zero_step with R = 1 (this GPU is its own peer): reduce + norm, in-kernel exchange (unscale / clip
    factor, overflow skip), AdamW on the fp32 master, cast + bulk-copy gather, clearing of the gradient bucket - against
    plain fp32 PyTorch. The multi-GPU behaviour is in tests/test_fused_zero_gpu.py / test_xgpu_protocol_gpu.py."""
    from deeperspeed_b200.parallel.fused_zero_single import SingleRankFusedZero
    torch.manual_seed(1)
    n = 8 * (512 * 37 + 5)              # not a multiple of the 512-vector CTA tile: ragged tails in both phases
    z = SingleRankFusedZero(n, dtype=dtype)
    lr, betas, eps, wd, scale, clip = 1e-2, (0.9, 0.95), 1e-8, 0.01, 64.0, 0.5
    w = torch.randn(n, device="cuda") * 0.5
    z.master.copy_(w)
    m, v = torch.zeros_like(w), torch.zeros_like(w)
    for step in (1, 2, 3):
        g16 = (torch.randn(n, device="cuda") * 0.05 * scale).to(dtype)
        z.grad.copy_(g16)
        out = z.whole_step(lr, betas, eps, wd, True, step, loss_scale=scale, clip=clip)
        g = g16.float()
        norm = g.norm()
        combined = scale * max(1.0, (float(norm) / scale + 1e-6) / clip)
        gu = g / combined
        m = betas[0] * m + (1 - betas[0]) * gu
        v = betas[1] * v + (1 - betas[1]) * gu * gu
        w = w - lr * ((m / (1 - betas[0] ** step)) / ((v / (1 - betas[1] ** step)).sqrt() + eps) + wd * w)
        torch.cuda.synchronize()
        assert abs(out[0].item() - float(norm)) < 1e-3 * float(norm) and out[1].item() == 0.0
        torch.testing.assert_close(z.master, w, rtol=2e-5, atol=2e-6)
        torch.testing.assert_close(z.param.float(), w.to(dtype).float(), rtol=0, atol=1e-2)
        assert float(z.grad.float().abs().max()) == 0.0          # the kernel cleared the gradient bucket
    # overflow: nothing is updated, the flag reaches the host slot, the gradients are still cleared
    z.grad.copy_((torch.randn(n, device="cuda") * 0.05).to(dtype))
    z.grad[11] = float("inf")
    before = z.master.clone()
    out = z.whole_step(lr, betas, eps, wd, True, 4, loss_scale=scale, clip=clip)
    torch.cuda.synchronize()
    assert out[1].item() == 1.0 and torch.equal(before, z.master) and float(z.grad.float().abs().max()) == 0.0

