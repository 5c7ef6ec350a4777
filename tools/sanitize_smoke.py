"""Small-shape pass over every sm_100a kernel family, meant to run under compute-sanitizer:

    compute-sanitizer --tool memcheck   python tools/sanitize_smoke.py
    compute-sanitizer --tool racecheck  python tools/sanitize_smoke.py      (shared-memory hazards)
    compute-sanitizer --tool synccheck  python tools/sanitize_smoke.py

(the reference has no sanitizer coverage at all). Also checks the results, so a plain run is a quick
numerics smoke test (GEMMs, transformer kernels, flash attention, fused ZeRO kernels). On the development GPU pool of this repository compute-sanitizer is administratively closed, so
only the plain run has been exercised there (profiles/README.md)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.ops import native as N
from deeperspeed_b200.ops.attention import causal_self_attention

dev = "cuda"
torch.manual_seed(0)


def rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-6)).item()


def check(name, a, b, tol):
    r = rel(a, b)
    print(f"{name:34s} rel err {r:.2e}", flush=True)
    assert r < tol, (name, r)


lib = N.cuda_lib()
dt = torch.float16
# ---- GEMM: single-CTA (ragged edges, all layouts), CTA pair, pair + split-K, epilogues
for (M, Nn, Kk, ta, tb) in [(136, 264, 200, False, True), (128, 256, 64, True, False), (200, 128, 72, False, False)]:
    a = torch.randn((Kk, M) if ta else (M, Kk), device=dev, dtype=dt)
    b = torch.randn((Nn, Kk) if tb else (Kk, Nn), device=dev, dtype=dt)
    K.GEMM_KERNEL = "single"
    ref = (a.float().t() if ta else a.float()) @ (b.float().t() if tb else b.float())
    check(f"gemm single {M}x{Nn}x{Kk} ta={int(ta)} tb={int(tb)}", K.gemm(a, b, trans_a=ta, trans_b=tb), ref, 3e-3)
a = torch.randn(520, 448, device=dev, dtype=dt) * 0.1
b = torch.randn(776, 448, device=dev, dtype=dt) * 0.1
bias = torch.randn(776, device=dev, dtype=dt)
aux = torch.randn(520, 776, device=dev, dtype=dt)
acc = a.float() @ b.float().t()
K.GEMM_KERNEL = "pair"
check("gemm pair bias+residual", K.gemm(a, b, epilogue=N.EPI_BIAS_RESIDUAL, bias=bias, aux=aux), acc + bias.float() + aux.float(), 3e-3)
pre = torch.empty(520, 776, device=dev, dtype=dt)
check("gemm pair bias+gelu", K.gemm(a, b, epilogue=N.EPI_BIAS_GELU, bias=bias, aux_out=pre), K.gelu_tanh(acc + bias.float()), 3e-3)
lib.dsb_gemm_pair_force_split(2)
a2 = torch.randn(512, 1024, device=dev, dtype=dt) * 0.1
b2 = torch.randn(512, 1024, device=dev, dtype=dt) * 0.1
check("gemm pair split-K 2", K.gemm(a2, b2), a2.float() @ b2.float().t(), 3e-3)
lib.dsb_gemm_pair_force_split(-1)
K.GEMM_KERNEL = "auto"

# ---- LayerNorm fwd/bwd (+residual, +dres), softmax, bias-gelu, dropout, transforms
x = torch.randn(70, 2048, device=dev, dtype=dt, requires_grad=True)
r = torch.randn(70, 2048, device=dev, dtype=dt, requires_grad=True)
g = torch.ones(2048, device=dev, dtype=dt, requires_grad=True)
be = torch.zeros(2048, device=dev, dtype=dt, requires_grad=True)
y, s = K.layer_norm(x, g, be, 1e-5, residual=r)
(y.float().square().sum() + s.float().sum()).backward()
sf = (x.detach().float() + r.detach().float())
check("layernorm fwd", y, torch.nn.functional.layer_norm(sf, (2048,), g.detach().float(), be.detach().float(), 1e-5), 5e-3)
sc = torch.randn(2, 4, 64, 64, device=dev, dtype=dt)
check("softmax causal", K.softmax_fwd(sc.clone(), scale=0.5, causal=True),
      torch.softmax((sc.float() * 0.5).masked_fill(torch.triu(torch.ones(64, 64, device=dev, dtype=torch.bool), 1), float("-inf")), -1), 5e-3)
h = torch.randn(64, 512, device=dev, dtype=dt)
hb = torch.randn(512, device=dev, dtype=dt)
check("bias gelu", K.bias_gelu_fwd(h, hb), K.gelu_tanh(h.float() + hb.float()), 5e-3)
out, mask = K.dropout_fwd(h, 0.1, seed=3, offset=0)
assert abs(mask.float().mean().item() - 0.9) < 0.05
check("column sum", K.column_sum(h), h.float().sum(0), 5e-3)
check("fused add", K.fused_add(h, h, h), 3 * h.float(), 5e-3)

# ---- cross entropy, rotary + attention op, optimizer kernels
logits = torch.randn(96, 1000, device=dev, dtype=dt, requires_grad=True)
tgt = torch.randint(0, 1000, (96,), device=dev)
ce_ref = torch.nn.functional.cross_entropy(logits.detach().float(), tgt)   # before backward: it reuses the logits buffer
loss = K.cross_entropy(logits, tgt)
loss.backward()
check("cross entropy", loss, ce_ref, 5e-3)
qkv = torch.randn(2, 128, 3 * 4 * 64, device=dev, dtype=dt, requires_grad=True)
ctx = causal_self_attention(qkv, 4, 64, 32)
ctx.float().sum().backward()
assert torch.isfinite(qkv.grad).all()
n = 1 << 16
p = torch.randn(n, device=dev)
m = torch.zeros(n, device=dev)
v = torch.zeros(n, device=dev)
gr = torch.randn(n, device=dev, dtype=dt)
lp = torch.empty(n, device=dev, dtype=dt)
pr = p.clone()
K.adam_flat(p, m, v, gr, lp=lp, lr=1e-2, beta1=0.9, beta2=0.99, eps=1e-8, weight_decay=0.0, step=1)
check("adam flat (first step = lr*sign)", p, pr - 1e-2 * torch.sign(gr.float()), 1e-3)
check("sumsq", K.sumsq(gr), gr.float().square().sum(), 1e-3)

# ---- flash attention (tcgen05), both forward block shapes + backward, against fp32 attention
from deeperspeed_b200.ops import attention as A
Bq, Sq, Hq, Dq = 1, 256, 2, 128
qkv5 = torch.randn(Bq, Sq, 3, Hq, Dq, device=dev, dtype=dt)
q_, k_, v_ = qkv5[:, :, 0], qkv5[:, :, 1], qkv5[:, :, 2]
qf, kf, vf = (t.float().transpose(1, 2).detach().requires_grad_() for t in (q_, k_, v_))
sc = (qf @ kf.transpose(-1, -2) / Dq ** 0.5).masked_fill(torch.triu(torch.ones(Sq, Sq, dtype=torch.bool, device=dev), 1), float("-inf"))
ref_o = (torch.softmax(sc, -1) @ vf).transpose(1, 2)
for variant in (128, 64, 2):      # 2 = the two-query-tile kernel with P in tensor memory (the default)
    A.set_flash_fwd_variant(variant)
    o_, lse_ = A.flash_attn_fwd(q_, k_, v_, causal=True, return_lse=True)
    check(f"flash attention fwd ({variant}-key blocks)", o_, ref_o, 5e-3)
do_ = torch.randn_like(o_)
dq_, dk_, dv_ = A.flash_attn_bwd(do_, q_, k_, v_, o_, lse_, causal=True)
rq, rk, rv = torch.autograd.grad(ref_o, (qf, kf, vf), do_.float())
check("flash attention bwd dq", dq_, rq.transpose(1, 2), 2e-2)
check("flash attention bwd dk", dk_, rk.transpose(1, 2), 2e-2)
check("flash attention bwd dv", dv_, rv.transpose(1, 2), 2e-2)

# ---- kernels that carry a collective, with this GPU as its own peer (barriers, flags, staging all exercised)
from deeperspeed_b200.parallel.fused_zero_single import SingleRankFusedZero
z = SingleRankFusedZero(1 << 16, tensor_sizes=[4096, 8192, 1 << 15])
z.master.normal_(0, 0.1)
z.grad.normal_(0, 0.01)
z.rs_norm()
check("fused zero: reduce + norm", z.sumsq[0], z.grad.float().square().sum(), 1e-3)
w_before = z.master.clone()
z.adam_ag(1e-2, (0.9, 0.99), 1e-8, 0.0, True, 1)
check("fused zero: adam + push", z.param, w_before - 1e-2 * torch.sign(z.grad.float()), 2e-3)
z.rs_norm()
z.lamb_ag(1e-2, (0.9, 0.99), 1e-8, 0.0, 10.0, 0.01, 2)
assert torch.isfinite(z.master).all() and (z.coeff[:3] > 0).all()
torch.cuda.synchronize()
print("sanitize_smoke: ok")
