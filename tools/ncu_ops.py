"""Driver for an ncu capture of the bandwidth-bound kernels (one launch each inside the profiler range):
LayerNorm fwd / fwd+residual / bwd(+dres), bias-GeLU bwd, column sum, fused add, softmax fwd/bwd (seq 128 and 2048),
Adam (64M elements), sumsq, cross entropy fwd/bwd, rotary. Run plainly first, then under
`ncu --set full --clock-control none --import-source on --profile-from-start off`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.ops import native as N
from deeperspeed_b200.ops.attention import rotary_oop

dev, dt = "cuda", torch.float16
ROWS, H = 8192, 2048
lib = N.cuda_lib()
x = torch.randn(ROWS, H, device=dev, dtype=dt); r = torch.randn_like(x); y = torch.empty_like(x); s = torch.empty_like(x)
gamma = torch.ones(H, device=dev, dtype=dt); beta = torch.zeros(H, device=dev, dtype=dt)
mean = torch.empty(ROWS, device=dev); rstd = torch.empty(ROWS, device=dev)
dg = torch.empty(H, device=dev, dtype=dt); db = torch.empty(H, device=dev, dtype=dt)
sc = N.scratch(lib.dsb_layernorm_bwd_scratch_bytes(H), x.device, "ln_bwd")
big = torch.randn(ROWS, 4 * H, device=dev, dtype=dt); big2 = torch.randn_like(big)
sm128 = torch.randn(64, 16, 128, 128, device=dev, dtype=dt); sm2k = torch.randn(4, 16, 2048, 2048, device=dev, dtype=dt)
n = 64 << 20
p = torch.randn(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
g = torch.randn(n, device=dev, dtype=dt); lp = torch.empty(n, device=dev, dtype=dt)
logits = torch.randn(ROWS, 50432, device=dev, dtype=dt, requires_grad=True); tgt = torch.randint(0, 50432, (ROWS,), device=dev)
qkv = torch.randn(4, 2048, 3, 16, 128, device=dev, dtype=dt); q = torch.empty(4, 2048, 16, 128, device=dev, dtype=dt); k = torch.empty_like(q)


def once():
    dc = N.dtype_code(dt)
    st = N.stream_ptr()
    lib.dsb_layernorm_fwd(N.ptr(x), 0, N.ptr(gamma), N.ptr(beta), N.ptr(y), 0, N.ptr(mean), N.ptr(rstd), ROWS, H, 1e-5, dc, st)
    lib.dsb_layernorm_fwd(N.ptr(x), N.ptr(r), N.ptr(gamma), N.ptr(beta), N.ptr(y), N.ptr(s), N.ptr(mean), N.ptr(rstd), ROWS, H, 1e-5, dc, st)
    lib.dsb_layernorm_bwd(N.ptr(r), N.ptr(x), N.ptr(mean), N.ptr(rstd), N.ptr(gamma), N.ptr(s), N.ptr(y), N.ptr(dg), N.ptr(db), ROWS, H, dc, N.ptr(sc), st)
    K.bias_gelu_bwd(big, big2, None)
    K.column_sum(big)
    K.fused_add(x, r, y, out=s)
    y1 = K.softmax_fwd(sm128, scale=0.125); K.softmax_bwd(sm128, y1, scale=0.125)
    y2 = K.softmax_fwd(sm2k, scale=0.125, causal=True); K.softmax_bwd(sm2k, y2, scale=0.125)
    K.adam_flat(p, m, v, g, lp=lp, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.01, step=3)
    K.sumsq(g)
    logits.grad = None
    K.cross_entropy(logits, tgt).backward()
    rotary_oop(qkv[:, :, 0], qkv[:, :, 1], q, k, 32, 10000.0)


once(); once()
torch.cuda.synchronize()
torch.cuda.profiler.start()
once()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
