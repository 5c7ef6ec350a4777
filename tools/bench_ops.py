"""Roofline check of the bandwidth-bound kernels on the GPT-NeoX-1.3B shapes (rows = 4 x 2048 tokens, hidden 2048).

Each op is timed with CUDA events over inputs that rotate through > 126 MB of distinct buffers (so nothing is served
from L2), and reported as achieved GB/s against the measured copy bandwidth in MEASURED_PEAKS.json.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.ops import native as N

dev = "cuda"
dt = torch.float16
ROWS, H = 8192, 2048
NBUF = 6   # 6 x (2-4 tensors of 32 MB) >> L2


def peak_gbs():
    try:
        p = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))
        for k in ("hbm_gbs", "hbm_copy_GBps", "hbm_GBps"):
            if k in p:
                return float(p[k])
        for k, v in p.items():
            if "hbm" in k.lower() and isinstance(v, (int, float)):
                return float(v)
    except Exception:
        pass
    return 6545.0


def timeit(fns, iters=30, warm=6, graph=True):
    """fns: list of closures over different buffers, called round-robin."""
    for i in range(warm):
        fns[i % len(fns)]()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graph:
        # replay the launches from a CUDA graph so the number is the kernels' time, not the Python launch rate
        g = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            with torch.cuda.graph(g, stream=side):
                for i in range(iters):
                    fns[i % len(fns)]()
        torch.cuda.current_stream().wait_stream(side)
        g.replay()
        torch.cuda.synchronize()
        a.record()
        g.replay()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / iters
    a.record()
    for i in range(iters):
        fns[i % len(fns)]()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    peak = peak_gbs()
    res = []

    def report(name, ms, nbytes):
        gbs = nbytes / ms / 1e6
        res.append(dict(op=name, ms=round(ms, 4), GBps=round(gbs, 1), pct_of_copy_peak=round(100 * gbs / peak, 1)))
        print(res[-1], flush=True)

    act = ROWS * H * 2
    xs = [torch.randn(ROWS, H, device=dev, dtype=dt) for _ in range(NBUF)]
    rs = [torch.randn(ROWS, H, device=dev, dtype=dt) for _ in range(NBUF)]
    gamma = torch.ones(H, device=dev, dtype=dt)
    beta = torch.zeros(H, device=dev, dtype=dt)
    lib = N.cuda_lib()

    # ---- LayerNorm fwd (plain: read x, write y; with residual: read x, r, write sum, y)
    ys = [torch.empty_like(x) for x in xs]
    ss = [torch.empty_like(x) for x in xs]
    mean = torch.empty(ROWS, device=dev, dtype=torch.float32)
    rstd = torch.empty(ROWS, device=dev, dtype=torch.float32)

    def ln_fwd(i, res_):
        N.check(lib.dsb_layernorm_fwd(N.ptr(xs[i]), N.ptr(rs[i]) if res_ else 0, N.ptr(gamma), N.ptr(beta), N.ptr(ys[i]),
                                      N.ptr(ss[i]) if res_ else 0, N.ptr(mean), N.ptr(rstd), ROWS, H, 1e-5,
                                      N.dtype_code(dt), N.stream_ptr()), "ln_fwd")
    report("layernorm_fwd", timeit([lambda i=i: ln_fwd(i, False) for i in range(NBUF)]), 2 * act)
    report("layernorm_fwd+residual", timeit([lambda i=i: ln_fwd(i, True) for i in range(NBUF)]), 4 * act)

    # ---- LayerNorm bwd (read dy, x, write dx; + dres read)
    sc = N.scratch(lib.dsb_layernorm_bwd_scratch_bytes(H), xs[0].device, "ln_bwd")
    dg, db = torch.empty(H, device=dev, dtype=dt), torch.empty(H, device=dev, dtype=dt)

    def ln_bwd(i, dres):
        N.check(lib.dsb_layernorm_bwd(N.ptr(rs[i]), N.ptr(xs[i]), N.ptr(mean), N.ptr(rstd), N.ptr(gamma),
                                      N.ptr(ss[i]) if dres else 0, N.ptr(ys[i]), N.ptr(dg), N.ptr(db), ROWS, H,
                                      N.dtype_code(dt), N.ptr(sc), N.stream_ptr()), "ln_bwd")
    report("layernorm_bwd", timeit([lambda i=i: ln_bwd(i, False) for i in range(NBUF)]), 3 * act)
    report("layernorm_bwd+dres", timeit([lambda i=i: ln_bwd(i, True) for i in range(NBUF)]), 4 * act)

    # ---- bias-gelu bwd on the 4h activation (read dy, pre; write dpre)
    big = [torch.randn(ROWS, 4 * H, device=dev, dtype=dt) for _ in range(3)]
    big2 = [torch.randn(ROWS, 4 * H, device=dev, dtype=dt) for _ in range(3)]
    report("bias_gelu_bwd(4h)", timeit([lambda i=i: K.bias_gelu_bwd(big[i], big2[i], None) for i in range(3)]), 3 * act * 4)
    # ---- column sum (bias gradient) on 4h
    outc = torch.zeros(4 * H, device=dev, dtype=dt)
    report("column_sum(4h)", timeit([lambda i=i: K.column_sum(big[i], out=outc, accumulate=True) for i in range(3)]), act * 4)
    report("column_sum(h)", timeit([lambda i=i: K.column_sum(xs[i]) for i in range(NBUF)]), act)
    # ---- fused add
    report("fused_add(3)", timeit([lambda i=i: K.fused_add(xs[i], rs[i], ys[i], out=ss[i]) for i in range(NBUF)]), 4 * act)

    # ---- attention softmax fwd / bwd: BERT (64 x 16 heads, 128 x 128) and long rows (4 x 16 heads, 2048 x 2048)
    for shape, tag in (((64, 16, 128, 128), "seq128"), ((4, 16, 2048, 2048), "seq2048")):
        sx = [torch.randn(*shape, device=dev, dtype=dt) for _ in range(2)]
        sy = [K.softmax_fwd(t, scale=0.125) for t in sx]
        nb = sx[0].numel() * 2
        report(f"softmax_fwd({tag})", timeit([lambda i=i: K.softmax_fwd(sx[i], scale=0.125, out=sy[i]) for i in range(2)]), 2 * nb)
        report(f"softmax_bwd({tag})", timeit([lambda i=i: K.softmax_bwd(sx[i], sy[i], scale=0.125) for i in range(2)]), 3 * nb)
        del sx, sy

    # ---- Adam on a 256M-element shard: fp32 p/m/v rw + fp16 grad read + fp16 param write
    n = 256 * 1024 * 1024
    p = torch.randn(n, device=dev)
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    g = torch.randn(n, device=dev, dtype=dt)
    lp = torch.empty(n, device=dev, dtype=dt)
    t = timeit([lambda: K.adam_flat(p, m, v, g, lp=lp, lr=1e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.01, step=3)],
               iters=6, warm=2, graph=False)
    report("adam_flat(256M)", t, n * (12 + 12 + 2 + 2))
    del p, m, v, g, lp

    # ---- sumsq (grad norm) over 256M fp16
    g = torch.randn(n, device=dev, dtype=dt)
    report("sumsq(256M fp16)", timeit([lambda: K.sumsq(g)], iters=6, warm=2, graph=False), n * 2)
    del g

    # ---- cross entropy fwd+bwd over [8192, 50432] logits
    V = 50432
    logits = torch.randn(ROWS, V, device=dev, dtype=dt, requires_grad=True)
    tgt = torch.randint(0, V, (ROWS,), device=dev)

    def ce():
        logits.grad = None
        loss = K.cross_entropy(logits, tgt)
        loss.backward()
    report("cross_entropy fwd+bwd", timeit([ce], iters=6, warm=2, graph=False), ROWS * V * 2 * 3)

    # ---- rotary: forward (q, k out of the fused QKV activation) and backward (+ dV copy into the dQKV buffer)
    from deeperspeed_b200.ops.attention import rotary_oop
    Bq, Sq, Hq, Dq = 4, 2048, 16, 128
    qkv = torch.randn(Bq, Sq, 3, Hq, Dq, device=dev, dtype=dt)
    q = torch.empty(Bq, Sq, Hq, Dq, device=dev, dtype=dt)
    k = torch.empty_like(q)
    dq, dk, dv = torch.randn_like(q), torch.randn_like(q), torch.randn_like(q)
    dqkv = torch.empty_like(qkv)
    one = Bq * Sq * Hq * Dq * 2
    report("rotary fwd (q,k)", timeit([lambda: rotary_oop(qkv[:, :, 0], qkv[:, :, 1], q, k, Dq, 10000.0)]), 4 * one)
    report("rotary bwd (dq,dk,+dv)", timeit([lambda: rotary_oop(dq, dk, dqkv[:, :, 0], dqkv[:, :, 1], Dq, 10000.0,
                                                               inverse=True, v_in=dv, v_out=dqkv[:, :, 2])]), 6 * one)

    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(dict(copy_peak_GBps=peak, ops=res), open("gpurun_out/bench_ops.json", "w"), indent=1)


if __name__ == "__main__":
    main()
