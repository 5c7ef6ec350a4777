"""All-gather + GEMM kernels with LOCAL stand-ins for the peers (every "shard" is a buffer on this GPU): separates the
kernel's own structure (copier warps, panel flags, k-loop rotation, two passes) from NVLink effects, and gives ncu a
single-process target. Prints CUDA-event times; under ncu use --profile-from-start off."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from deeperspeed_b200.ops import kernels as K
from deeperspeed_b200.ops import native as N
from deeperspeed_b200.parallel import ag_gemm  # noqa: F401

lib = N.cuda_lib()
R = int(os.environ.get("R", 2))
tokens = 8192
dc = N.dtype_code(torch.float16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=10, warm=3):
    ts = []
    for it in range(warm + iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        if it >= warm:
            ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


prof = []
for tag, out_f, in_f in (("ff1", 8192, 2048), ("ff2", 2048, 8192)):
    w = (torch.randn(out_f, in_f, device="cuda") * 0.02).half()
    shards = [w[r * (out_f // R):(r + 1) * (out_f // R)].clone() for r in range(R)]
    ptrs = (C.c_void_p * R)(*[s.data_ptr() for s in shards])
    gathered = torch.empty_like(w)
    pflags = torch.zeros(out_f // 256, dtype=torch.int32, device="cuda")
    x = (torch.randn(tokens, in_f, device="cuda") * 0.1).half()
    go = (torch.randn(tokens, out_f, device="cuda") * 0.1).half()
    y = torch.empty(tokens, out_f, device="cuda", dtype=torch.float16)
    gi = torch.empty(tokens, in_f, device="cuda", dtype=torch.float16)
    st = N.stream_ptr()

    def fwd():
        N.check(lib.dsb_gemm_ag_overlap(N.ptr(x), ptrs, R, 0, N.ptr(gathered), N.ptr(pflags), N.ptr(y), tokens, out_f,
                                        in_f, x.stride(0), y.stride(0), dc, 0, 1.0, None, None, 0, st), "fwd")

    def dgrad():
        N.check(lib.dsb_gemm_ag_overlap_k(N.ptr(go), ptrs, R, 0, N.ptr(gathered), N.ptr(pflags), N.ptr(gi), tokens,
                                          in_f, out_f, go.stride(0), gi.stride(0), dc, 1.0, st), "dgrad")

    def single(a, b, d, M, Nn, Kk, bmn):
        N.check(lib.dsb_gemm(N.ptr(a), N.ptr(b), N.ptr(d), M, Nn, Kk, 1, a.stride(0), b.stride(0), d.stride(0), 0, 0, 0,
                             0, bmn, dc, 0, 1.0, None, None, 0, 0, None, 0, 0, st), "gemm")

    fwd(); dgrad(); torch.cuda.synchronize()
    e1 = ((y.float() - x.float() @ w.float().t()).abs().max() / (x.float() @ w.float().t()).abs().max()).item()
    e2 = ((gi.float() - go.float() @ w.float()).abs().max() / (go.float() @ w.float()).abs().max()).item()
    print(f"{tag} R={R}: fwd fused {timed(fwd):.4f} ms | single-CTA gemm {timed(lambda: single(x, w, y, tokens, out_f, in_f, 0)):.4f}"
          f" | pair {timed(lambda: K.gemm(x, w)):.4f}   ||   dgrad fused {timed(dgrad):.4f} ms | single-CTA gemm "
          f"{timed(lambda: single(go, w, gi, tokens, in_f, out_f, 1)):.4f} | pair {timed(lambda: K.gemm(go, w, trans_b=False)):.4f}"
          f"   rel err {e1:.1e} {e2:.1e}")
    prof.append((fwd, dgrad))
torch.cuda.profiler.start()
for f, d in prof:
    f(); d()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok")
