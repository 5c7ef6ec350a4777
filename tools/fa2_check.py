"""Numerics + timing of the second-generation flash-attention kernels against fp32 references and the library SDPA.
Usage: python tools/fa2_check.py [fwd] [bwd] [bench]   (default: all). Writes gpurun_out/fa2_check.json."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F
from deeperspeed_b200.ops import attention as A

what = set(sys.argv[1:]) or {"fwd", "bwd", "bench"}
res = {}


def ref_attn(q, k, v, causal):
    D = q.shape[-1]
    S = q.shape[1]
    qf, kf, vf = (t.float().transpose(1, 2) for t in (q, k, v))
    scores = qf @ kf.transpose(-1, -2) / D ** 0.5
    if causal:
        scores = scores.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device="cuda"), 1), float("-inf"))
    return (torch.softmax(scores, -1) @ vf).transpose(1, 2), torch.logsumexp(scores, -1)


if "fwd" in what:
    for dtype in (torch.float16, torch.bfloat16):
        for causal in (True, False):
            for (B, S, H) in ((2, 512, 3), (1, 2048, 2)):
                torch.manual_seed(0)
                qkv = torch.randn(B, S, 3, H, 128, device="cuda", dtype=dtype)
                q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
                A.set_flash_fwd_variant(2)
                out, lse = A.flash_attn_fwd(q, k, v, causal=causal, return_lse=True)
                torch.cuda.synchronize()
                ref, rl = ref_attn(q, k, v, causal)
                e = (out.float() - ref).abs().max().item()
                el = (lse - rl).abs().max().item()
                key = "fwd %s causal=%d S=%d" % (str(dtype)[6:], causal, S)
                res[key] = {"out_err": e, "lse_err": el}
                print(key, res[key], flush=True)

if "bwd" in what:
    for fwd_variant in (64, 2):
        for dtype in (torch.float16, torch.bfloat16):
            for causal in (True, False):
                for (B, S, H) in ((2, 512, 3), (1, 2048, 2)):
                    torch.manual_seed(1)
                    qkv = torch.randn(B, S, 3, H, 128, device="cuda", dtype=dtype)
                    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
                    A.set_flash_fwd_variant(fwd_variant)
                    out, lse = A.flash_attn_fwd(q, k, v, causal=causal, return_lse=True)
                    dout = torch.randn(B, S, H, 128, device="cuda", dtype=dtype)
                    A.set_flash_bwd_variant(2)
                    dq, dk, dv = A.flash_attn_bwd(dout, q, k, v, out, lse, causal=causal)
                    torch.cuda.synchronize()
                    qf, kf, vf = (t.float().detach().requires_grad_() for t in (q, k, v))
                    sc = qf.transpose(1, 2) @ kf.transpose(1, 2).transpose(-1, -2) / 128 ** 0.5
                    if causal:
                        sc = sc.masked_fill(torch.triu(torch.ones(S, S, dtype=torch.bool, device="cuda"), 1), float("-inf"))
                    ref = (torch.softmax(sc, -1) @ vf.transpose(1, 2)).transpose(1, 2)
                    rq, rk, rv = torch.autograd.grad(ref, (qf, kf, vf), dout.float())
                    key = "bwd fwdv=%d %s causal=%d S=%d" % (fwd_variant, str(dtype)[6:], causal, S)
                    res[key] = {n: ((g.float() - w).abs().max() / w.abs().max()).item()
                                for n, g, w in (("dq", dq, rq), ("dk", dk, rk), ("dv", dv, rv))}
                    print(key, res[key], flush=True)

if "bench" in what:
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

    for (B, S, H, D) in ((4, 2048, 16, 128), (1, 8192, 16, 128)):
        qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.float16)
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        fl = 4.0 * B * H * S * S * D / 2
        row = {"B": B, "S": S}
        for var in (2, 64):
            A.set_flash_fwd_variant(var)
            row["fwd_v%d_ms" % var] = round(timed(lambda: A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)), 4)
        qt, kt, vt = (t.transpose(1, 2) for t in (q, k, v))
        row["fwd_sdpa_ms"] = round(timed(lambda: F.scaled_dot_product_attention(qt, kt, vt, is_causal=True)), 4)
        A.set_flash_fwd_variant(2)
        out, lse = A.flash_attn_fwd(q, k, v, causal=True, return_lse=True)
        do = torch.randn_like(out)
        for var in (2, 1):
            A.set_flash_bwd_variant(var)
            row["bwd_v%d_ms" % var] = round(timed(lambda: A.flash_attn_bwd(do, q, k, v, out, lse, causal=True)), 4)
        A.set_flash_bwd_variant(2)
        ql, kl, vl = (t.detach().clone().requires_grad_() for t in (qt, kt, vt))
        o = F.scaled_dot_product_attention(ql, kl, vl, is_causal=True)
        dot = do.transpose(1, 2)
        row["bwd_sdpa_ms"] = round(timed(lambda: torch.autograd.grad(o, (ql, kl, vl), dot, retain_graph=True)), 4)
        row["fwd_v2_tflops"] = round(fl / row["fwd_v2_ms"] / 1e9, 1)
        row["bwd_v2_tflops"] = round(2.5 * fl / row["bwd_v2_ms"] / 1e9, 1)
        res["bench S=%d" % S] = row
        print(row, flush=True)

os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/fa2_check.json", "w"), indent=1)
