"""Block-sparse attention as ONE tcgen05 flash kernel per direction.

The reference computes sparse attention as three Triton kernels (SDD Q K^T on the layout's blocks, block-sparse softmax,
DSD P V; deepspeed/ops/sparse_attention/matmul.py:98-501, softmax.py, trsrc/*.tr) and materialises the sparse score
matrix between them. Here the layout is compiled into a *schedule* - per (head, 256-row query block) the list of 64-key
blocks that contain anything visible, with one 32-bit visibility word per 128 x 64 tile at 16 x 16 granularity - and the
flash-attention kernels of csrc/transformer/flash_attn_v2.cu walk that list: tiles that are not listed are never loaded
or multiplied, listed tiles run on the tensor cores (tcgen05.mma, operands by TMA, P kept in tensor memory), and the 16 x
16 blocks inside a tile that the layout leaves out are masked in the softmax. The backward uses the transposed schedule
(per 128-key block, the 64-row query blocks that see it). Key-padding masks ride along as an additive per-key bias.

Supported: fp16 / bf16, sequence length a multiple of 256, head dim <= 128 (smaller head dims are zero-padded to the
kernel's 128 by TMA itself, without copies), layout blocks of 16 / 32 / 64 / 128, key-padding mask in "add" or "mul" mode. Relative position
embeddings and dense attention masks go through the unfused operators (`SparseSelfAttention` falls back by itself).
"""
import ctypes as C

import torch

from .. import kernels as K
from .. import native as N

_LOG2E = 1.4426950408889634
_p, _i, _f = C.c_void_p, C.c_int, C.c_float
N.register_sigs({
    "dsb_flash_attn_fwd_ex": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _f, _i, _i, _p, _i, _i, _p, _p],
    "dsb_flash_attn_bwd2_ex": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _f, _i, _i, _p, _i, _i, _p,
                               _p],
    "dsb_flash_attn_fwd_drop": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _f, _i, _i, _p, _i, _i, _p, _f, C.c_uint, _p],
    "dsb_flash_attn_bwd2_drop": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _f, _i, _i, _p, _i, _i, _p,
                                 _f, C.c_uint, _p],
})


def _records(words, index_dim_size):
    """words: int64 [Hs, G, T] (G groups, T candidate tiles, 0 = nothing visible). Returns int32 [Hs, G, 1 + w * maxn]
    records [n, (tile index, word...) * n] where `words` may carry several words per tile in a trailing dim."""
    if words.dim() == 3:
        words = words.unsqueeze(-1)
    Hs, G, T, W = words.shape
    live = (words != 0).any(-1)                                   # [Hs, G, T]
    counts = live.sum(-1)
    maxn = max(int(counts.max()), 1)
    rec = torch.zeros(Hs, G, 1 + (1 + W) * maxn, dtype=torch.int64)
    rec[..., 0] = counts
    order = torch.argsort((~live).to(torch.int8), dim=-1, stable=True)[..., :maxn]      # live tiles first, in order
    picked = torch.gather(words, 2, order.unsqueeze(-1).expand(-1, -1, -1, W))         # [Hs, G, maxn, W]
    body = torch.cat([order.unsqueeze(-1), picked], dim=-1)                             # [Hs, G, maxn, 1 + W]
    valid = (torch.arange(maxn).view(1, 1, -1) < counts.unsqueeze(-1)).unsqueeze(-1)
    rec[..., 1:] = (body * valid).reshape(Hs, G, -1)
    # visibility words use all 32 bits: store their two's complement in int32
    rec = torch.where(rec >= 2 ** 31, rec - 2 ** 32, rec)
    return rec.to(torch.int32).contiguous()


def build_schedules(layout, block, seq_len):
    """(forward schedule, backward schedule) as CPU int32 tensors for a [heads, nb, nb] 0/1 layout with `block`-sized
    blocks over `seq_len` tokens. Heads with identical layouts share one record set (first dim 1)."""
    assert block % 16 == 0, "layout blocks must be multiples of 16"
    assert seq_len % 256 == 0
    nb = seq_len // block
    lay = layout[:, :nb, :nb].to(torch.int64)
    if bool((lay == lay[:1]).all()):
        lay = lay[:1]
    rep = block // 16
    sub = lay.repeat_interleave(rep, 1).repeat_interleave(rep, 2)       # [Hs, S / 16, S / 16]
    Hs, n16 = sub.shape[0], sub.shape[1]
    P, KB = seq_len // 256, seq_len // 64
    # forward: bit = qgroup * 4 + kgroup inside a (128 query rows) x (64 keys) tile; two tiles per 256-row block
    f = sub.view(Hs, P, 2, 8, KB, 4).permute(0, 1, 4, 2, 3, 5)         # [Hs, P, KB, tile, qg, kg]
    wf = (f * (2 ** (torch.arange(8).view(8, 1) * 4 + torch.arange(4).view(1, 4)))).sum((-1, -2))     # [Hs, P, KB, 2]
    # backward: bit = kgroup * 4 + qgroup inside a (128 keys) x (64 query rows) tile
    JB, QB = seq_len // 128, seq_len // 64
    g = sub.view(Hs, QB, 4, JB, 8).permute(0, 3, 1, 4, 2)              # [Hs, JB, QB, kg, qg]
    wb = (g * (2 ** (torch.arange(8).view(8, 1) * 4 + torch.arange(4).view(1, 4)))).sum((-1, -2))     # [Hs, JB, QB]
    return _records(wf, KB), _records(wb, QB)


def _tile_work(fwd, bwd):
    """Relative kernel work of a schedule pair: live 128 x 64 forward tiles + 2.5 x live backward tiles."""
    return float(fwd[..., 0].sum()) + 2.5 * float(bwd[..., 0].sum())


def plan_key_order(layout, block, seq_len, min_gain=0.15, min_saved=1000.0):
    """Attention does not care in which ORDER the keys of a row are visited, only which ones: permuting K, V (and the
    layout's columns) together changes nothing in the result. The kernels skip work at the granularity of 64-key
    (forward) / 128-key (backward) tiles, so a layout whose "global" columns are sprinkled one 16-key block per window
    (the fixed pattern of Sparse Transformers, Longformer's global tokens, BigBird's) keeps nearly every tile alive
    for 16 useful keys. This plans a key-block permutation that moves the columns seen by many query blocks to the
    front, in index order (so a causal prefix of them stays a prefix), and leaves the rest where they were.

    Returns (token permutation [seq_len] as int64, permuted layout) when all heads share one layout and the permuted
    schedule does at least `min_gain` less tile work and saves at least `min_saved` tile units per head (the two gathers
    of K / V per direction cost about that much: measured break-even between 2k and 4k tokens for the fixed pattern),
    else (None, layout)."""
    nb = seq_len // block
    lay = layout[:, :nb, :nb]
    if not bool((lay == lay[:1]).all()):
        return None, layout
    seen_by = lay[0].sum(0)                                   # query blocks that see each key block
    typical = max(float(seen_by.float().median()), 1.0)
    wide = seen_by.float() >= max(4.0 * typical, nb / 8.0)
    if not bool(wide.any()) or bool(wide.all()):
        return None, layout
    idx = torch.arange(nb)
    order = torch.cat([idx[wide], idx[~wide]])
    permuted = lay[:, :, order].contiguous()
    before, after = _tile_work(*build_schedules(lay, block, seq_len)), _tile_work(*build_schedules(permuted, block, seq_len))
    if after > (1.0 - min_gain) * before or before - after < min_saved:
        return None, layout
    tokens = (order.view(-1, 1) * block + torch.arange(block).view(1, -1)).reshape(-1)
    return tokens, permuted


class PermuteKeysFn(torch.autograd.Function):
    """x[:, :, order] along the sequence axis of a [B, H, S, D] tensor; `order` is a PERMUTATION, so the gradient is a
    gather with the inverse permutation (no scatter-add)."""

    @staticmethod
    def forward(ctx, x, order, inverse):
        ctx.save_for_backward(inverse)
        return x.index_select(2, order)

    @staticmethod
    def backward(ctx, grad):
        (inverse,) = ctx.saved_tensors
        return grad.index_select(2, inverse), None, None


def _as_bshd(x):
    """[B, H, S, D] -> a [B, S, H, D] view with unit last stride and 16-byte-aligned strides (a copy only if needed).
    Head dims below 128 need no padding: the kernels' tensor maps carry the true D and TMA zero-fills the rest."""
    v = x.transpose(1, 2)
    if v.stride(-1) == 1 and all(st % 8 == 0 for st in v.stride()[:3]) and v.data_ptr() % 16 == 0:
        return v
    return v.contiguous()


def _strides(*tensors):
    return (C.c_long * (3 * len(tensors)))(*[t.stride(i) for t in tensors for i in range(3)])


def _sched_strides(sched):
    """(head stride, block stride) of a schedule tensor in int32 elements; (0, 0) without a schedule (dense attention),
    head stride 0 when all heads share one layout."""
    if sched is None:
        return 0, 0
    return (sched.stride(0) if sched.shape[0] > 1 else 0), sched.stride(1)


class BlockSparseFlashFn(torch.autograd.Function):
    """context = dropout(softmax(scale * Q K^T restricted to the layout (+ key mask))) V for [B, H, S, D] operands.
    Without schedules it is plain (non-causal) attention with an optional additive key mask. `dropout` = (p, seed) or
    None: the keep mask is a hash inside the kernels (csrc/transformer/flash_v2.cuh), regenerated in backward."""

    @staticmethod
    def forward(ctx, q, k, v, sched_f, sched_b, kpm, scale, dropout=None):
        B, H, S, D = q.shape
        q4, k4, v4 = _as_bshd(q), _as_bshd(k), _as_bshd(v)
        out = torch.empty(B, S, H, D, dtype=q.dtype, device=q.device)
        lse = torch.empty(B, H, S, dtype=torch.float32, device=q.device)
        hs, ps = _sched_strides(sched_f)
        p_drop, seed = dropout if dropout is not None else (0.0, 0)
        N.check(N.cuda_lib().dsb_flash_attn_fwd_drop(
            N.ptr(q4), N.ptr(k4), N.ptr(v4), N.ptr(out), N.ptr(lse), B, S, H, D, _strides(q4, k4, v4, out),
            float(scale), 0, N.dtype_code(q.dtype), N.ptr(sched_f), hs, ps, N.ptr(kpm), float(p_drop),
            int(seed) & 0xffffffff, N.stream_ptr()), "dsb_flash_attn_fwd_drop")
        K._count()
        ctx.save_for_backward(q4, k4, v4, out, lse, sched_b, kpm)
        ctx.meta = (D, float(scale), float(p_drop), int(seed) & 0xffffffff)
        return out.transpose(1, 2)

    @staticmethod
    def backward(ctx, dctx):
        q4, k4, v4, out, lse, sched_b, kpm = ctx.saved_tensors
        D, scale, p_drop, seed = ctx.meta
        B, S, H, _ = q4.shape
        do4 = _as_bshd(dctx)
        dq, dk, dv = (torch.empty(B, S, H, D, dtype=q4.dtype, device=q4.device) for _ in range(3))
        scratch = torch.empty((2 + 128) * B * H * S, dtype=torch.float32, device=q4.device)
        hs, ps = _sched_strides(sched_b)
        N.check(N.cuda_lib().dsb_flash_attn_bwd2_drop(
            N.ptr(do4), N.ptr(q4), N.ptr(k4), N.ptr(v4), N.ptr(out), N.ptr(lse), N.ptr(scratch), N.ptr(dq), N.ptr(dk),
            N.ptr(dv), B, S, H, D, _strides(do4, q4, k4, v4, out, dq, dk, dv), scale, 0, N.dtype_code(q4.dtype),
            N.ptr(sched_b), hs, ps, N.ptr(kpm), p_drop, seed, N.stream_ptr()), "dsb_flash_attn_bwd2_drop")
        K._count(3)
        return dq.transpose(1, 2), dk.transpose(1, 2), dv.transpose(1, 2), None, None, None, None, None


def dropout_keep_mask(B, H, S, p, seed, device="cpu"):
    """The [B, H, S, S] keep mask (bool) the kernels derive for dropout probability `p` and `seed` - a restatement of
    the device hash for tests and debugging; returns (mask, 1 / keep probability as the kernels quantise it)."""
    thr = min(int(p * 65536.0 + 0.5), 65535)
    M = 0xffffffff

    def lowbias32(x):
        x = x ^ (x >> 16)
        x = (x * 0x7feb352d) & M
        x = x ^ (x >> 15)
        x = (x * 0x846ca68b) & M
        return x ^ (x >> 16)

    bh = torch.arange(B * H, dtype=torch.int64, device=device)
    seed_bh = lowbias32((int(seed) & M) ^ ((bh * 0x27d4eb2f) & M))                          # [BH]
    q = (torch.arange(S, dtype=torch.int64, device=device) * 0x9E3779B1) & M                # [S]
    key = torch.arange(S, dtype=torch.int64, device=device)
    kp = ((key >> 1) * 0x85EBCA77) & M                                                      # [S]
    h = lowbias32((seed_bh.view(-1, 1, 1) + q.view(1, -1, 1) + kp.view(1, 1, -1)) & M)      # [BH, S, S]
    bits = (h >> ((key & 1) * 16).view(1, 1, -1)) & 0xffff
    return (bits >= thr).view(B, H, S, S), 65536.0 / (65536 - thr)


def key_mask_bias(key_padding_mask, mode, batch, seq_len):
    """fp32 [B, S] additive bias in log2 units from a key-padding mask ("add": the mask values themselves, "mul":
    0 where the mask is non-zero and -inf where it is zero)."""
    if key_padding_mask is None:
        return None
    m = key_padding_mask.reshape(batch, seq_len).float()
    if mode == "mul":
        return torch.where(m == 0, torch.full_like(m, float("-inf")), torch.zeros_like(m)).contiguous()
    return (m * _LOG2E).contiguous()


def supported(query, block, rpe, attn_mask):
    B, H, S, D = query.shape
    return (query.is_cuda and query.dtype in (torch.float16, torch.bfloat16) and rpe is None and attn_mask is None
            and S % 256 == 0 and D <= 128 and D % 8 == 0 and block % 16 == 0)
