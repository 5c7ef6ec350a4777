"""SparseSelfAttention: scores = SDD(Q K^T) on the layout's blocks, block-sparse softmax (scale, relative position
embedding, key-padding and attention masks), context = DSD(P V).
API per reference deepspeed/ops/sparse_attention/sparse_self_attention.py:14-174 (constructor arguments, `get_layout`,
`get_ops`, mask handling, [batch, heads, seq, head_dim] operands). Unlike the reference, the per-length operator cache
is keyed by the layout as well, so modules with different sparsity configs never pick up each other's operators.
Calls without rpe / dense attention mask on sequences that are multiples of 256 run as ONE block-sparse flash kernel per
direction on the tensor cores (fused.py); everything else uses the three separate operators."""
import torch
import torch.distributed as dist
import torch.nn as nn

from . import fused
from .matmul import MatMul
from .softmax import Softmax
from .sparsity_config import SparsityConfig


class SparseSelfAttention(nn.Module):
    ops = dict()      # (seq_len, block, layout fingerprint) -> (sdd, dsd, softmax)
    schedules = dict()  # (seq_len, block, layout fingerprint, device) -> (forward, backward) schedules of the fused kernel
    use_fused = True          # one tcgen05 flash kernel per direction when the call qualifies (ops/sparse_attention/fused.py)
    reorder_keys = True       # fused path: visit keys in the order planned by fused.plan_key_order

    def __init__(self, sparsity_config=SparsityConfig(num_heads=4), key_padding_mask_mode="add", attn_mask_mode="mul",
                 max_seq_length=2048, mpu=None):
        super().__init__()
        self.mpu = mpu          # kept for the reference's signature (sparse_self_attention.py:28); the layout is a pure
                                # function of the configuration, so there is nothing to synchronise over its groups
        self.sparsity_config = sparsity_config
        self.max_seq_length = max_seq_length
        self.key_padding_mask_mode = key_padding_mask_mode
        self.attn_mask_mode = attn_mask_mode
        # the layout for the longest sequence; shorter sequences use its top-left corner
        self.register_buffer("master_layout", sparsity_config.make_layout(max_seq_length))
        self._need_layout_synchronization = True

    # ---- layout / operators
    def _sync_layout_once(self):
        """Layouts with random blocks (Variable, BigBird) must be identical on every rank: take rank 0's."""
        if not self._need_layout_synchronization or not dist.is_initialized():
            return
        on_gpu = torch.cuda.is_available() and dist.get_backend() == "nccl"
        buf = self.master_layout.to(torch.device("cuda", torch.cuda.current_device()) if on_gpu else "cpu")
        dist.broadcast(buf, src=0)
        self.master_layout.copy_(buf.to(self.master_layout.device))
        self._need_layout_synchronization = False

    def get_layout(self, L):
        self._sync_layout_once()
        block = self.sparsity_config.block
        if L % block:
            raise ValueError(f"Sequence Length, {L}, needs to be dividable by Block size {block}!")
        n = L // block
        return self.master_layout[..., :n, :n].cpu()

    def get_ops(self, H, L):
        memo = self.__dict__.setdefault("_ops_memo", {})
        if L in memo:           # the layout of a module never changes after its one-time synchronisation: skip the
            return memo[L]      # copy + fingerprint of the (possibly tens of MB) layout on every call
        memo[L] = self._get_ops_uncached(H, L)
        return memo[L]

    def _get_ops_uncached(self, H, L):
        layout = self.get_layout(L)
        block = self.sparsity_config.block
        key = (L, block, hash(layout.numpy().tobytes()))
        cached = SparseSelfAttention.ops.get(key)
        if cached is None:
            cached = (MatMul(layout, block, "sdd", trans_a=False, trans_b=True),      # Q K^T -> sparse scores
                      MatMul(layout, block, "dsd", trans_a=False, trans_b=False),     # sparse P, dense V -> context
                      Softmax(layout, block))
            SparseSelfAttention.ops[key] = cached
        return cached

    def get_schedules(self, L, device):
        memo = self.__dict__.setdefault("_sched_memo", {})
        if (L, str(device)) not in memo:
            memo[(L, str(device))] = self._get_schedules_uncached(L, device)
        return memo[(L, str(device))]

    def _get_schedules_uncached(self, L, device):
        layout = self.get_layout(L)
        block = self.sparsity_config.block
        key = (L, block, hash(layout.numpy().tobytes()), str(device), bool(self.reorder_keys))
        cached = SparseSelfAttention.schedules.get(key)
        if cached is None:
            # keys are visited in an order that packs widely seen ("global") columns into dense tiles when that pays
            order, planned = fused.plan_key_order(layout, block, L) if self.reorder_keys else (None, layout)
            fwd, bwd = fused.build_schedules(planned, block, L)
            if order is not None:      # (order, inverse order) as device index tensors
                inverse = torch.empty_like(order)
                inverse[order] = torch.arange(order.numel())
                order = (order.to(device), inverse.to(device))
            cached = SparseSelfAttention.schedules[key] = (fwd.to(device), bwd.to(device), order)
        return cached

    # ---- operand helpers
    def transpose_key_for_scores(self, x, L):
        """Keys may arrive already transposed ([batch, heads, head_dim, seq]); bring them to [.., seq, head_dim]."""
        return x if x.size(2) == L else x.transpose(2, 3)

    def transpose_mask_for_sparse(self, qtype, x, is_key_padding_mask=False):
        """Masks in the operand dtype with their singleton dimensions removed (key padding mask -> [batch, seq])."""
        x = x.type(qtype)
        if not is_key_padding_mask:
            return x.squeeze()
        for d in range(x.dim() - 1, 0, -1):
            x = x.squeeze(dim=d)
        return x

    def forward(self, query, key, value, rpe=None, key_padding_mask=None, attn_mask=None):
        """query / key / value: [batch, heads, seq, head_dim]; returns the attention context of the same shape."""
        if query.is_cuda and query.dtype not in (torch.half, torch.bfloat16):
            raise AssertionError("sparse attention only supports training in fp16/bf16 on GPU")
        _, heads, seq, head_dim = query.shape
        key = self.transpose_key_for_scores(key, seq)
        if not (query.shape == key.shape == value.shape):
            raise NotImplementedError("only self-attention is supported for now")
        if key_padding_mask is not None:
            key_padding_mask = self.transpose_mask_for_sparse(query.dtype, key_padding_mask, is_key_padding_mask=True)
        if attn_mask is not None:
            attn_mask = self.transpose_mask_for_sparse(query.dtype, attn_mask)
        if self.use_fused and fused.supported(query, self.sparsity_config.block, rpe, attn_mask):
            sched_f, sched_b, key_order = self.get_schedules(seq, query.device)
            bias = fused.key_mask_bias(key_padding_mask, self.key_padding_mask_mode, query.shape[0], seq)
            if key_order is not None:      # the schedules were built for this key order (autograd undoes the gathers)
                key = fused.PermuteKeysFn.apply(key, *key_order)
                value = fused.PermuteKeysFn.apply(value, *key_order)
                bias = None if bias is None else bias.index_select(1, key_order[0]).contiguous()
            return fused.BlockSparseFlashFn.apply(query, key, value, sched_f, sched_b, bias, float(head_dim) ** -0.5)
        scores_op, context_op, softmax_op = self.get_ops(heads, seq)
        probs = softmax_op(scores_op(query, key), scale=float(head_dim) ** -0.5, rpe=rpe,
                           key_padding_mask=key_padding_mask, attn_mask=attn_mask,
                           key_padding_mask_mode=self.key_padding_mask_mode, attn_mask_mode=self.attn_mask_mode)
        return context_op(probs, value)
