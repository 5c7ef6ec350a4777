"""SparseSelfAttention: SDD(QK^T) -> block-sparse softmax(scale, rpe, masks) -> DSD(P V).
API per reference deepspeed/ops/sparse_attention/sparse_self_attention.py:14-174."""
import torch
import torch.nn as nn
import torch.distributed as dist

from .matmul import MatMul
from .softmax import Softmax
from .sparsity_config import SparsityConfig


class SparseSelfAttention(nn.Module):
    ops = dict()

    def __init__(self, sparsity_config=SparsityConfig(num_heads=4), key_padding_mask_mode="add", attn_mask_mode="mul",
                 max_seq_length=2048):
        super().__init__()
        self.sparsity_config = sparsity_config
        self.max_seq_length = max_seq_length
        master_layout = self.sparsity_config.make_layout(max_seq_length)
        self.register_buffer("master_layout", master_layout)
        self._need_layout_synchronization = True
        self.key_padding_mask_mode = key_padding_mask_mode
        self.attn_mask_mode = attn_mask_mode

    def get_layout(self, L):
        # random layouts (Variable / BigBird) must agree across ranks: rank 0's layout is broadcast once
        if self._need_layout_synchronization and dist.is_initialized():
            dev = torch.device("cuda", torch.cuda.current_device()) if (torch.cuda.is_available() and dist.get_backend() == "nccl") else torch.device("cpu")
            t = self.master_layout.to(dev)
            dist.broadcast(t, src=0)
            self.master_layout.copy_(t.to(self.master_layout.device))
            self._need_layout_synchronization = False
        if (L % self.sparsity_config.block != 0):
            raise ValueError(f"Sequence Length, {L}, needs to be dividable by Block size {self.sparsity_config.block}!")
        num_blocks = L // self.sparsity_config.block
        return self.master_layout[..., :num_blocks, :num_blocks].cpu()

    def get_ops(self, H, L):
        if L not in SparseSelfAttention.ops:
            sparsity_layout = self.get_layout(L)
            sparse_dot_sdd_nt = MatMul(sparsity_layout, self.sparsity_config.block, "sdd", trans_a=False, trans_b=True)
            sparse_dot_dsd_nn = MatMul(sparsity_layout, self.sparsity_config.block, "dsd", trans_a=False, trans_b=False)
            sparse_softmax = Softmax(sparsity_layout, self.sparsity_config.block)
            SparseSelfAttention.ops[L] = (sparse_dot_sdd_nt, sparse_dot_dsd_nn, sparse_softmax)
        return SparseSelfAttention.ops[L]

    def transpose_key_for_scores(self, x, L):
        bsz, num_heads, seq_len, head_dim = x.size()
        if seq_len != L:
            return x.permute(0, 1, 3, 2)
        return x

    def transpose_mask_for_sparse(self, qtype, x, is_key_padding_mask=False):
        x = x.type(qtype)
        if is_key_padding_mask:
            xdim = x.dim()
            for d in range(xdim - 1, 0, -1):
                x = x.squeeze(dim=d)
            return x
        return x.squeeze()

    def forward(self, query, key, value, rpe=None, key_padding_mask=None, attn_mask=None):
        """query/key/value: [batch, heads, seq, head_dim]; returns the attention context of the same shape."""
        assert query.dtype in (torch.half, torch.bfloat16) or not query.is_cuda, "sparse attention only supports training in fp16/bf16 on GPU"
        bsz, num_heads, tgt_len, head_dim = query.size()
        key = self.transpose_key_for_scores(key, tgt_len)
        if query.shape != key.shape or key.shape != value.shape:
            raise NotImplementedError("only self-attention is supported for now")
        if key_padding_mask is not None:
            key_padding_mask = self.transpose_mask_for_sparse(query.dtype, key_padding_mask, is_key_padding_mask=True)
        if attn_mask is not None:
            attn_mask = self.transpose_mask_for_sparse(query.dtype, attn_mask)
        sparse_dot_sdd_nt, sparse_dot_dsd_nn, sparse_softmax = self.get_ops(num_heads, tgt_len)
        scaling = float(head_dim) ** -0.5
        attn_output_weights = sparse_dot_sdd_nt(query, key)
        attn_output_weights = sparse_softmax(attn_output_weights, scale=scaling, rpe=rpe,
                                             key_padding_mask=key_padding_mask, attn_mask=attn_mask,
                                             key_padding_mask_mode=self.key_padding_mask_mode,
                                             attn_mask_mode=self.attn_mask_mode)
        return sparse_dot_dsd_nn(attn_output_weights, value)
