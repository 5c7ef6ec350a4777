"""BertSparseSelfAttention: drop-in for the self-attention block of a BERT layer with block-sparse attention inside.
Keeps the attribute names (`query`, `key`, `value`) that SparseAttentionUtils re-points at the original projections
(reference deepspeed/ops/sparse_attention/bert_sparse_self_attention.py:16-78)."""
from torch import nn

from .sparse_self_attention import SparseSelfAttention
from .sparsity_config import FixedSparsityConfig


class BertSparseSelfAttention(nn.Module):
    def __init__(self, config, sparsity_config=FixedSparsityConfig(num_heads=4)):
        super().__init__()
        hidden, heads = config.hidden_size, config.num_attention_heads
        if hidden % heads:
            raise ValueError("The hidden size (%d) is not a multiple of the number of attention heads (%d)" % (hidden, heads))
        self.num_attention_heads = heads
        self.attention_head_size = hidden // heads
        self.all_head_size = hidden
        for name in ("query", "key", "value"):
            setattr(self, name, nn.Linear(hidden, hidden))
        self.sparse_self_attention = SparseSelfAttention(sparsity_config)

    def transpose_for_scores(self, x):
        """[batch, seq, hidden] -> [batch, heads, seq, head_size]"""
        b, s, _ = x.shape
        return x.view(b, s, self.num_attention_heads, self.attention_head_size).transpose(1, 2)

    def forward(self, hidden_states, attention_mask):
        """hidden_states [batch, seq, hidden]; attention_mask is used as the key padding mask. Returns the context in
        the same [batch, seq, hidden] layout."""
        q, k, v = (self.transpose_for_scores(proj(hidden_states)).contiguous()
                   for proj in (self.query, self.key, self.value))
        ctx = self.sparse_self_attention(q, k, v, key_padding_mask=attention_mask)     # [batch, heads, seq, head_size]
        b, _, s, _ = ctx.shape
        return ctx.transpose(1, 2).reshape(b, s, self.all_head_size)
