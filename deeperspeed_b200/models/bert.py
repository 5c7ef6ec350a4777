"""Minimal HuggingFace-style BERT encoder layers (post-LN and pre-LN) in plain PyTorch. They are the numerical oracle
for DeepSpeedTransformerLayer (role of reference tests/unit/modeling.py / modelingpreln.py) and the injection target
of module_inject; `BertLargeConfig` matches the BASELINE BERT-large pre-training config."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class BertConfig:
    def __init__(self, hidden_size=1024, num_hidden_layers=24, num_attention_heads=16, intermediate_size=None,
                 hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, initializer_range=0.02,
                 layer_norm_eps=1e-12, vocab_size=30528, max_position_embeddings=512):
        self.hidden_size = hidden_size
        self.num_hidden_layers = num_hidden_layers
        self.num_attention_heads = num_attention_heads
        self.intermediate_size = intermediate_size or 4 * hidden_size
        self.hidden_dropout_prob = hidden_dropout_prob
        self.attention_probs_dropout_prob = attention_probs_dropout_prob
        self.initializer_range = initializer_range
        self.layer_norm_eps = layer_norm_eps
        self.vocab_size = vocab_size
        self.max_position_embeddings = max_position_embeddings


def BertLargeConfig():
    return BertConfig(1024, 24, 16)


def gelu_tanh(x):
    return 0.5 * x * (1.0 + torch.tanh(0.7978845608028654 * x * (1.0 + 0.044715 * x * x)))


class BertSelfAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.num_attention_heads = config.num_attention_heads
        self.attention_head_size = config.hidden_size // config.num_attention_heads
        self.query = nn.Linear(config.hidden_size, config.hidden_size)
        self.key = nn.Linear(config.hidden_size, config.hidden_size)
        self.value = nn.Linear(config.hidden_size, config.hidden_size)
        self.dropout = nn.Dropout(config.attention_probs_dropout_prob)

    def _heads(self, x):
        B, S, _ = x.shape
        return x.view(B, S, self.num_attention_heads, self.attention_head_size).permute(0, 2, 1, 3)

    def forward(self, hidden_states, attention_mask):
        q, k, v = self._heads(self.query(hidden_states)), self._heads(self.key(hidden_states)), self._heads(self.value(hidden_states))
        scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(self.attention_head_size)
        if attention_mask is not None:
            scores = scores + attention_mask
        probs = self.dropout(F.softmax(scores, dim=-1))
        ctx = torch.matmul(probs, v).permute(0, 2, 1, 3).contiguous()
        return ctx.view(ctx.shape[0], ctx.shape[1], -1)


class BertSelfOutput(nn.Module):
    def __init__(self, config, with_ln=True):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        if with_ln:
            self.LayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)


class BertAttention(nn.Module):
    def __init__(self, config, with_ln=True):
        super().__init__()
        self.self = BertSelfAttention(config)
        self.output = BertSelfOutput(config, with_ln)


class BertIntermediate(nn.Module):
    def __init__(self, config, preln=False):
        super().__init__()
        lin = nn.Linear(config.hidden_size, config.intermediate_size)
        if preln:
            self.dense_act = lin
        else:
            self.dense = lin


class BertOutput(nn.Module):
    def __init__(self, config, with_ln=True):
        super().__init__()
        self.dense = nn.Linear(config.intermediate_size, config.hidden_size)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        if with_ln:
            self.LayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)


class BertLayer(nn.Module):
    """Post-LN BERT layer."""

    def __init__(self, config):
        super().__init__()
        self.attention = BertAttention(config)
        self.intermediate = BertIntermediate(config)
        self.output = BertOutput(config)

    def forward(self, hidden_states, attention_mask=None):
        a = self.attention.self(hidden_states, attention_mask)
        a = self.attention.output.dropout(self.attention.output.dense(a))
        r1 = self.attention.output.LayerNorm(a + hidden_states)
        f = gelu_tanh(self.intermediate.dense(r1))
        o = self.output.dropout(self.output.dense(f))
        return self.output.LayerNorm(o + r1)


class BertLayerPreLN(nn.Module):
    """Pre-LN BERT layer (PreAttentionLayerNorm / PostAttentionLayerNorm naming of the reference's modelingpreln.py)."""

    def __init__(self, config):
        super().__init__()
        self.attention = BertAttention(config, with_ln=False)
        self.PreAttentionLayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.PostAttentionLayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)
        self.intermediate = BertIntermediate(config, preln=True)
        self.output = BertOutput(config, with_ln=False)

    def forward(self, hidden_states, attention_mask=None):
        x = hidden_states
        a = self.attention.self(self.PreAttentionLayerNorm(x), attention_mask)
        r1 = x + self.attention.output.dropout(self.attention.output.dense(a))
        f = gelu_tanh(self.intermediate.dense_act(self.PostAttentionLayerNorm(r1)))
        return r1 + self.output.dropout(self.output.dense(f))


class BertEncoder(nn.Module):
    def __init__(self, config, preln=False):
        super().__init__()
        cls = BertLayerPreLN if preln else BertLayer
        self.layer = nn.ModuleList([cls(config) for _ in range(config.num_hidden_layers)])
        self.preln = preln
        if preln:
            self.FinalLayerNorm = nn.LayerNorm(config.hidden_size, eps=config.layer_norm_eps)

    def forward(self, hidden_states, attention_mask=None):
        for layer in self.layer:
            hidden_states = layer(hidden_states, attention_mask)
        if self.preln:
            hidden_states = self.FinalLayerNorm(hidden_states)
        return hidden_states
