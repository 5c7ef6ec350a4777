"""GPT-NeoX-style decoder as a PipelineModule (the reference ships no NeoX model, only string hooks for it —
reference deepspeed/runtime/pipe/engine.py:185, pipe/module.py:585-587 — so the flagship model lives here).

Layer math on the sm_100a kernel stack: fused LayerNorm(+residual) kernels, tcgen05 GEMMs with bias / bias+GeLU /
bias+residual epilogues, rotary kernel, our embedding gather / scatter-add, fused cross-entropy (in-place gradient).
Causal attention is the attention block op of ops/attention.py: our tcgen05 flash-attention kernels (the library
kernel only with DSB_ATTN=sdpa, for comparison). 1.3B preset: 24 layers, d_model 2048, 16 heads, seq 2048, vocab 50432."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import kernels as K
from ..ops.attention import causal_self_attention
from ..runtime.pipe.module import LayerSpec, PipelineModule


class NeoXConfig:
    def __init__(self, vocab_size=50432, hidden_size=2048, num_layers=24, num_heads=16, seq_len=2048,
                 rotary_pct=1.0, rotary_base=10000.0, layernorm_eps=1e-5, init_std=0.02, parallel_residual=False):
        self.vocab_size = vocab_size
        self.hidden_size = hidden_size
        self.num_layers = num_layers
        self.num_heads = num_heads
        self.seq_len = seq_len
        self.rotary_pct = rotary_pct
        self.rotary_base = rotary_base
        self.layernorm_eps = layernorm_eps
        self.init_std = init_std
        self.parallel_residual = parallel_residual

    @staticmethod
    def neox_1p3b():
        return NeoXConfig()

    @staticmethod
    def tiny():
        return NeoXConfig(vocab_size=1024, hidden_size=256, num_layers=2, num_heads=4, seq_len=128)

    def num_params(self):
        h = self.hidden_size
        return 2 * self.vocab_size * h + self.num_layers * (12 * h * h + 13 * h) + 2 * h


class EmbeddingPipe(nn.Module):
    pipe_compute_weight = 0      # a row gather: no GEMM work when stages are balanced by compute
    def __init__(self, cfg):
        super().__init__()
        self.word_embeddings = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        nn.init.normal_(self.word_embeddings.weight, std=cfg.init_std)

    def forward(self, tokens):
        return K.embedding(tokens, self.word_embeddings.weight)


class ParallelTransformerLayerPipe(nn.Module):
    """One pre-LN decoder block. The class name matches the reference's checkpointable-layer hook."""

    def __init__(self, cfg, layer_idx=0):
        super().__init__()
        h = cfg.hidden_size
        self.cfg = cfg
        self.heads = cfg.num_heads
        self.head_dim = h // cfg.num_heads
        self.rot_dim = int(self.head_dim * cfg.rotary_pct) // 2 * 2
        out_std = cfg.init_std / math.sqrt(2.0 * cfg.num_layers)
        self.input_layernorm_w = nn.Parameter(torch.ones(h))
        self.input_layernorm_b = nn.Parameter(torch.zeros(h))
        self.qkv_w = nn.Parameter(torch.empty(3 * h, h).normal_(std=cfg.init_std))
        self.qkv_b = nn.Parameter(torch.zeros(3 * h))
        self.dense_w = nn.Parameter(torch.empty(h, h).normal_(std=out_std))
        self.dense_b = nn.Parameter(torch.zeros(h))
        self.post_attention_layernorm_w = nn.Parameter(torch.ones(h))
        self.post_attention_layernorm_b = nn.Parameter(torch.zeros(h))
        self.h_to_4h_w = nn.Parameter(torch.empty(4 * h, h).normal_(std=cfg.init_std))
        self.h_to_4h_b = nn.Parameter(torch.zeros(4 * h))
        self.four_h_to_h_w = nn.Parameter(torch.empty(h, 4 * h).normal_(std=out_std))
        self.four_h_to_h_b = nn.Parameter(torch.zeros(h))

    def _attention(self, xn):
        B, S, h = xn.shape
        H, D = self.heads, self.head_dim
        qkv = K.linear(xn, self.qkv_w, self.qkv_b)
        return causal_self_attention(qkv, H, D, self.rot_dim, self.cfg.rotary_base)

    def forward(self, x):
        cfg = self.cfg
        # `xr` aliases x: the residual stream is threaded through the LayerNorm ops so that its gradient is added
        # inside their backward kernels (no stand-alone accumulation adds in autograd)
        xn, xr = K.layer_norm(x, self.input_layernorm_w, self.input_layernorm_b, cfg.layernorm_eps)
        ctx = self._attention(xn)
        if cfg.parallel_residual:
            yn, xr = K.layer_norm(xr, self.post_attention_layernorm_w, self.post_attention_layernorm_b,
                                  cfg.layernorm_eps)
            a = K.linear_residual(ctx, self.dense_w, self.dense_b, xr)
            return K.mlp_residual(yn, self.h_to_4h_w, self.h_to_4h_b, self.four_h_to_h_w, self.four_h_to_h_b, a)
        a = K.linear(ctx, self.dense_w, self.dense_b)
        # LN(x + a) and the new residual stream x + a come out of one kernel
        yn, r1 = K.layer_norm(a, self.post_attention_layernorm_w, self.post_attention_layernorm_b, cfg.layernorm_eps,
                              residual=xr)
        return K.mlp_residual(yn, self.h_to_4h_w, self.h_to_4h_b, self.four_h_to_h_w, self.four_h_to_h_b, r1)


class NormPipe(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.eps = cfg.layernorm_eps
        self.weight = nn.Parameter(torch.ones(cfg.hidden_size))
        self.bias = nn.Parameter(torch.zeros(cfg.hidden_size))

    def forward(self, x):
        y, _ = K.layer_norm(x, self.weight, self.bias, self.eps)
        return y


class LMHeadPipe(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(cfg.vocab_size, cfg.hidden_size).normal_(std=cfg.init_std))

    def forward(self, x):
        return K.linear(x, self.weight, None)


def neox_loss_fn(logits, labels):
    return K.cross_entropy(logits, labels)


def neox_layer_specs(cfg):
    specs = [LayerSpec(EmbeddingPipe, cfg)]
    for i in range(cfg.num_layers):
        specs.append(LayerSpec(ParallelTransformerLayerPipe, cfg, i))
    specs.append(LayerSpec(NormPipe, cfg))
    specs.append(LayerSpec(LMHeadPipe, cfg))
    return specs


class GPTNeoXPipe(PipelineModule):
    def __init__(self, cfg, num_stages=1, topology=None, activation_checkpoint_interval=0,
                 partition_method="compute", **kwargs):
        self.neox_config = cfg
        super().__init__(layers=neox_layer_specs(cfg), num_stages=num_stages if topology is None else None,
                         topology=topology, loss_fn=neox_loss_fn, partition_method=partition_method,
                         activation_checkpoint_interval=activation_checkpoint_interval,
                         checkpointable_layers=["ParallelTransformerLayerPipe"], **kwargs)
