"""GPT-2 family (learned positions, pre-LN, tied LM head) on the sm_100a op stack, with optional block-sparse
attention (`sparsity_config`) for the long-sequence config of BASELINE.json (GPT-2 medium, fixed layout, seq 8192)
and presets up to the 13B ZeRO-3 + Offload sizing config."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..ops import kernels as K
from ..ops.attention import causal_self_attention
from ..ops.sparse_attention import SparseSelfAttention


class GPT2Config:
    def __init__(self, vocab_size=50304, n_positions=1024, n_embd=768, n_layer=12, n_head=12, layer_norm_eps=1e-5,
                 init_std=0.02, sparsity_config=None, attention_impl="flash", checkpoint_blocks=False):
        self.vocab_size = vocab_size
        self.n_positions = n_positions
        self.n_embd = n_embd
        self.n_layer = n_layer
        self.n_head = n_head
        self.layer_norm_eps = layer_norm_eps
        self.init_std = init_std
        self.sparsity_config = sparsity_config
        # dense attention: "flash" (fused causal kernel) or "masked" (scores materialised, -inf mask, softmax: what the
        # reference's dense layers do and the stand-in its sparse-attention numbers are quoted against)
        self.attention_impl = attention_impl
        self.checkpoint_blocks = checkpoint_blocks      # recompute each block in backward (activation checkpointing)

    @staticmethod
    def medium(**kw):
        return GPT2Config(n_embd=1024, n_layer=24, n_head=16, **kw)

    @staticmethod
    def xl(**kw):
        return GPT2Config(n_embd=1600, n_layer=48, n_head=25, **kw)

    @staticmethod
    def gpt2_13b(**kw):
        return GPT2Config(n_embd=5120, n_layer=40, n_head=40, **kw)

    def num_params(self):
        h = self.n_embd
        return self.vocab_size * h + self.n_positions * h + self.n_layer * (12 * h * h + 13 * h) + 2 * h


class GPT2Block(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        h = cfg.n_embd
        self.cfg = cfg
        self.heads, self.head_dim = cfg.n_head, h // cfg.n_head
        out_std = cfg.init_std / math.sqrt(2.0 * cfg.n_layer)
        self.ln_1_w, self.ln_1_b = nn.Parameter(torch.ones(h)), nn.Parameter(torch.zeros(h))
        self.c_attn_w = nn.Parameter(torch.empty(3 * h, h).normal_(std=cfg.init_std))
        self.c_attn_b = nn.Parameter(torch.zeros(3 * h))
        self.c_proj_w = nn.Parameter(torch.empty(h, h).normal_(std=out_std))
        self.c_proj_b = nn.Parameter(torch.zeros(h))
        self.ln_2_w, self.ln_2_b = nn.Parameter(torch.ones(h)), nn.Parameter(torch.zeros(h))
        self.c_fc_w = nn.Parameter(torch.empty(4 * h, h).normal_(std=cfg.init_std))
        self.c_fc_b = nn.Parameter(torch.zeros(4 * h))
        self.mlp_proj_w = nn.Parameter(torch.empty(h, 4 * h).normal_(std=out_std))
        self.mlp_proj_b = nn.Parameter(torch.zeros(h))
        self.sparse_attn = None
        if cfg.sparsity_config is not None:
            self.sparse_attn = SparseSelfAttention(cfg.sparsity_config, attn_mask_mode="add",
                                                   max_seq_length=cfg.n_positions)

    def _attention(self, xn):
        B, S, h = xn.shape
        H, D = self.heads, self.head_dim
        fused = K.linear(xn, self.c_attn_w, self.c_attn_b)
        if self.sparse_attn is None and self.cfg.attention_impl == "flash":
            # our tcgen05 flash kernels for head dim 128 (reads q / k / v in place out of the fused activation)
            return causal_self_attention(fused, H, D, 0)
        qkv = fused.view(B, S, 3, H, D)
        q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))       # [B,H,S,D]
        if self.sparse_attn is not None:
            ctx = self.sparse_attn(q, k, v)
        elif self.cfg.attention_impl == "masked":
            scores = torch.matmul(q, k.transpose(-1, -2)) * (D ** -0.5)
            causal = torch.ones(S, S, dtype=torch.bool, device=xn.device).tril_()
            ctx = torch.matmul(torch.softmax(scores.masked_fill(~causal, float("-inf")), -1), v)
        else:
            ctx = F.scaled_dot_product_attention(q, k, v, is_causal=True)
        return ctx.transpose(1, 2).reshape(B, S, h)

    def forward(self, x):
        cfg = self.cfg
        xn, xr = K.layer_norm(x, self.ln_1_w, self.ln_1_b, cfg.layer_norm_eps)   # xr aliases x (see K.layer_norm)
        a = K.linear(self._attention(xn), self.c_proj_w, self.c_proj_b)
        yn, r1 = K.layer_norm(a, self.ln_2_w, self.ln_2_b, cfg.layer_norm_eps, residual=xr)
        return K.mlp_residual(yn, self.c_fc_w, self.c_fc_b, self.mlp_proj_w, self.mlp_proj_b, r1)


def _init_normal(param, std):
    """N(0, std) initialisation that also works for a parameter `zero.Init` has already partitioned (nn.Embedding is
    partitioned the moment its own constructor returns): gather it, let rank 0 draw, broadcast on the way out."""
    if hasattr(param, "ds_id"):
        import torch.distributed as dist
        from ..runtime.zero.partition_parameters import GatheredParameters
        with GatheredParameters(param, modifier_rank=0):
            if dist.get_rank() == 0:
                nn.init.normal_(param, std=std)
    else:
        nn.init.normal_(param, std=std)


class GPT2Model(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        _init_normal(self.wte.weight, cfg.init_std)
        _init_normal(self.wpe.weight, cfg.init_std)
        self.h = nn.ModuleList([GPT2Block(cfg) for _ in range(cfg.n_layer)])
        self.ln_f_w, self.ln_f_b = nn.Parameter(torch.ones(cfg.n_embd)), nn.Parameter(torch.zeros(cfg.n_embd))

    def forward(self, tokens, labels=None):
        B, S = tokens.shape
        pos = torch.arange(S, device=tokens.device)
        x = K.embedding(tokens, self.wte.weight) + self.wpe(pos)[None]
        for blk in self.h:
            if self.cfg.checkpoint_blocks and torch.is_grad_enabled():
                from torch.utils.checkpoint import checkpoint
                x = checkpoint(blk, x, use_reentrant=False)
            else:
                x = blk(x)
        x, _ = K.layer_norm(x, self.ln_f_w, self.ln_f_b, self.cfg.layer_norm_eps)
        logits = K.linear(x, self.wte.weight, None)      # tied LM head
        if labels is None:
            return logits
        return K.cross_entropy(logits, labels)
