"""GPT-NeoX-style model in plain PyTorch (stock torch ops only) on the REFERENCE's LayerSpec/PipelineModule.
Same architecture and sizes as deeperspeed_b200.models.gpt_neox; none of our kernels, models or engine is used."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F


class RefEmbedding(nn.Module):
    def __init__(self, vocab, hidden, std):
        super().__init__()
        self.word_embeddings = nn.Embedding(vocab, hidden)
        nn.init.normal_(self.word_embeddings.weight, std=std)

    def forward(self, tokens):
        return self.word_embeddings(tokens)


def _rotary(q, k, rot_dim, base):
    S = q.shape[1]
    half = rot_dim // 2
    inv_freq = 1.0 / (base ** (torch.arange(0, half, device=q.device, dtype=torch.float32) * 2 / rot_dim))
    ang = torch.arange(S, device=q.device, dtype=torch.float32)[:, None] * inv_freq[None]
    cs, sn = ang.cos()[None, :, None, :].to(q.dtype), ang.sin()[None, :, None, :].to(q.dtype)

    def rot(t):
        x0, x1, rest = t[..., :half], t[..., half:rot_dim], t[..., rot_dim:]
        return torch.cat((x0 * cs - x1 * sn, x1 * cs + x0 * sn, rest), dim=-1)

    return rot(q), rot(k)


class ParallelTransformerLayerPipe(nn.Module):
    def __init__(self, hidden, heads, num_layers, std, rotary_pct=1.0, base=10000.0, eps=1e-5):
        super().__init__()
        self.heads, self.head_dim = heads, hidden // heads
        self.rot_dim = int(self.head_dim * rotary_pct) // 2 * 2
        self.base = base
        out_std = std / math.sqrt(2.0 * num_layers)
        self.input_layernorm = nn.LayerNorm(hidden, eps=eps)
        self.query_key_value = nn.Linear(hidden, 3 * hidden)
        self.dense = nn.Linear(hidden, hidden)
        self.post_attention_layernorm = nn.LayerNorm(hidden, eps=eps)
        self.dense_h_to_4h = nn.Linear(hidden, 4 * hidden)
        self.dense_4h_to_h = nn.Linear(4 * hidden, hidden)
        for lin, s in ((self.query_key_value, std), (self.dense, out_std), (self.dense_h_to_4h, std),
                       (self.dense_4h_to_h, out_std)):
            nn.init.normal_(lin.weight, std=s)
            nn.init.zeros_(lin.bias)

    def forward(self, x):
        B, S, h = x.shape
        H, D = self.heads, self.head_dim
        qkv = self.query_key_value(self.input_layernorm(x)).view(B, S, 3, H, D)
        q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
        if self.rot_dim > 0:
            q, k = _rotary(q, k, self.rot_dim, self.base)
        ctx = F.scaled_dot_product_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), is_causal=True)
        x = x + self.dense(ctx.transpose(1, 2).reshape(B, S, h))
        y = self.post_attention_layernorm(x)
        return x + self.dense_4h_to_h(F.gelu(self.dense_h_to_4h(y), approximate="tanh"))


class RefNorm(nn.Module):
    def __init__(self, hidden, eps=1e-5):
        super().__init__()
        self.norm = nn.LayerNorm(hidden, eps=eps)

    def forward(self, x):
        return self.norm(x)


class RefLMHead(nn.Module):
    def __init__(self, vocab, hidden, std):
        super().__init__()
        self.lm_head = nn.Linear(hidden, vocab, bias=False)
        nn.init.normal_(self.lm_head.weight, std=std)

    def forward(self, x):
        return self.lm_head(x)


def ref_loss_fn(logits, labels):
    return F.cross_entropy(logits.float().view(-1, logits.shape[-1]), labels.view(-1))


def build_specs(LayerSpec, vocab=50432, hidden=2048, layers=24, heads=16, std=0.02):
    specs = [LayerSpec(RefEmbedding, vocab, hidden, std)]
    for _ in range(layers):
        specs.append(LayerSpec(ParallelTransformerLayerPipe, hidden, heads, layers, std))
    specs.append(LayerSpec(RefNorm, hidden))
    specs.append(LayerSpec(RefLMHead, vocab, hidden, std))
    return specs
