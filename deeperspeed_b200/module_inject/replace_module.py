"""Swap HuggingFace-style BERT layers for DeepSpeedTransformerLayer (and back), copying q/k/v/o, LayerNorm and FFN
weights (reference deepspeed/module_inject/replace_module.py, deepspeed/ops/module_inject.py:8-216)."""
import copy

import torch

from ..ops.transformer import DeepSpeedTransformerLayer, DeepSpeedTransformerConfig


def _attr(obj, path):
    for p in path.split("."):
        obj = getattr(obj, p)
    return obj


def replace_transformer_layer(orig_layer_impl, model, micro_batch_size, bert_config, seed=-1, preln=True, fp16=True,
                              training=True, huggingface=False, local_rank=-1, max_seq_length=None):
    """Replace every `orig_layer_impl` instance in `model` by a DeepSpeedTransformerLayer carrying the same weights.
    Argument order and defaults of the reference (module_inject/replace_module.py:6-15); `max_seq_length` (an addition,
    keyword) defaults to the model's `max_position_embeddings`."""
    if max_seq_length is None:
        max_seq_length = getattr(bert_config, "max_position_embeddings", 512)
    cfg = DeepSpeedTransformerConfig(batch_size=micro_batch_size, max_seq_length=max_seq_length,
                                     hidden_size=bert_config.hidden_size, heads=bert_config.num_attention_heads,
                                     intermediate_size=getattr(bert_config, "intermediate_size", 4 * bert_config.hidden_size),
                                     attn_dropout_ratio=bert_config.attention_probs_dropout_prob,
                                     hidden_dropout_ratio=bert_config.hidden_dropout_prob,
                                     num_hidden_layers=bert_config.num_hidden_layers,
                                     initializer_range=bert_config.initializer_range, seed=seed, fp16=fp16,
                                     pre_layer_norm=preln, huggingface=huggingface, local_rank=local_rank,
                                     layer_norm_eps=getattr(bert_config, "layer_norm_eps", 1e-12))
    cfg.training = training
    return replace_layers_with_config(orig_layer_impl, model, cfg)


def replace_layers_with_config(orig_layer_impl, model, transformer_config):
    """The same replacement driven by a ready DeepSpeedTransformerConfig (what `deepspeed.ops.replace_module` takes,
    reference ops/module_inject.py:8-88)."""
    import copy
    preln = bool(transformer_config.pre_layer_norm)

    def replace_fn(child):
        cfg = copy.copy(transformer_config)          # every layer gets its own config object (it carries a layer id)
        new_module = DeepSpeedTransformerLayer(cfg)
        qw, qb = child.attention.self.query.weight, child.attention.self.query.bias
        kw, kb = child.attention.self.key.weight, child.attention.self.key.bias
        vw, vb = child.attention.self.value.weight, child.attention.self.value.bias
        with torch.no_grad():
            new_module.attn_qkvw.data = torch.cat((qw.data, kw.data, vw.data), 0).clone()
            new_module.attn_qkvb.data = torch.cat((qb.data, kb.data, vb.data), 0).clone()
            new_module.attn_ow.data = child.attention.output.dense.weight.data.clone()
            new_module.attn_ob.data = child.attention.output.dense.bias.data.clone()
            if preln:
                attention_layernorm = child.PostAttentionLayerNorm
                transformer_layernorm = child.PreAttentionLayerNorm
                inter = child.intermediate.dense_act
            else:
                attention_layernorm = child.attention.output.LayerNorm
                transformer_layernorm = child.output.LayerNorm
                inter = child.intermediate.dense
            new_module.attn_nw.data = attention_layernorm.weight.data.clone()
            new_module.attn_nb.data = attention_layernorm.bias.data.clone()
            new_module.inter_w.data = inter.weight.data.clone()
            new_module.inter_b.data = inter.bias.data.clone()
            new_module.output_w.data = child.output.dense.weight.data.clone()
            new_module.output_b.data = child.output.dense.bias.data.clone()
            new_module.norm_w.data = transformer_layernorm.weight.data.clone()
            new_module.norm_b.data = transformer_layernorm.bias.data.clone()
        return new_module.to(qw.device, qw.dtype)

    return replace_module(model=model, orig_class=orig_layer_impl, replace_fn=replace_fn)


def revert_transformer_layer(orig_layer_impl, model, bert_config, preln=False):
    """Inverse of replace_transformer_layer: rebuild `orig_layer_impl(bert_config)` layers from the fused weights."""

    def replace_fn(child):
        orig_module = orig_layer_impl(bert_config)
        qkvw, qkvb = child.attn_qkvw.data, child.attn_qkvb.data
        qw, kw, vw = torch.chunk(qkvw, 3, dim=0)
        qb, kb, vb = torch.chunk(qkvb, 3, dim=0)
        with torch.no_grad():
            orig_module.attention.self.query.weight.data = qw.clone()
            orig_module.attention.self.query.bias.data = qb.clone()
            orig_module.attention.self.key.weight.data = kw.clone()
            orig_module.attention.self.key.bias.data = kb.clone()
            orig_module.attention.self.value.weight.data = vw.clone()
            orig_module.attention.self.value.bias.data = vb.clone()
            orig_module.attention.output.dense.weight.data = child.attn_ow.data.clone()
            orig_module.attention.output.dense.bias.data = child.attn_ob.data.clone()
            if preln:
                attention_layernorm = orig_module.PostAttentionLayerNorm
                transformer_layernorm = orig_module.PreAttentionLayerNorm
                inter = orig_module.intermediate.dense_act
            else:
                attention_layernorm = orig_module.attention.output.LayerNorm
                transformer_layernorm = orig_module.output.LayerNorm
                inter = orig_module.intermediate.dense
            attention_layernorm.weight.data = child.attn_nw.data.clone()
            attention_layernorm.bias.data = child.attn_nb.data.clone()
            inter.weight.data = child.inter_w.data.clone()
            inter.bias.data = child.inter_b.data.clone()
            orig_module.output.dense.weight.data = child.output_w.data.clone()
            orig_module.output.dense.bias.data = child.output_b.data.clone()
            transformer_layernorm.weight.data = child.norm_w.data.clone()
            transformer_layernorm.bias.data = child.norm_b.data.clone()
        return orig_module.to(qkvw.device, qkvw.dtype)

    return replace_module(model=model, orig_class=DeepSpeedTransformerLayer, replace_fn=replace_fn)


def replace_module(model, orig_class, replace_fn):
    """Recursively replace all instances of `orig_class` using `replace_fn(child) -> new module`."""
    policy = {orig_class: replace_fn}
    return _replace_module(model, policy)


def _replace_module(model, policies):
    for name, child in model.named_children():
        if child.__class__ in policies:
            setattr(model, name, policies[child.__class__](child))
        else:
            _replace_module(child, policies)
    return model
