"""Inject DeepSpeedTransformerLayer into a HuggingFace-style encoder by layer path
(reference deepspeed/module_inject/inject.py)."""
import copy

import torch

from ..ops.transformer import DeepSpeedTransformerLayer, DeepSpeedTransformerConfig


def module_inject(layer_obj, model, config, micro_batch_size, max_seq_length, seed, preln, fp16=True):
    for name, child in model.named_children():
        if isinstance(child, layer_obj):
            cuda_config = DeepSpeedTransformerConfig(batch_size=micro_batch_size, max_seq_length=max_seq_length,
                                                     hidden_size=config.hidden_size, heads=config.num_attention_heads,
                                                     attn_dropout_ratio=config.attention_probs_dropout_prob,
                                                     hidden_dropout_ratio=config.hidden_dropout_prob,
                                                     num_hidden_layers=config.num_hidden_layers,
                                                     initializer_range=config.initializer_range, seed=seed, fp16=fp16,
                                                     pre_layer_norm=preln)
            new_module = DeepSpeedTransformerLayer(cuda_config)
            qw, qb = child.attention.self.query.weight, child.attention.self.query.bias
            kw, kb = child.attention.self.key.weight, child.attention.self.key.bias
            vw, vb = child.attention.self.value.weight, child.attention.self.value.bias
            with torch.no_grad():
                new_module.attn_qkvw.data = torch.cat((qw, kw, vw), 0).data.clone()
                new_module.attn_qkvb.data = torch.cat((qb, kb, vb), 0).data.clone()
                new_module.attn_ow.data = child.attention.output.dense.weight.data.clone()
                new_module.attn_ob.data = child.attention.output.dense.bias.data.clone()
                new_module.attn_nw.data = child.attention.output.LayerNorm.weight.data.clone()
                new_module.attn_nb.data = child.attention.output.LayerNorm.bias.data.clone()
                new_module.inter_w.data = child.intermediate.dense.weight.data.clone()
                new_module.inter_b.data = child.intermediate.dense.bias.data.clone()
                new_module.output_w.data = child.output.dense.weight.data.clone()
                new_module.output_b.data = child.output.dense.bias.data.clone()
                new_module.norm_w.data = child.output.LayerNorm.weight.data.clone()
                new_module.norm_b.data = child.output.LayerNorm.bias.data.clone()
            setattr(model, name, copy.deepcopy(new_module))
        else:
            module_inject(layer_obj, child, config, micro_batch_size, max_seq_length, seed, preln, fp16)
    return model
