"""Helpers to adopt sparse attention in existing HuggingFace-style models
(reference deepspeed/ops/sparse_attention/sparse_attention_utils.py:19-225)."""
import torch
from torch.nn import functional as F

from .bert_sparse_self_attention import BertSparseSelfAttention
from .sparsity_config import SparsityConfig


class SparseAttentionUtils:
    @staticmethod
    def extend_position_embedding(model, max_position):
        """Tile the learned position embeddings up to `max_position` (bert / roberta layouts)."""
        if hasattr(model, "bert"):
            original_max_position = model.bert.embeddings.position_embeddings.weight.size(0)
            assert max_position > original_max_position
            extend_multiples = max(1, max_position // original_max_position)
            model.bert.embeddings.position_embeddings.weight.data = \
                model.bert.embeddings.position_embeddings.weight.repeat(extend_multiples, 1)
        elif hasattr(model, "roberta"):
            original_max_position, embed_size = model.roberta.embeddings.position_embeddings.weight.shape
            original_max_position -= 2      # roberta reserves two positions
            extend_multiples = max(1, max_position // original_max_position)
            assert max_position > original_max_position
            max_position += 2
            extended = model.roberta.embeddings.position_embeddings.weight.new_empty(max_position, embed_size)
            k = 2
            for i in range(extend_multiples):
                extended[k:(k + original_max_position)] = model.roberta.embeddings.position_embeddings.weight[2:]
                k += original_max_position
            model.roberta.embeddings.position_embeddings.weight.data = extended
        else:
            raise ValueError("Please extend \"extend_position_embedding\" function to support your model type. It "
                             "currently only supports \"bert\" & \"roberta\"!")
        model.config.max_position_embeddings = max_position
        print(f"Extended position embeddings to {original_max_position * extend_multiples}")
        return model

    @staticmethod
    def update_tokenizer_model_max_length(tokenizer, max_position):
        tokenizer.model_max_length = max_position
        tokenizer.init_kwargs["model_max_length"] = max_position
        print(f"updated tokenizer model max imum length to {max_position}")
        return tokenizer

    @staticmethod
    def replace_model_self_attention_with_sparse_self_attention(model, max_position,
                                                                sparsity_config=SparsityConfig(num_heads=4)):
        if hasattr(model, "bert"):
            model.config.max_position_embeddings = max_position
            SparseAttentionUtils.replace_self_attention_layer_with_sparse_self_attention_layer(
                model.config, model.bert.encoder.layer, sparsity_config)
        elif hasattr(model, "roberta"):
            model.config.max_position_embeddings = max_position + 2
            SparseAttentionUtils.replace_self_attention_layer_with_sparse_self_attention_layer(
                model.config, model.roberta.encoder.layer, sparsity_config)
        else:
            raise ValueError("Please extend \"update_model_self_attention_to_sparse_self_attention\" function to "
                             "support your model type. It currently only supports \"bert\" & \"roberta\"!")
        return model

    @staticmethod
    def replace_self_attention_layer_with_sparse_self_attention_layer(config, layers,
                                                                      sparsity_config=SparsityConfig(num_heads=4)):
        for layer in layers:
            deepspeed_sparse_self_attn = BertSparseSelfAttention(config, sparsity_config)
            deepspeed_sparse_self_attn.query = layer.attention.self.query
            deepspeed_sparse_self_attn.key = layer.attention.self.key
            deepspeed_sparse_self_attn.value = layer.attention.self.value
            layer.attention.self = deepspeed_sparse_self_attn
        return layers

    @staticmethod
    def pad_to_block_size(block_size, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds,
                          pad_token_id, model_embeddings):
        """Pad the sequence dimension to a multiple of the block size; returns pad_len and the padded tensors."""
        batch_size, seq_len = input_ids.shape if input_ids is not None else inputs_embeds.shape[:-1]
        pad_len = (block_size - seq_len % block_size) % block_size
        if pad_len > 0:
            if inputs_embeds is not None:
                pad_input_ids = inputs_embeds.new_full((batch_size, pad_len), pad_token_id, dtype=torch.long)
                pad_inputs_embeds = model_embeddings(pad_input_ids)
                inputs_embeds = torch.cat([inputs_embeds, pad_inputs_embeds], dim=-2)
            if input_ids is not None:
                input_ids = F.pad(input_ids, (0, pad_len), value=pad_token_id)
            if position_ids is not None:
                position_ids = F.pad(position_ids, (0, pad_len), value=pad_token_id)
            attention_mask = F.pad(attention_mask, (0, pad_len), value=False)
            token_type_ids = F.pad(token_type_ids, (0, pad_len), value=0)
        return pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds

    @staticmethod
    def unpad_sequence_output(pad_len, sequence_output):
        if pad_len > 0:
            sequence_output = sequence_output[:, :-pad_len]
        return sequence_output
