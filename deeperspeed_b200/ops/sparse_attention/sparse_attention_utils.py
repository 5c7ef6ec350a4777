"""Helpers for switching an existing HuggingFace-style BERT / RoBERTa model to sparse self-attention.

Static-method surface and messages per reference deepspeed/ops/sparse_attention/sparse_attention_utils.py:19-225. The
two supported families differ only in where the encoder lives and in RoBERTa's two reserved position slots, so both go
through one small description table instead of parallel branches.
"""
import torch
from torch.nn import functional as F

from .bert_sparse_self_attention import BertSparseSelfAttention
from .sparsity_config import SparsityConfig

# model attribute -> number of position slots reserved before the first real position
_FAMILIES = {"bert": 0, "roberta": 2}


def _family(model, caller):
    for name, reserved in _FAMILIES.items():
        if hasattr(model, name):
            return getattr(model, name), reserved
    raise ValueError(f"Please extend \"{caller}\" function to support your model type. It currently only supports "
                     "\"bert\" & \"roberta\"!")


class SparseAttentionUtils:
    """Position-embedding extension, attention-layer replacement and block-size padding."""

    @staticmethod
    def extend_position_embedding(model, max_position):
        """Repeat the learned position table until it covers `max_position` positions (reserved slots kept once)."""
        body, reserved = _family(model, "extend_position_embedding")
        table = body.embeddings.position_embeddings.weight
        learned = table.size(0) - reserved
        assert max_position > learned
        copies = max(1, max_position // learned)
        if reserved == 0:
            new_table = table.repeat(copies, 1)
        else:
            # rows beyond copies * learned stay uninitialised, as in the reference implementation
            new_table = table.new_empty(max_position + reserved, table.size(1))
            new_table[:reserved] = table[:reserved]
            for i in range(copies):
                new_table[reserved + i * learned:reserved + (i + 1) * learned] = table[reserved:]
        body.embeddings.position_embeddings.weight.data = new_table
        model.config.max_position_embeddings = max_position + reserved
        print(f"Extended position embeddings to {learned * copies}")
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
        body, reserved = _family(model, "update_model_self_attention_to_sparse_self_attention")
        model.config.max_position_embeddings = max_position + reserved
        SparseAttentionUtils.replace_self_attention_layer_with_sparse_self_attention_layer(
            model.config, body.encoder.layer, sparsity_config)
        return model

    @staticmethod
    def replace_self_attention_layer_with_sparse_self_attention_layer(config, layers,
                                                                      sparsity_config=SparsityConfig(num_heads=4)):
        """Swap `layer.attention.self` for the block-sparse module, keeping the trained q / k / v projections."""
        for layer in layers:
            dense = layer.attention.self
            sparse = BertSparseSelfAttention(config, sparsity_config)
            sparse.query, sparse.key, sparse.value = dense.query, dense.key, dense.value
            layer.attention.self = sparse
        return layers

    @staticmethod
    def pad_to_block_size(block_size, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds,
                          pad_token_id, model_mbeddings):
        """Right-pad the sequence dimension to a multiple of `block_size` (padding is masked out of attention).
        Returns (pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds).
        `model_mbeddings` is the reference's spelling of the keyword (sparse_attention_utils.py:158)."""
        model_embeddings = model_mbeddings
        batch, seq_len = (input_ids.shape if input_ids is not None else inputs_embeds.shape[:-1])
        pad_len = -seq_len % block_size
        if pad_len == 0:
            return pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds

        def right_pad(t, value):
            return None if t is None else F.pad(t, (0, pad_len), value=value)

        if inputs_embeds is not None:
            filler = inputs_embeds.new_full((batch, pad_len), pad_token_id, dtype=torch.long)
            inputs_embeds = torch.cat([inputs_embeds, model_embeddings(filler)], dim=-2)
        input_ids = right_pad(input_ids, pad_token_id)
        position_ids = right_pad(position_ids, pad_token_id)
        attention_mask = right_pad(attention_mask, False)
        token_type_ids = right_pad(token_type_ids, 0)
        return pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds

    @staticmethod
    def unpad_sequence_output(pad_len, sequence_output):
        return sequence_output[:, :-pad_len] if pad_len > 0 else sequence_output
