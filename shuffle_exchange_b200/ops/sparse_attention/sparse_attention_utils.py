"""Helpers for retrofitting sparse attention into BERT/RoBERTa-style models.

Reference: ``deepspeed/ops/sparse_attention/sparse_attention_utils.py:14-208`` — position-embedding extension,
tokenizer max-length update, self-attention replacement, pad/unpad to a multiple of the block size."""
from __future__ import annotations

import torch
from torch.nn import functional as F

from .bert_sparse_self_attention import BertSparseSelfAttention
from .sparsity_config import SparsityConfig


class SparseAttentionUtils:

    @staticmethod
    def extend_position_embedding(model, max_position):
        """Tile the learned position table until it covers ``max_position`` (RoBERTa keeps its 2 reserved rows)."""
        if hasattr(model, 'bert'):
            emb = model.bert.embeddings.position_embeddings
            original = emb.weight.size(0)
            assert max_position > original
            reps = max(1, max_position // original)
            emb.weight.data = emb.weight.repeat(reps, 1)
        elif hasattr(model, 'roberta'):
            emb = model.roberta.embeddings.position_embeddings
            original, width = emb.weight.shape
            original -= 2
            reps = max(1, max_position // original)
            assert max_position > original
            max_position += 2
            table = emb.weight.new_empty(max_position, width)
            table[:2] = emb.weight[:2]
            body = emb.weight[2:].repeat(reps, 1)
            table[2:2 + body.shape[0]] = body[:max_position - 2]
            emb.weight.data = table
        else:
            raise ValueError('Please extend "extend_position_embedding" function to support your model type. It '
                             'currently only supports "bert" & "roberta"!')
        model.config.max_position_embeddings = max_position
        print(f'Extended position embeddings to {original * reps}')
        return model

    @staticmethod
    def update_tokenizer_model_max_length(tokenizer, max_position):
        tokenizer.model_max_length = max_position
        tokenizer.init_kwargs['model_max_length'] = max_position
        print(f'updated tokenizer model max imum length to {max_position}')
        return tokenizer

    @staticmethod
    def replace_model_self_attention_with_sparse_self_attention(model, max_position,
                                                                sparsity_config=SparsityConfig(num_heads=4)):
        if hasattr(model, 'bert'):
            model.config.max_position_embeddings = max_position
            layers = model.bert.encoder.layer
        elif hasattr(model, 'roberta'):
            model.config.max_position_embeddings = max_position + 2
            layers = model.roberta.encoder.layer
        else:
            raise ValueError('Please extend "update_model_self_attention_to_sparse_self_attention" function to '
                             'support your model type. It currently only supports "bert" & "roberta"!')
        SparseAttentionUtils.replace_self_attention_layer_with_sparse_self_attention_layer(model.config, layers,
                                                                                          sparsity_config)
        return model

    @staticmethod
    def replace_self_attention_layer_with_sparse_self_attention_layer(config, layers,
                                                                      sparsity_config=SparsityConfig(num_heads=4)):
        for layer in layers:
            sparse = BertSparseSelfAttention(config, sparsity_config)
            dense = layer.attention.self
            sparse.query, sparse.key, sparse.value = dense.query, dense.key, dense.value
            layer.attention.self = sparse
        return layers

    @staticmethod
    def pad_to_block_size(block_size, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds,
                          pad_token_id, model_embeddings):
        """Right-pad every model input so the sequence length is a multiple of ``block_size``; returns
        ``(pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds)``."""
        batch_size, seq_len = input_ids.shape if input_ids is not None else inputs_embeds.shape[:-1]
        pad_len = (block_size - seq_len % block_size) % block_size
        if pad_len > 0:
            if inputs_embeds is not None:
                pad_ids = inputs_embeds.new_full((batch_size, pad_len), pad_token_id, dtype=torch.long)
                inputs_embeds = torch.cat([inputs_embeds, model_embeddings(pad_ids)], dim=-2)
            if input_ids is not None:
                input_ids = F.pad(input_ids, (0, pad_len), value=pad_token_id)
            if position_ids is not None:
                position_ids = F.pad(position_ids, (0, pad_len), value=pad_token_id)
            attention_mask = F.pad(attention_mask, (0, pad_len), value=False)   # no attention on padding
            token_type_ids = F.pad(token_type_ids, (0, pad_len), value=0)
        return pad_len, input_ids, attention_mask, token_type_ids, position_ids, inputs_embeds

    @staticmethod
    def unpad_sequence_output(pad_len, sequence_output):
        if pad_len > 0:
            sequence_output = sequence_output[:, :-pad_len]
        return sequence_output
