"""``RandomLayerTokenDrop``: run a transformer layer on a random, sorted subset of the tokens and scatter its
output back (reference: data_routing/basic_layer.py:14-113).  Layer 0 draws the indices for all wrapped layers
(one multinomial + one sort kernel) and parks them in the scheduler state; gather/scatter are the native
row-copy kernels behind ``ops.random_ltd``."""
from __future__ import annotations

import torch
from torch import Tensor
from torch.nn import Module

from ....ops.random_ltd import GatherTokens, ScatterTokens, bert_sample_tokens, gpt_sample_tokens
from ..constants import *  # noqa: F401,F403


class RandomLayerTokenDrop(Module):
    def __init__(self, layer: Module):
        super().__init__()
        self.random_ltd_layer = layer
        self.random_ltd_scheduler = None
        self.max_length = None
        self.reserved_length = -1
        self.curr_seq = -1
        self.batch_first = False

    def init_config(self, config, scheduler, random_ltd_layer_id):
        self.random_ltd_scheduler = scheduler
        self.random_ltd_layer_id = random_ltd_layer_id
        self.max_length = scheduler.state[RANDOM_LTD_MAX_VALUE]
        self.mask_name = config[RANDOM_LTD_MODEL_MASK_NAME]
        self.micro_bs = config[RANDOM_LTD_MICRO_BATCH_SIZE]
        self.random_ltd_num_layer = scheduler.random_ltd_layer_num
        order = config[RANDOM_LTD_HIDDEN_STATE_ORDER]
        if order not in ("batch_seq_dim", "seq_batch_dim"):
            raise NotImplementedError("hidden_state_order must be batch_seq_dim or seq_batch_dim")
        self.batch_first = order == "batch_seq_dim"
        self.model_type = config[RANDOM_LTD_MODEL_TYPE]
        if self.model_type not in ("encoder", "decoder"):
            raise NotImplementedError("model_type must be encoder or decoder")
        self.index_generator = bert_sample_tokens if self.model_type == "encoder" else gpt_sample_tokens

    def _shape(self, h):
        if self.batch_first:
            self.curr_micro_batch, self.curr_seq = h.size(0), h.size(1)
        else:
            self.curr_seq, self.curr_micro_batch = h.size(0), h.size(1)

    def forward(self, hidden_states, **kwargs) -> Tensor:
        sched = self.random_ltd_scheduler
        if sched is not None:
            self.reserved_length = sched.get_current_seq()
            self._shape(hidden_states)
        if not (self.training and sched is not None and self.reserved_length < self.curr_seq):
            return self.random_ltd_layer(hidden_states, **kwargs)
        mask = kwargs[self.mask_name] if self.mask_name is not None else None
        if self.random_ltd_layer_id == 0:
            idx, part_mask = self.index_generator(self.reserved_length, self.curr_seq, self.curr_micro_batch,
                                                  self.random_ltd_num_layer, hidden_states.device, mask)
            sched.state[RANDOM_LTD_SAMPLE_INDEX] = idx
            sched.state[RANDOM_LTD_ATTENTION_MASK] = part_mask
        else:
            idx = sched.state[RANDOM_LTD_SAMPLE_INDEX]
            part_mask = sched.state[RANDOM_LTD_ATTENTION_MASK]
        mine = idx[self.random_ltd_layer_id, :, :]
        hidden_states, part = GatherTokens.apply(hidden_states, mine, self.batch_first)
        if self.mask_name is not None:
            kwargs[self.mask_name] = part_mask[self.random_ltd_layer_id] if self.model_type == "encoder" else part_mask
        outputs = self.random_ltd_layer(part, **kwargs)
        if isinstance(outputs, tuple):
            merged = ScatterTokens.apply(hidden_states, outputs[0], mine, self.batch_first)
            return (merged,) + tuple(outputs[1:])
        if isinstance(outputs, Tensor):
            return ScatterTokens.apply(hidden_states, outputs, mine, self.batch_first)
        raise NotImplementedError("the wrapped layer must return a tensor or a tuple whose first item is one")
