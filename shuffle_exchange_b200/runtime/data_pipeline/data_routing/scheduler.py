"""Kept-token-count schedule for random layer-token drop (reference: data_routing/scheduler.py:15-107).

``fixed_linear``: the number of tokens the wrapped layers keep grows linearly from ``min_value`` to ``max_value``
over ``require_steps`` steps in multiples of ``seq_per_step``; the scheduler also accounts the layer-tokens
consumed so far (used to scale the LR schedule by compute actually spent)."""
from __future__ import annotations

import math

from ..constants import *  # noqa: F401,F403


class BaseScheduler:
    def __init__(self):
        self.state = {}

    def _root_value(self, global_steps, root_degree=None):
        sc = self.state[RANDOM_LTD_SCHEDULE_CONFIG]
        degree = sc["root_degree"] if root_degree is None else root_degree
        lo, hi = self.state[RANDOM_LTD_MIN_VALUE], self.state[RANDOM_LTD_MAX_VALUE]
        frac = (float(global_steps) / sc[RANDOM_LTD_REQUIRE_STEP]) ** (1.0 / degree)
        seq = math.floor(frac * (hi - lo) + lo)
        seq -= seq % sc[RANDOM_LTD_INCREASE_STEP]
        return min(seq, hi)

    def get_value(self, global_steps):
        if self.state[RANDOM_LTD_SCHEDULER_TYPE] == "fixed_linear":
            return self._root_value(global_steps, 1)
        raise RuntimeError("Unsupported random LTD schedule type")


class RandomLTDScheduler(BaseScheduler):
    def __init__(self, config):
        super().__init__()
        self.model_layer_num = config[RANDOM_LTD_TOTAL_LAYER_NUM]
        self.random_ltd_layer_num = config[RANDOM_LTD_LAYER_NUM]
        self.config_schedule = config[RANDOM_LTD_SCHEDULER]
        self.global_batch_size = config[RANDOM_LTD_GLOBAL_BATCH_SIZE]
        self.reset_to_init()
        if config.get(RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE, {}).get(RANDOM_LTD_LAYER_TOKEN_LR_ENABLED, False):
            raise NotImplementedError("layer_token_lr_schedule is not implemented (nor in the reference)")
        self.state[RANDOM_LTD_CONSUMED_LAYER_TOKENS] = 0

    def reset_to_init(self):
        cs = self.config_schedule
        if cs is not None:
            for key in (RANDOM_LTD_MIN_VALUE, RANDOM_LTD_MAX_VALUE, RANDOM_LTD_SCHEDULE_CONFIG, RANDOM_LTD_SCHEDULER_TYPE):
                self.state[key] = cs[key]
            self.state[RANDOM_LTD_CURRENT_VALUE] = cs[RANDOM_LTD_MIN_VALUE]
        self.state[RANDOM_LTD_CONSUMED_LAYER_TOKENS] = 0
        self.state[RANDOM_LTD_CURR_STEP] = -1

    def get_total_layer_tokens(self, train_iters):
        for step in range(train_iters):
            self.update_seq(step)
        return self.state[RANDOM_LTD_CONSUMED_LAYER_TOKENS]

    def get_current_seq(self):
        return self.state[RANDOM_LTD_CURRENT_VALUE]

    def set_current_seq(self, seq_length):
        self.state[RANDOM_LTD_CURRENT_VALUE] = seq_length

    def get_random_ltd_layer_num(self):
        return self.random_ltd_layer_num

    def get_state(self):
        return self.state

    def set_state(self, state):
        self.state = state

    def update_seq(self, global_steps):
        st = self.state
        if st[RANDOM_LTD_CURRENT_VALUE] < st[RANDOM_LTD_MAX_VALUE]:
            st[RANDOM_LTD_CURRENT_VALUE] = self.get_value(global_steps)
        if global_steps != st[RANDOM_LTD_CURR_STEP]:
            dropped = st[RANDOM_LTD_CURRENT_VALUE] * self.random_ltd_layer_num
            full = st[RANDOM_LTD_MAX_VALUE] * (self.model_layer_num - self.random_ltd_layer_num)
            st[RANDOM_LTD_CONSUMED_LAYER_TOKENS] += self.global_batch_size * (dropped + full)
            st[RANDOM_LTD_CURR_STEP] = global_steps

    _SAVED = (RANDOM_LTD_CONSUMED_LAYER_TOKENS, RANDOM_LTD_CURR_STEP, RANDOM_LTD_CURRENT_VALUE, RANDOM_LTD_MIN_VALUE,
              RANDOM_LTD_MAX_VALUE)

    def state_dict(self):
        return {k: self.state[k] for k in self._SAVED}

    def load_state_dict(self, state_dict):
        for k in self._SAVED:
            self.state[k] = state_dict[k]
