"""Normalise the ``"data_efficiency"`` block: fill every default so downstream code can index without ``.get``
(reference: ``runtime/data_pipeline/config.py:11-193``).

Layout::

    data_efficiency: {enabled, seed,
        data_sampling: {enabled, num_epochs, num_workers, pin_memory,
            curriculum_learning: {enabled, curriculum_metrics: {...}, ...},
            dynamic_batching: {enabled, lr_scaling_method, min_batch_size, max_batch_size, sequence_picking_order, ...}},
        data_routing: {enabled, random_ltd: {enabled, layer_token_lr_schedule: {enabled, ...}, ...}}}

plus the pre-data-efficiency top-level ``"curriculum_learning"`` section (``*_legacy`` accessors).
"""
from __future__ import annotations

import copy

from . import constants as K


def _section(d, key):
    v = d.get(key, {}) if d else {}
    return copy.copy(v) if isinstance(v, dict) else {}


def get_data_efficiency_enabled(param_dict) -> bool:
    return bool(_section(param_dict, K.DATA_EFFICIENCY).get(K.DATA_EFFICIENCY_ENABLED, K.DATA_EFFICIENCY_ENABLED_DEFAULT))


def get_data_efficiency_seed(param_dict) -> int:
    return _section(param_dict, K.DATA_EFFICIENCY).get(K.DATA_EFFICIENCY_SEED, K.DATA_EFFICIENCY_SEED_DEFAULT)


def get_data_sampling_enabled(param_dict) -> bool:
    return bool(_section(param_dict, K.DATA_SAMPLING).get(K.DATA_SAMPLING_ENABLED, K.DATA_SAMPLING_ENABLED_DEFAULT))


def get_data_sampling_num_epochs(param_dict) -> int:
    return _section(param_dict, K.DATA_SAMPLING).get(K.DATA_SAMPLING_NUM_EPOCHS, K.DATA_SAMPLING_NUM_EPOCHS_DEFAULT)


def get_data_sampling_num_workers(param_dict) -> int:
    return _section(param_dict, K.DATA_SAMPLING).get(K.DATA_SAMPLING_NUM_WORKERS, K.DATA_SAMPLING_NUM_WORKERS_DEFAULT)


def get_data_sampling_pin_memory(param_dict) -> bool:
    return _section(param_dict, K.DATA_SAMPLING).get(K.DATA_SAMPLING_PIN_MEMORY, K.DATA_SAMPLING_PIN_MEMORY_DEFAULT)


def get_curriculum_learning_enabled(param_dict) -> bool:
    return bool(_section(param_dict, K.CURRICULUM_LEARNING).get(K.CURRICULUM_LEARNING_ENABLED,
                                                                K.CURRICULUM_LEARNING_ENABLED_DEFAULT))


def get_curriculum_learning_params(param_dict) -> dict:
    params = _section(param_dict, K.CURRICULUM_LEARNING)
    params.pop(K.CURRICULUM_LEARNING_ENABLED, None)
    return params


def get_curriculum_learning(param_dict) -> dict:
    """``param_dict`` is the ``data_sampling`` section."""
    out = {K.CURRICULUM_LEARNING_ENABLED: get_curriculum_learning_enabled(param_dict)}
    if K.CURRICULUM_LEARNING not in (param_dict or {}):
        return out
    out.update(get_curriculum_learning_params(param_dict))
    if out[K.CURRICULUM_LEARNING_ENABLED] and K.CURRICULUM_LEARNING_METRICS not in out:
        raise ValueError(f"Curriculum learning is enabled, {K.CURRICULUM_LEARNING_METRICS} must be specified")
    return out


def get_dynamic_batching(param_dict) -> dict:
    out = _section(param_dict, K.DYNAMIC_BATCHING)
    out.setdefault(K.DYNAMIC_BATCHING_ENABLED, K.DYNAMIC_BATCHING_ENABLED_DEFAULT)
    out.setdefault(K.DYNAMIC_BATCHING_LR_SCALING_METHOD, K.DYNAMIC_BATCHING_LR_SCALING_METHOD_DEFAULT)
    out.setdefault(K.DYNAMIC_BATCHING_MIN_BATCH_SIZE, K.DYNAMIC_BATCHING_MIN_BATCH_SIZE_DEFAULT)
    out.setdefault(K.DYNAMIC_BATCHING_MAX_BATCH_SIZE, K.DYNAMIC_BATCHING_MAX_BATCH_SIZE_DEFAULT)
    out.setdefault(K.DYNAMIC_BATCHING_SEQUENCE_PICKING_ORDER, K.DYNAMIC_BATCHING_SEQUENCE_PICKING_ORDER_DEFAULT)
    out.setdefault(K.DYNAMIC_BATCHING_VERBOSE, False)
    if out[K.DYNAMIC_BATCHING_ENABLED] and K.DYNAMIC_BATCHING_MAX_TOKENS not in out:
        raise ValueError(f"dynamic batching is enabled, {K.DYNAMIC_BATCHING_MAX_TOKENS} must be specified")
    return out


def get_data_sampling(param_dict) -> dict:
    """``param_dict`` is the ``data_efficiency`` section."""
    sub = _section(param_dict, K.DATA_SAMPLING)
    wrapped = {K.DATA_SAMPLING: sub}
    out = {K.DATA_SAMPLING_ENABLED: get_data_sampling_enabled(wrapped),
           K.DATA_SAMPLING_NUM_EPOCHS: get_data_sampling_num_epochs(wrapped),
           K.DATA_SAMPLING_NUM_WORKERS: get_data_sampling_num_workers(wrapped),
           K.DATA_SAMPLING_PIN_MEMORY: get_data_sampling_pin_memory(wrapped)}
    out[K.CURRICULUM_LEARNING] = get_curriculum_learning(sub)
    out[K.DYNAMIC_BATCHING] = get_dynamic_batching(sub)
    return out


def get_random_ltd_enabled(param_dict) -> bool:
    return bool(_section(param_dict, K.RANDOM_LTD).get(K.RANDOM_LTD_ENABLED, K.RANDOM_LTD_ENABLED_DEFAULT))


def get_random_ltd_params(param_dict) -> dict:
    params = _section(param_dict, K.RANDOM_LTD)
    params.pop(K.RANDOM_LTD_ENABLED, None)
    return params


def get_random_ltd(param_dict) -> dict:
    """``param_dict`` is the ``data_routing`` section."""
    out = {K.RANDOM_LTD_ENABLED: get_random_ltd_enabled(param_dict),
           K.RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE: {K.RANDOM_LTD_LAYER_TOKEN_LR_ENABLED:
                                                  K.RANDOM_LTD_LAYER_TOKEN_LR_ENABLED_DEFAULT}}
    if get_random_ltd_enabled(param_dict):
        params = get_random_ltd_params(param_dict)
        sched = dict(params.pop(K.RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE, {}))
        sched.setdefault(K.RANDOM_LTD_LAYER_TOKEN_LR_ENABLED, K.RANDOM_LTD_LAYER_TOKEN_LR_ENABLED_DEFAULT)
        out.update(params)
        out[K.RANDOM_LTD_LAYER_TOKEN_LR_SCHEDULE] = sched
    return out


def get_data_routing_enabled(param_dict) -> bool:
    return bool(_section(param_dict, K.DATA_ROUTING).get(K.DATA_ROUTING_ENABLED, K.DATA_ROUTING_ENABLED_DEFAULT))


def get_data_routing(param_dict) -> dict:
    sub = _section(param_dict, K.DATA_ROUTING)
    return {K.DATA_ROUTING_ENABLED: get_data_routing_enabled(param_dict), K.RANDOM_LTD: get_random_ltd(sub)}


def get_data_efficiency_config(param_dict) -> dict:
    """Whole config dictionary in, fully defaulted ``data_efficiency`` block out."""
    sub = _section(param_dict, K.DATA_EFFICIENCY)
    return {K.DATA_EFFICIENCY_ENABLED: get_data_efficiency_enabled(param_dict),
            K.DATA_EFFICIENCY_SEED: get_data_efficiency_seed(param_dict),
            K.DATA_SAMPLING: get_data_sampling(sub), K.DATA_ROUTING: get_data_routing(sub)}


# ---- the older top-level "curriculum_learning" section
def get_curriculum_enabled_legacy(param_dict) -> bool:
    return bool(_section(param_dict, K.CURRICULUM_LEARNING_LEGACY).get(K.CURRICULUM_ENABLED_LEGACY,
                                                                       K.CURRICULUM_ENABLED_DEFAULT_LEGACY))


def get_curriculum_params_legacy(param_dict):
    if K.CURRICULUM_LEARNING_LEGACY not in (param_dict or {}):
        return False
    params = _section(param_dict, K.CURRICULUM_LEARNING_LEGACY)
    params.pop(K.CURRICULUM_ENABLED_LEGACY, None)
    return params
