"""Parse ``compression_training`` into a normalised dict (reference: ``deepspeed/compression/config.py:12-452``):
every technique gets ``shared_parameters`` (with defaults) and ``different_groups`` = {group name: {params, modules,
related_modules}}."""
from __future__ import annotations

from .constants import *  # noqa: F401,F403

_SHARED_DEFAULTS = {
    WEIGHT_QUANTIZATION: {TECHNIQUE_ENABLED: False, WEIGHT_QUANTIZE_KERNEL: False, TECHNIQUE_SCHEDULE_OFFSET: 0,
                          WEIGHT_QUANTIZE_GROUPS: 1, WEIGHT_QUANTIZE_VERBOSE: False, WEIGHT_QUANTIZE_TYPE: "symmetric",
                          WEIGHT_QUANTIZE_IN_FORWARD_ENABLED: False, WEIGHT_QUANTIZE_ROUNDING: "nearest",
                          WEIGHT_QUANTIZE_FP16_MIXED_QUANTIZE: {"enabled": False, WEIGHT_QUANTIZE_CHANGE_RATIO: 0.001}},
    ACTIVATION_QUANTIZATION: {TECHNIQUE_ENABLED: False, ACTIVATION_QUANTIZE_TYPE: "symmetric",
                              ACTIVATION_QUANTIZE_RANGE: "dynamic", TECHNIQUE_SCHEDULE_OFFSET: 1000},
    SPARSE_PRUNING: {TECHNIQUE_ENABLED: False, SPARSE_PRUNING_METHOD: "l1", TECHNIQUE_SCHEDULE_OFFSET: 1000,
                     TECHNIQUE_SCHEDULE_OFFSET_END: 1000, SPARSE_PRUNING_SCHEDULE_OFFSET_STRIDE: 1,
                     SPARSE_PRUNING_BLOCK_PATTERN: "4x1", SPARSE_PRUNING_DENSE_RATIO: 0.1, SPARSE_PRUNING_EXCLUDED_MODULES: []},
    ROW_PRUNING: {TECHNIQUE_ENABLED: False, ROW_PRUNING_METHOD: "l1", TECHNIQUE_SCHEDULE_OFFSET: 1000},
    HEAD_PRUNING: {TECHNIQUE_ENABLED: False, HEAD_PRUNING_METHOD: "topk", TECHNIQUE_SCHEDULE_OFFSET: 1000},
    CHANNEL_PRUNING: {TECHNIQUE_ENABLED: False, CHANNEL_PRUNING_METHOD: "l1", TECHNIQUE_SCHEDULE_OFFSET: 1000},
}
_REQUIRED_GROUP_PARAMS = {
    WEIGHT_QUANTIZATION: (WEIGHT_QUANTIZE_START_BITS, WEIGHT_QUANTIZE_TARGET_BITS),
    ACTIVATION_QUANTIZATION: (ACTIVATION_QUANTIZE_BITS,),
    SPARSE_PRUNING: (), ROW_PRUNING: (ROW_PRUNING_DENSE_RATIO,), HEAD_PRUNING: (HEAD_PRUNING_DENSE_RATIO,),
    CHANNEL_PRUNING: (CHANNEL_PRUNING_DENSE_RATIO,),
}


def get_layer_reduction(param_dict):
    lr = dict(param_dict.get(LAYER_REDUCTION, {}))
    lr.setdefault(LAYER_REDUCTION_ENABLED, LAYER_REDUCTION_ENABLED_DEFAULT)
    return lr


def _technique(sub, name):
    out = {SHARED_PARAMETERS: dict(_SHARED_DEFAULTS[name]), DIFFERENT_GROUPS: {}}
    block = sub.get(name, {})
    shared = block.get(SHARED_PARAMETERS, {})
    out[SHARED_PARAMETERS].update(shared)
    if name == HEAD_PRUNING and out[SHARED_PARAMETERS][TECHNIQUE_ENABLED]:
        assert HEAD_PRUNING_NUM_HEADS in shared, f"{HEAD_PRUNING_NUM_HEADS} must be specified for head pruning"
    if name == SPARSE_PRUNING and TECHNIQUE_SCHEDULE_OFFSET_END not in shared:
        out[SHARED_PARAMETERS][TECHNIQUE_SCHEDULE_OFFSET_END] = out[SHARED_PARAMETERS][TECHNIQUE_SCHEDULE_OFFSET]
    for gname, g in block.get(DIFFERENT_GROUPS, {}).items():
        params = dict(g.get(DIFFERENT_GROUPS_PARAMETERS, {}))
        for req in _REQUIRED_GROUP_PARAMS[name]:
            assert req in params, f"{req} must be specified for {name} group {gname}"
        if name == WEIGHT_QUANTIZATION:
            params.setdefault(WEIGHT_QUANTIZATION_PERIOD, WEIGHT_QUANTIZATION_PERIOD_DEFAULT)
        if name == SPARSE_PRUNING and out[SHARED_PARAMETERS][SPARSE_PRUNING_METHOD] != SPARSE_PRUNING_METHOD_SNIP_MOMENTUM:
            params.setdefault(SPARSE_PRUNING_DENSE_RATIO, out[SHARED_PARAMETERS][SPARSE_PRUNING_DENSE_RATIO])
        out[DIFFERENT_GROUPS][gname] = {
            DIFFERENT_GROUPS_PARAMETERS: params,
            DIFFERENT_GROUPS_MODULE_SCOPE: g.get(DIFFERENT_GROUPS_MODULE_SCOPE, DIFFERENT_GROUPS_MODULE_SCOPE_DEFAULT),
            DIFFERENT_GROUPS_RELATED_MODULE_SCOPE: g.get(DIFFERENT_GROUPS_RELATED_MODULE_SCOPE,
                                                         DIFFERENT_GROUPS_RELATED_MODULE_SCOPE_DEFAULT),
        }
    return out


def get_compression_config(param_dict):
    sub = param_dict.get(COMPRESSION_TRAINING, {})
    out = {name: _technique(sub, name) for name in TECHNIQUES}
    out[LAYER_REDUCTION] = get_layer_reduction(sub)
    return out


# ---- per-technique getters under the reference's names (``compression/config.py:60-450``); each takes the
# ``compression_training`` sub-dictionary and returns the section with defaults filled, as ``_technique`` does
def get_layer_reduction_enabled(param_dict):
    return param_dict.get(LAYER_REDUCTION, {}).get(LAYER_REDUCTION_ENABLED, LAYER_REDUCTION_ENABLED_DEFAULT) \
        if LAYER_REDUCTION in param_dict else False


def get_layer_reduction_params(param_dict):
    if LAYER_REDUCTION not in param_dict:
        return False
    params = dict(param_dict[LAYER_REDUCTION])
    params.pop(LAYER_REDUCTION_ENABLED, None)
    return params


def get_quantize_enabled(param_dict):
    return bool(param_dict.get(COMPRESSION_TRAINING, {}).get(WEIGHT_QUANTIZATION, {}).get(SHARED_PARAMETERS, {})
                .get(TECHNIQUE_ENABLED, False))


def _make_getters(name):
    def whole(param_dict): return _technique(param_dict, name)
    def shared(param_dict): return _technique(param_dict, name)[SHARED_PARAMETERS]
    def groups(param_dict): return _technique(param_dict, name)[DIFFERENT_GROUPS]
    return whole, shared, groups


(get_weight_quantization, get_weight_quantization_shared_parameters,
 get_weight_quantization_different_groups) = _make_getters(WEIGHT_QUANTIZATION)
(get_activation_quantization, get_activation_quantization_shared_parameters,
 get_activation_quantization_different_groups) = _make_getters(ACTIVATION_QUANTIZATION)
(get_sparse_pruning, get_sparse_pruning_shared_parameters,
 get_sparse_pruning_different_groups) = _make_getters(SPARSE_PRUNING)
(get_row_pruning, get_row_pruning_shared_parameters, get_row_pruning_different_groups) = _make_getters(ROW_PRUNING)
(get_head_pruning, get_head_pruning_shared_parameters, get_head_pruning_different_groups) = _make_getters(HEAD_PRUNING)
(get_channel_pruning, get_channel_pruning_shared_parameters,
 get_channel_pruning_different_groups) = _make_getters(CHANNEL_PRUNING)
