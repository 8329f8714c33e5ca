"""Sampling helpers (reference: ``deepspeed/autotuning/tuner/utils.py``)."""
import numpy as np

from ..utils import dict_to_feature, flatten, get_all_configs, get_tuning_keys  # noqa: F401


def index_to_feature(p, dims):
    """Mixed-radix decomposition of a flat index."""
    feature = []
    for d in dims:
        feature.append(p % d)
        p //= d
    return feature


def feature_to_index(feature, dims):
    p, mult = 0, 1
    for f, d in zip(feature, dims):
        p += f * mult
        mult *= d
    return p


def gen_combinations(d: dict):
    return get_all_configs(d)


def dict_to_dims(tuning_space):
    dims = []
    for v in tuning_space.values():
        if isinstance(v, dict):
            dims.extend(dict_to_dims(v))
        elif isinstance(v, list):
            dims.append(len(v))
        else:
            dims.append(1)
    return dims
