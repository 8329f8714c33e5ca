"""Validated view of the ``"elasticity"`` config block.

The block is checked against a small declarative schema (key, default, predicate, complaint) instead of a chain of
``if`` statements; the attribute names and the exception types are those of the reference
(``deepspeed/elasticity/config.py``), so ``compute_elastic_config`` and user code are interchangeable."""
import json
from typing import Any, Callable, NamedTuple

from .constants import *  # noqa: F401,F403


class ElasticityError(Exception):
    """Base exception of the elasticity module."""


class ElasticityConfigError(ElasticityError):
    """The elasticity configuration is missing, disabled or inconsistent."""


class ElasticityIncompatibleWorldSize(ElasticityError):
    """The current world size is not among the GPU counts the elastic batch size supports."""


class _Field(NamedTuple):
    attr: str
    key: str
    default: Any
    ok: Callable[[Any], bool]
    complaint: str          # formatted with value=<what was given>


def _is_count(v) -> bool:
    return isinstance(v, int) and not isinstance(v, bool) and v >= 1


def _is_micro_batch_list(v) -> bool:
    return isinstance(v, list) and all(isinstance(m, int) and not isinstance(m, bool) and m > 0 for m in v)


_SCHEMA = (
    _Field("max_acceptable_batch_size", MAX_ACCEPTABLE_BATCH_SIZE, MAX_ACCEPTABLE_BATCH_SIZE_DEFAULT, lambda v: True, ""),
    _Field("micro_batches", MICRO_BATCHES, MICRO_BATCHES_DEFAULT, _is_micro_batch_list,
           f"{MICRO_BATCHES} must be a list of positive integers, got {{value!r}}"),
    _Field("min_gpus", MIN_GPUS, MIN_GPUS_DEFAULT, _is_count, f"{MIN_GPUS} must be a positive integer, got {{value!r}}"),
    _Field("max_gpus", MAX_GPUS, MAX_GPUS_DEFAULT, _is_count, f"{MAX_GPUS} must be a positive integer, got {{value!r}}"),
    _Field("model_parallel_size", MODEL_PARALLEL_SIZE, MODEL_PARALLEL_SIZE_DEFAULT, _is_count,
           f"{MODEL_PARALLEL_SIZE} cannot be less than 1, got {{value!r}}"),
    _Field("num_gpus_per_node", NUM_GPUS_PER_NODE, NUM_GPUS_PER_NODE_DEFAULT, _is_count,
           f"{NUM_GPUS_PER_NODE} cannot be less than 1, got {{value!r}}"),
    _Field("min_time", MIN_TIME, MIN_TIME_DEFAULT, lambda v: v >= 0, f"{MIN_TIME} needs to be >= 0, got {{value!r}}"),
    _Field("version", VERSION, VERSION_DEFAULT, lambda v: True, ""),
    _Field("prefer_larger_batch_size", PREFER_LARGER_BATCH, PREFER_LARGER_BATCH_DEFAULT, lambda v: True, ""),
    _Field("ignore_non_elastic_batch_info", IGNORE_NON_ELASTIC_BATCH_INFO, IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT,
           lambda v: True, ""),
)
_REQUIRED_WHEN_ENABLED = (MAX_ACCEPTABLE_BATCH_SIZE, MICRO_BATCHES)


class ElasticityConfig:
    def __init__(self, param_dict):
        self.enabled = param_dict.get(ENABLED, ENABLED_DEFAULT)
        if self.enabled:
            missing = [k for k in _REQUIRED_WHEN_ENABLED if k not in param_dict]
            if missing:
                raise ElasticityConfigError(f"Elasticity config missing {', '.join(missing)}")
        for f in _SCHEMA:
            value = param_dict.get(f.key, f.default)
            if not f.ok(value):
                raise ElasticityConfigError("Elasticity: " + f.complaint.format(value=value))
            setattr(self, f.attr, value)
        if self.max_gpus < self.min_gpus:
            raise ElasticityConfigError(f"Elasticity: {MIN_GPUS} ({self.min_gpus}) cannot be greater than {MAX_GPUS} "
                                        f"({self.max_gpus})")

    def repr(self):
        return self.__dict__

    def __repr__(self):
        return json.dumps(self.__dict__, sort_keys=True, indent=4)
