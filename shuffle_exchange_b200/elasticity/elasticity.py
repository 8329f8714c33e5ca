"""Elastic batch-size planning: pick ONE global batch size that stays valid (``micro_batch x gas x gpus``) for as many
GPU counts as possible, so a job can be re-scheduled on a different number of GPUs without changing convergence.

Reference: ``deepspeed/elasticity/elasticity.py:27-378``.  Candidates are ``base x h`` for bases = every allowed
micro-batch and their LCM, ``h`` the largest highly-composite number with ``base x h <= max_train_batch_size``; the
candidate compatible with the most GPU counts in ``[min_gpus, max_gpus]`` wins (ties: larger/smaller batch by
``prefer_larger_batch``).  v0.2 plans at node granularity and accounts for model parallelism.  The highly composite
numbers are generated (records of the divisor-count function) instead of being tabulated.
"""
from __future__ import annotations

import json
import math
import os
from functools import lru_cache, reduce
from typing import List, Tuple

from ..utils.logging import logger
from ..version import __version__
from .config import ElasticityConfig, ElasticityConfigError, ElasticityError, ElasticityIncompatibleWorldSize
from .constants import (DEEPSPEED_ELASTICITY_CONFIG, ELASTICITY, ENABLED, ENABLED_DEFAULT, IGNORE_NON_ELASTIC_BATCH_INFO,
                        IGNORE_NON_ELASTIC_BATCH_INFO_DEFAULT, LATEST_ELASTICITY_VERSION, MINIMUM_DEEPSPEED_VERSION)


@lru_cache(maxsize=None)
def _highly_composite(limit: int) -> Tuple[int, ...]:
    """Numbers <= ``limit`` with more divisors than every smaller number (sieve of divisor counts)."""
    limit = max(int(limit), 1)
    counts = [0] * (limit + 1)
    for d in range(1, limit + 1):
        for m in range(d, limit + 1, d):
            counts[m] += 1
    out, best = [], 0
    for n in range(1, limit + 1):
        if counts[n] > best:
            best = counts[n]
            out.append(n)
    return tuple(out)


HCN_LIST = list(_highly_composite(720720))


def get_candidate_batch_sizes(base_list: List[int], max_acceptable_batch_size: int) -> List[int]:
    out = set()
    for base in base_list:
        if base >= max_acceptable_batch_size:
            out.add(base)
            continue
        budget = max_acceptable_batch_size // base
        h = max(x for x in HCN_LIST if x <= budget)
        out.add(h * base)
    cands = sorted(out)
    logger.info(f"Candidate batch size: {cands}")
    return cands


def get_valid_gpus(batch_size: int, micro_batches: List[int], min_valid_gpus: int, max_valid_gpus: int) -> List[int]:
    """GPU counts g for which some allowed micro-batch mb satisfies ``batch_size = mb * gas * g`` with integer gas."""
    valid = set()
    for mb in micro_batches:
        if batch_size % mb:
            continue
        slots = batch_size // mb            # = gas * g
        for g in range(max(1, min_valid_gpus), min(slots, max_valid_gpus) + 1):
            if slots % g == 0:
                valid.add(g)
    return sorted(valid)


def get_best_candidates(candidate_batch_sizes, micro_batches, min_gpus, max_gpus, prefer_larger):
    best_n, best_gpus, best_bs = 0, None, int(min(micro_batches))
    for bs in candidate_batch_sizes:
        gpus = get_valid_gpus(bs, micro_batches, min_gpus, max_gpus)
        better_tie = (prefer_larger and bs > best_bs) or (not prefer_larger and bs < best_bs)
        if len(gpus) > best_n or (len(gpus) == best_n and better_tie):
            best_n, best_gpus, best_bs = len(gpus), gpus, bs
    return best_bs, best_gpus


def _get_compatible_gpus_v01(micro_batches, max_acceptable_batch_size, min_gpus=None, max_gpus=None, prefer_larger=True):
    min_gpus = min_gpus or 1
    max_gpus = max_gpus or max_acceptable_batch_size // min(micro_batches)
    if not all(mb <= max_acceptable_batch_size for mb in micro_batches):
        raise ValueError(f"All micro batches must be less than or equal to max_acceptable_batch_size: {max_acceptable_batch_size}")
    lcm = reduce(lambda a, b: a * b // math.gcd(a, b), micro_batches)
    cands = get_candidate_batch_sizes(list(micro_batches) + [lcm], max_acceptable_batch_size)
    return get_best_candidates(cands, micro_batches, min_gpus, max_gpus, prefer_larger)


def _get_compatible_gpus_v02(micro_batches, max_acceptable_batch_size, current_num_gpus, min_gpus=None, max_gpus=None,
                             prefer_larger=True, num_gpus_per_node=1, model_parallel_size=1):
    if num_gpus_per_node % model_parallel_size != 0:
        raise ElasticityError(f"In Elasticity v0.2, number of GPUs per node:{num_gpus_per_node} should be divisible by "
                              f"model parallel size {model_parallel_size}")

    def pick_microbatch(batch):
        chosen = None
        for mb in micro_batches:
            if (batch // current_num_gpus) % mb == 0:
                if chosen is None or (prefer_larger and mb > chosen):
                    chosen = mb
        return chosen

    dp_per_node = num_gpus_per_node // model_parallel_size
    bs, valid_nodes = _get_compatible_gpus_v01(micro_batches, int(max_acceptable_batch_size / dp_per_node),
                                               int(min_gpus / num_gpus_per_node), int(max_gpus / num_gpus_per_node),
                                               prefer_larger=prefer_larger)
    bs = int(bs) * dp_per_node
    valid_dp = [n * dp_per_node for n in valid_nodes]
    if current_num_gpus // model_parallel_size in valid_dp:
        return bs, valid_dp, pick_microbatch(bs)
    # the current allocation is not elastic-compatible: fall back to the largest batch it can run by itself
    current_dp = (current_num_gpus / num_gpus_per_node) * dp_per_node
    cands = [math.floor(max_acceptable_batch_size / float(mb * current_dp)) * mb * current_dp for mb in micro_batches]
    bs = max(cands) if prefer_larger else min(cands)
    return bs, [int(current_dp)], pick_microbatch(bs)


def _version_tuple(v: str):
    return tuple(int(x) for x in str(v).split("+")[0].split(".")[:3] if x.isdigit())


def _compatible_ds_version_check(target_deepspeed_version: str) -> bool:
    if _version_tuple(target_deepspeed_version) < _version_tuple(MINIMUM_DEEPSPEED_VERSION) and \
            not str(target_deepspeed_version).startswith("0.1.0"):
        raise ElasticityError(f"Target deepspeed version of {target_deepspeed_version} is not compatible with minimum "
                              f"version {MINIMUM_DEEPSPEED_VERSION} supporting elasticity.")
    return True


def elasticity_enabled(ds_config: dict) -> bool:
    return bool(ds_config.get(ELASTICITY, {}).get(ENABLED, ENABLED_DEFAULT))


def ensure_immutable_elastic_config(runtime_elastic_config_dict: dict) -> None:
    """The resource scheduler and the runtime must have planned with the same elastic config."""
    if DEEPSPEED_ELASTICITY_CONFIG not in os.environ:
        logger.warning("Unable to find DEEPSPEED_ELASTICITY_CONFIG environment variable, cannot guarantee resource "
                       "scheduler will scale this job using compatible GPU counts.")
        return
    sched = ElasticityConfig(json.loads(os.environ[DEEPSPEED_ELASTICITY_CONFIG]))
    run = ElasticityConfig(runtime_elastic_config_dict)
    for attr in ("max_acceptable_batch_size", "micro_batches", "version"):
        if getattr(run, attr) != getattr(sched, attr):
            raise ElasticityConfigError(f"Elastic config '{attr}={getattr(sched, attr)}' seen by resource scheduler does not "
                                        f"match config passed to runtime {attr}={getattr(run, attr)}")


def compute_elastic_config(ds_config: dict, target_deepspeed_version: str, world_size=0, return_microbatch=False):
    """Returns ``(final_batch_size, valid_gpus[, micro_batch_size])`` for the ``elasticity`` block of ``ds_config``."""
    if not isinstance(ds_config, dict):
        raise ValueError(f"Expected ds_config to be a dictionary but received a {type(ds_config)}, containing: {ds_config}")
    if ELASTICITY not in ds_config:
        raise ElasticityConfigError(f"'{ELASTICITY}' is missing from config json, please add it if running an elastic training job.")
    block = ds_config[ELASTICITY]
    if not block.get(ENABLED, ENABLED_DEFAULT):
        raise ElasticityConfigError("Elasticity is disabled, please enable it ('enabled':true) if running an elastic training job.")
    cfg = ElasticityConfig(block)
    version = float(cfg.version)
    if cfg.model_parallel_size > 1 and version != 0.2:
        raise ElasticityConfigError(f"Elasticity V{cfg.version} does not support model-parallel training. Given model-parallel size: {cfg.model_parallel_size}")
    if version > LATEST_ELASTICITY_VERSION:
        raise ElasticityConfigError(f"Attempting to run elasticity version {cfg.version} but runtime only supports up to {LATEST_ELASTICITY_VERSION}")
    _compatible_ds_version_check(target_deepspeed_version)
    micro = None
    if version == 0.1:
        bs, gpus = _get_compatible_gpus_v01(cfg.micro_batches, cfg.max_acceptable_batch_size, cfg.min_gpus, cfg.max_gpus,
                                            cfg.prefer_larger_batch_size)
        bs = int(bs)
    elif version == 0.2:
        if world_size != 0:
            current = world_size
        elif os.getenv("WORLD_SIZE", "").isnumeric():
            current = int(os.environ["WORLD_SIZE"])
        else:
            raise ElasticityConfigError("Elasticity V 0.2 needs WORLD_SIZE to compute valid batch size. Either give it as "
                                        "argument to function compute_elastic_config or set it as an environment variable. "
                                        f"Value of WORLD_SIZE as environment variable is {os.getenv('WORLD_SIZE')}")
        bs, gpus, micro = _get_compatible_gpus_v02(cfg.micro_batches, cfg.max_acceptable_batch_size, current, cfg.min_gpus,
                                                   cfg.max_gpus, cfg.prefer_larger_batch_size, cfg.num_gpus_per_node,
                                                   cfg.model_parallel_size)
        bs = int(bs)
    else:
        raise NotImplementedError(f"Unable to find elastic logic for version: {cfg.version}")
    logger.info(f"Valid World Size (GPUs / Model Parallel Size): {gpus}")
    if world_size > 0:
        if world_size not in gpus:
            raise ElasticityIncompatibleWorldSize(f"World size ({world_size}) is not valid with the current list of valid GPU counts: {gpus}")
        if micro is None:
            for mb in sorted(set(cfg.micro_batches), reverse=True):
                if (bs // world_size) % mb == 0:
                    micro = mb
                    break
        assert micro is not None, f"Unable to find divisible micro batch size world_size={world_size}, final_batch_size={bs}, and  micro_batches={cfg.micro_batches}."
        return bs, gpus, micro
    if return_microbatch:
        return bs, gpus, micro
    return bs, gpus
