"""Config-space utilities (reference: ``deepspeed/autotuning/utils.py``)."""
from __future__ import annotations

import collections.abc
import copy
import itertools
import json
import os
import re
import subprocess

from .constants import *  # noqa: F401,F403


def search_error(filename):
    if not os.path.exists(filename):
        return "stderr.log does not exist"
    with open(filename) as f:
        for line in f:
            for s in ("Error", "error", "ERROR"):
                idx = line.find(s)
                if idx != -1:
                    return line[idx + len(s):].lstrip(": ")
    return None


def was_interruptted(filename):
    if not os.path.exists(filename):
        return "stderr.log does not exist"
    with open(filename) as f:
        return any(s in line for line in f for s in ("KeyboardInterrupt", "SIGINT"))


def find_replace_str(value, replace_dict):
    if not isinstance(value, str):
        return str(value)
    for m in re.findall(r"\$[\w]+", value):
        if m in replace_dict:
            value = value.replace(m, str(replace_dict[m]))
    return value


def find_replace(target, replace_dict):
    if isinstance(target, dict):
        for k, v in target.items():
            if isinstance(v, str):
                target[k] = find_replace_str(v, replace_dict)
            elif isinstance(v, list):
                target[k] = [find_replace_str(x, replace_dict) for x in v]
            elif isinstance(v, dict):
                find_replace(v, replace_dict)
    elif isinstance(target, list):
        for i, v in enumerate(target):
            target[i] = str(replace_dict.get(v, v))


def get_list(val):
    return val if isinstance(val, list) else [val]


def combine_dict(d, u):
    for k, v in u.items():
        if isinstance(v, collections.abc.Mapping):
            d[k] = combine_dict(d.get(k, {}), v)
        elif k not in d:
            d[k] = v
        else:
            d[k] = get_list(d[k]) + [x for x in get_list(v) if x not in get_list(d[k])]
    return d


def del_if_exists(t, d):
    if t in d:
        del d[t]
        return
    for v in d.values():
        if isinstance(v, collections.abc.Mapping):
            del_if_exists(t, v)


def replace_dict(d, u, ignored_keys=[]):
    """Overwrite values of ``d`` with those of ``u`` (recursively), skipping ``ignored_keys``."""
    if u is not None:
        for k, v in u.items():
            if k in ignored_keys:
                continue
            if v is None:
                del_if_exists(k, d)
            elif isinstance(v, collections.abc.Mapping):
                d[k] = replace_dict(d.get(k, {}), v, ignored_keys)
            else:
                d[k] = v
    return d


def get_val_by_key(d: dict, k):
    if k in d:
        return d[k]
    for v in d.values():
        if isinstance(v, dict):
            r = get_val_by_key(v, k)
            if r is not None:
                return r
    return None


def set_val_by_key(d: dict, k, vv):
    if k in d:
        d[k] = vv
    for v in d.values():
        if isinstance(v, dict):
            set_val_by_key(v, k, vv)


def fetch_hostfile(hostfile_path):
    from ..launcher.runner import fetch_hostfile as f
    return f(hostfile_path)


def validate_ds_config(config: dict):
    def is_false(cfg, key):
        return not cfg.get(key, None)
    z = config.get("zero_optimization", {})
    if not z:
        return True
    stage = z.get("stage")
    offload = False
    if stage == 1:
        return True
    if stage == 2:
        if z.get("cpu_offload") is True:
            offload = True
    elif stage == 3:
        if z.get("offload_optimizer", {}).get("device") == "nvme" or z.get("offload_param", {}).get("device") == "nvme":
            return "aio" in config
    return True


def remove_dupe_dicts(l):
    seen, out = set(), []
    for d in l:
        key = json.dumps(d, sort_keys=True)
        if key not in seen:
            seen.add(key)
            out.append(d)
    return out


def prune_config(config, ignored_keys=[]):
    if ignored_keys:
        for k in ignored_keys:
            config.pop(k, None)
        for v in config.values():
            if isinstance(v, dict):
                prune_config(v, ignored_keys)


def prune_configs(configs, ignored_keys=[]):
    pruned = []
    for c in configs:
        c = copy.deepcopy(c)
        prune_config(c, ignored_keys)
        pruned.append(c)
    return remove_dupe_dicts(pruned)


def get_tuning_keys(tuning_space: dict):
    """Keys whose value is a list with more than one choice."""
    keys = []
    for k, v in tuning_space.items():
        if isinstance(v, dict):
            keys.extend(get_tuning_keys(v))
        if isinstance(v, list) and len(v) > 1:
            keys.append(k)
    return keys


def flatten(d, parent_key="", sep="_"):
    items = []
    for k, v in d.items():
        nk = parent_key + sep + k if parent_key else k
        if isinstance(v, collections.abc.MutableMapping):
            items.extend(flatten(v, nk, sep=sep).items())
        else:
            items.append((nk, v))
    return dict(items)


def dict_to_feature(feature_dict, keys, max_value=None):
    feature = []
    for k, v in sorted(feature_dict.items()):
        if k in keys:
            if isinstance(v, bool):
                v = float(v)
            if isinstance(v, (int, float)):
                feature.append(float(v) / max_value[k] if max_value and max_value.get(k) else float(v))
    return feature


def get_all_configs(tuning_space: dict, ignore_keys=None):
    """Cartesian product over every list-valued entry of the (nested) tuning space."""
    def gen(d):
        keys, values = list(d.keys()), list(d.values())
        for v in values:
            if not isinstance(v, list):
                v = [v]
        expanded = [v if isinstance(v, list) else [v] for v in values]
        for combo in itertools.product(*expanded):
            yield dict(zip(keys, combo))

    all_configs = []
    space = copy.deepcopy(tuning_space)
    ignored = {}
    if ignore_keys:
        for ik in ignore_keys:
            ignored[ik] = space.pop(ik, {})
    for cfg in gen(space):
        nested = [(k, v) for k, v in cfg.items() if isinstance(v, dict)]
        if not nested:
            subs = [cfg]
        else:
            subs = [cfg]
            for k, v in nested:
                subs = [dict(s, **{k: sub}) for s in subs for sub in get_all_configs(v)]
        for s in subs:
            for ik, iv in ignored.items():
                if iv:
                    s[ik] = iv
            all_configs.append(s)
    return all_configs


def canonical_name(config: dict, tuning_keys=None, prefix="", omit_val=False):
    """Short experiment name, e.g. ``z2_gas1_tmbspg4_rbs5e8``."""
    if TRAIN_MICRO_BATCH_SIZE_PER_GPU not in (tuning_keys or []):
        tuning_keys = list(tuning_keys or []) + [TRAIN_MICRO_BATCH_SIZE_PER_GPU]
    if GRADIENT_ACCUMULATION_STEPS not in tuning_keys:
        tuning_keys = tuning_keys + [GRADIENT_ACCUMULATION_STEPS]
    tuning_keys = sorted(tuning_keys)

    def short(k):
        return "".join(w[0] for w in k.split("_"))

    def walk(cfg):
        parts = []
        for k, v in sorted(cfg.items()):
            if isinstance(v, dict):
                n = walk(v)
                if n:
                    parts.append(n)
            elif k in tuning_keys:
                parts.append(short(k) if omit_val else f"{short(k)}{v}")
        return "_".join(parts)

    name = walk(config)
    return f"{prefix}_{name}" if prefix else name


def get_first_config(config: dict):
    if not config:
        return None
    cfg = copy.deepcopy(config)
    for k, v in cfg.items():
        if isinstance(v, dict):
            cfg[k] = get_first_config(v)
        elif isinstance(v, list):
            cfg[k] = v[0]
    return cfg


def write_experiments(exps: list, exps_dir: str):
    paths = []
    for exp in exps:
        path = os.path.join(exps_dir, f"{exp['name']}.json")
        with open(path, "w") as fd:
            json.dump(exp, fd)
        paths.append(path)
    return paths


def memory_to_string(n, postfix="", units=None, precision=2):
    for scale, u in ((1024 ** 4, "T"), (1024 ** 3, "G"), (1024 ** 2, "M"), (1024, "K")):
        if (units is None and n >= scale) or units == u:
            return f"{round(n / scale, precision)} {u}{postfix}"
    return f"{n} {postfix}"


def number_to_string(n, postfix="", units=None, precision=2):
    for scale, u in ((1e9, "B"), (1e6, "M"), (1e3, "K")):
        if (units is None and n >= scale) or units == u:
            return f"{round(n / scale, precision)} {u}{postfix}"
    return f"{n} {postfix}"
