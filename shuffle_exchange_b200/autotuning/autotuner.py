"""Autotuner: explore ZeRO stage x micro-batch size x stage-specific knobs by running short profiling jobs and keep the
fastest configuration.

Reference: ``deepspeed/autotuning/autotuner.py:42-1110``.  Flow kept: (1) a *model-info* run (ZeRO-3, micro-batch 1)
records parameter count and activation memory; (2) for each ZeRO stage whose estimated state memory fits, find the
micro-batch range that fits and measure a few sizes; (3) tune the stage's template space with the chosen tuner at the
best micro-batch sizes; (4) write ``ds_config_optimal.json`` + ``cmd_optimal.txt``.
"""
from __future__ import annotations

import copy
import datetime
import json
import math
import os
import shutil

import hjson

from ..runtime.zero.config import ZeroStageEnum
from ..utils.logging import logger
from .config import DeepSpeedAutotuningConfig
from .constants import *  # noqa: F401,F403
from .scheduler import ResourceManager
from .tuner import GridSearchTuner, ModelBasedTuner, RandomTuner
from .utils import (canonical_name, combine_dict, get_all_configs, get_first_config, get_tuning_keys, memory_to_string,
                    number_to_string, replace_dict, write_experiments)

ZERO_OPTIMIZATION_DISABLED, ZERO_OPTIMIZATION_OPTIMIZER_STATES, ZERO_OPTIMIZATION_GRADIENTS, ZERO_OPTIMIZATION_WEIGHTS = 0, 1, 2, 3


class Autotuner:
    def __init__(self, args, active_resources):
        self.args = args
        self.selected_exp_dir = None
        logger.debug(f"autotuning args={args}")
        self.user_config = self._get_user_config(args.user_args)
        assert self.user_config is not None, "DeepSpeed configuration is not provided"
        self.autotuning_config = DeepSpeedAutotuningConfig(self.user_config)
        if self.user_config.get(AUTOTUNING):
            self.user_config[AUTOTUNING].setdefault(AUTOTUNING_EXPS_DIR, self.autotuning_config.exps_dir)
            self.user_config[AUTOTUNING].setdefault(AUTOTUNING_RESULTS_DIR, self.autotuning_config.results_dir)
        self.exps_dir, self.results_dir = self.autotuning_config.exps_dir, self.autotuning_config.results_dir
        for d in (self.exps_dir, self.results_dir):
            if self.autotuning_config.overwrite and os.path.exists(d):
                shutil.rmtree(d, ignore_errors=True)
            os.makedirs(d, exist_ok=True)
        self.start_profile_step, self.end_profile_step = self.autotuning_config.start_profile_step, self.autotuning_config.end_profile_step
        self.metric_path = self.autotuning_config.metric_path
        self.model_info_path = self.autotuning_config.model_info_path
        self.rm = self._get_resource_manager(active_resources)
        self.exp_num_nodes, self.exp_num_gpus = len(self.rm.nodes), self.rm.nodes[0].max_slots if self.rm.nodes else 0
        assert self.exp_num_gpus > 0 and self.exp_num_nodes > 0
        self.records = {}
        self.optimal_cmd = None
        self.optimal_ds_config = None
        self.mlflow_parent_id = None
        self.model_info = None
        self.gpu_mem = self.get_gpu_memory_info()
        self.activation_mem = None
        self.best_exp_info = None

    # ------------------------------------------------------------------------- configuration
    def _get_user_config(self, user_args):
        for flag in ("--deepspeed_config", "--deepspeed"):
            if flag in user_args:
                idx = user_args.index(flag)
                if idx + 1 < len(user_args) and not user_args[idx + 1].startswith("--"):
                    path = user_args[idx + 1]
                    assert os.path.isfile(path), f"DeepSpeed configuration file: {path} is not an existing file"
                    with open(path) as f:
                        return hjson.load(f)
        return None

    def _get_resource_manager(self, active_resources):
        hosts, ngpus = [], []
        for host, slots in active_resources.items():
            hosts.append(host)
            ngpus.append(len(slots))
        assert len(set(ngpus)) == 1, "Autotuning requires the same number of GPUs on every node"
        return ResourceManager(args=self.args, hosts=hosts, num_gpus_per_node=ngpus[0], results_dir=self.results_dir,
                               exps_dir=self.exps_dir, arg_mappings=self.autotuning_config.arg_mappings)

    def metric(self): return self.autotuning_config.metric
    def fast_enabled(self): return self.autotuning_config.fast
    def max_train_batch_size(self): return self.autotuning_config.max_train_batch_size
    def mp_size(self): return self.autotuning_config.mp_size
    def min_train_micro_batch_size_per_gpu(self): return self.autotuning_config.min_train_micro_batch_size_per_gpu
    def num_tuning_micro_batch_sizes(self): return self.autotuning_config.num_tuning_micro_batch_sizes

    def max_train_micro_batch_size_per_gpu(self):
        mtbs = self.max_train_batch_size()
        if mtbs and mtbs > 0:
            cap = mtbs * self.mp_size() // (self.exp_num_gpus * self.exp_num_nodes)
            return min(self.autotuning_config.max_train_micro_batch_size_per_gpu, cap)
        return self.autotuning_config.max_train_micro_batch_size_per_gpu

    def fp16_enabled(self):
        return bool(self.user_config.get("fp16", {}).get("enabled", False) or self.user_config.get("bf16", {}).get("enabled", False))

    def get_gpu_memory_info(self):
        import torch
        if torch.cuda.is_available():
            return torch.cuda.get_device_properties(0).total_memory
        return 180 * 1024 ** 3

    def get_activation_memory_per_gpu(self):
        if self.model_info and "activation_mem_per_gpu" in self.model_info:
            return self.model_info["activation_mem_per_gpu"]
        return None

    def get_model_num_params(self):
        if self.model_info and "num_params" in self.model_info:
            return self.model_info["num_params"]
        return None

    def get_instantiation_memory_required_per_gpu(self, zero_stage):
        """Model states per GPU: params + grads in the compute dtype, fp32 master + Adam moments, divided by the dp degree
        for whatever the stage shards."""
        n = self.get_model_num_params()
        total_gpus = self.exp_num_nodes * self.exp_num_gpus
        if n is None:
            return 0
        half = self.fp16_enabled()
        params_mem = n * (2 if half else 4)
        grads_mem = n * (2 if half else 4)
        optim_mem = n * (12 if half else 8)
        dp = max(1, total_gpus // self.mp_size())
        if zero_stage >= ZERO_OPTIMIZATION_OPTIMIZER_STATES:
            optim_mem /= dp
        if zero_stage >= ZERO_OPTIMIZATION_GRADIENTS:
            grads_mem /= dp
        if zero_stage >= ZERO_OPTIMIZATION_WEIGHTS:
            params_mem /= dp
        return (params_mem + grads_mem + optim_mem) / self.mp_size()

    # ------------------------------------------------------------------------------ experiments
    def _generate_experiments(self, tuning_space, max_train_batch_size_per_gpu):
        exps = []
        configs = get_all_configs(tuning_space, ignore_keys=["optimizer"])
        keys = get_tuning_keys(tuning_space)
        stage = tuning_space.get(ZERO_OPTIMIZATION, {}).get(ZERO_OPTIMIZATION_STAGE, 0)
        for cfg in configs:
            ds = copy.deepcopy(self.user_config)
            replace_dict(ds, cfg, ignored_keys=[])
            mbs = ds.get(TRAIN_MICRO_BATCH_SIZE_PER_GPU, 1)
            gas = ds.get(GRADIENT_ACCUMULATION_STEPS, 1)
            gas = 1 if gas == "auto" else gas
            ds[GRADIENT_ACCUMULATION_STEPS] = gas
            ds[TRAIN_BATCH_SIZE] = mbs * gas * self.exp_num_gpus * self.exp_num_nodes // self.mp_size()
            if max_train_batch_size_per_gpu and mbs * gas > max_train_batch_size_per_gpu:
                continue
            ds.setdefault(AUTOTUNING, {})
            ds[AUTOTUNING].update({AUTOTUNING_ENABLED: True, AUTOTUNING_START_PROFILE_STEP: self.start_profile_step,
                                   AUTOTUNING_END_PROFILE_STEP: self.end_profile_step})
            ds["flops_profiler"] = {"enabled": False}
            exps.append({"name": canonical_name(ds, keys, prefix=f"z{stage}"), "ds_config": ds,
                         "num_gpus": self.exp_num_gpus, "num_nodes": self.exp_num_nodes})
        return exps

    def model_info_profile_run(self):
        """One short ZeRO-3 job that records the model's parameter count and activation memory."""
        if self.model_info_path and os.path.exists(self.model_info_path):
            with open(self.model_info_path) as f:
                return json.load(f)
        info_cfg = self.autotuning_config.model_info
        if info_cfg and not info_cfg.get(MODEL_INFO_PROFILE, False) and info_cfg.get(MODEL_INFO_NUM_PARAMS):
            return {"num_params": info_cfg[MODEL_INFO_NUM_PARAMS]}
        ds = copy.deepcopy(self.user_config)
        replace_dict(ds, DEFAULT_MIN_MEM_CONFIG)
        path = os.path.join(self.results_dir, "profile_model_info", "model_info.json")
        ds.setdefault(AUTOTUNING, {})
        ds[AUTOTUNING].update({"enabled": True, "model_info_path": path, "model_info": {"profile": True}})
        exp = {"name": "profile_model_info", "ds_config": ds, "num_gpus": self.exp_num_gpus, "num_nodes": self.exp_num_nodes}
        paths = write_experiments([exp], self.exps_dir)
        self.rm.schedule_experiments(paths)
        self.rm.run()
        for exp_id, (e, err) in self.rm.finished_experiments.items():
            if err:
                logger.error(f"The model is not runnable with DeepSpeed with error = {err}")
                self.rm.clear()
                return None
        self.rm.clear()
        if os.path.exists(path):
            with open(path) as f:
                return json.load(f)
        return None

    def get_tuning_micro_batch_size_list(self, min_mbs, max_mbs, num=3):
        """Up to ``num`` sizes between min and max, preferring ones that divide ``max_train_batch_size``."""
        if min_mbs <= 0 or max_mbs < min_mbs:
            return []
        if max_mbs == min_mbs or num == 1:
            return [min_mbs]
        stride = max(1, (max_mbs - min_mbs) // max(1, num - 1))
        sizes = sorted({min(max_mbs, min_mbs + i * stride) for i in range(num)} | {max_mbs})
        mtbs = self.max_train_batch_size()
        if mtbs:
            world = self.exp_num_gpus * self.exp_num_nodes // self.mp_size()
            ok = [m for m in sizes if mtbs // world >= m]
            sizes = ok or sizes[:1]
        return sizes[-num:]

    def run_ds_config(self, ds_config, exp_name):
        exp = {"name": exp_name, "ds_config": ds_config, "num_gpus": self.exp_num_gpus, "num_nodes": self.exp_num_nodes}
        paths = write_experiments([exp], self.exps_dir)
        self.rm.schedule_experiments(paths)
        self.rm.run()
        exp, metric_val = self.rm.parse_results(self.metric())
        self.rm.clear()
        return exp, metric_val

    def update_records(self, space_name, exp, metric_val, num_exps):
        rec = self.records.setdefault(space_name, [])
        rec.append((exp, metric_val, num_exps))

    def get_best_space_record(self, space_name):
        best = None
        for exp, val, n in self.records.get(space_name, []):
            if val is not None and (best is None or val > best[1]):
                best = (exp, val, n)
        if best is not None:
            total = sum(n for _, _, n in self.records[space_name])
            best = (best[0], best[1], total)
        return best

    def get_best_space_records(self):
        out, glob = {}, None
        for name in self.records:
            b = self.get_best_space_record(name)
            if b:
                out[name] = b
                if glob is None or b[1] > glob[1]:
                    glob = b
        if glob:
            out[GLOBAL_TUNING_SPACE] = glob
        return out

    # ------------------------------------------------------------------- micro-batch-size search
    def get_gas_from_user_config(self) -> int:
        gas = self.user_config.get(GRADIENT_ACCUMULATION_STEPS, 1)
        if isinstance(gas, int):
            return max(1, gas)
        if gas == "auto":                     # the launcher script decides: read it from the mapped user argument
            v = self.get_val_from_user_args(GRADIENT_ACCUMULATION_STEPS)
            return max(1, int(v)) if v is not None else 1
        raise ValueError(f"gradient_accumulation_steps must be an int or 'auto', got {gas!r}")

    def get_val_from_user_args(self, ds_name: str):
        """Value the user's command line gives to the script argument mapped (``arg_mappings``) to a config key."""
        mappings = self.autotuning_config.arg_mappings or {}
        arg_name = mappings.get(ds_name)
        user_args = list(getattr(self.args, "user_args", []) or [])
        if not arg_name:
            return None
        for i, tok in enumerate(user_args):
            if tok == arg_name and i + 1 < len(user_args):
                return user_args[i + 1]
            if tok.startswith(arg_name + "="):
                return tok.split("=", 1)[1]
        return None

    def _mbs_config(self, stage: int, mbs: int, gas: int) -> dict:
        cfg = copy.deepcopy(self.user_config)
        z = dict(cfg.get(ZERO_OPTIMIZATION, {}) or {})
        z[ZERO_OPTIMIZATION_STAGE] = stage
        cfg[ZERO_OPTIMIZATION] = z
        cfg[TRAIN_MICRO_BATCH_SIZE_PER_GPU] = mbs
        cfg[GRADIENT_ACCUMULATION_STEPS] = gas
        world = max(1, self.exp_num_gpus * self.exp_num_nodes // self.mp_size())
        cfg[TRAIN_BATCH_SIZE] = mbs * gas * world
        return cfg

    def run_tuning_micro_batch_sizes(self, tuning_micro_batch_sizes, max_train_batch_size_per_gpu, min_gas, stage,
                                     tuning_space_name=None):
        """Run one experiment per candidate micro-batch size (ascending); stops at the first size that fails or that
        would exceed ``max_train_batch_size_per_gpu``.  Returns the largest size that ran."""
        name = tuning_space_name or f"{TUNING_MICRO_BATCH_SIZE_PREFIX}{stage}"
        largest = 0
        for mbs in sorted(set(tuning_micro_batch_sizes)):
            gas = min_gas
            if max_train_batch_size_per_gpu and mbs * gas > max_train_batch_size_per_gpu:
                break
            exp, metric_val = self.run_ds_config(self._mbs_config(stage, mbs, gas), f"z{stage}_gas{gas}_tmbspg{mbs}")
            if not metric_val:
                logger.info(f"micro batch size {mbs} did not run at ZeRO stage {stage}")
                break
            self.update_records(name, exp, metric_val, 1)
            largest = mbs
        return largest

    def get_min_max_micro_batch_size(self, stage, min_micro_batch_size, calculated_max_micro_batch_size,
                                     max_train_batch_size_per_gpu=None):
        """Smallest and largest micro-batch size that actually run: the minimum is tried first (``-1, -1`` if even that
        fails), then the estimate from the memory model; if the estimate fails the answer is found by bisection, if it
        succeeds sizes keep doubling until one fails or the batch-size cap is hit."""
        name = f"{TUNING_MICRO_BATCH_SIZE_PREFIX}{stage}"
        gas = self.get_gas_from_user_config()
        cap = max_train_batch_size_per_gpu or float("inf")
        ran = {}

        def runs(mbs: int) -> bool:
            if mbs not in ran:
                if mbs * gas > cap:
                    ran[mbs] = False
                else:
                    exp, val = self.run_ds_config(self._mbs_config(stage, mbs, gas), f"z{stage}_gas{gas}_tmbspg{mbs}")
                    ran[mbs] = bool(val)
                    if val:
                        self.update_records(name, exp, val, 1)
            return ran[mbs]

        lo = max(1, int(min_micro_batch_size))
        if not runs(lo):
            return -1, -1
        hi = max(lo, int(calculated_max_micro_batch_size))
        if hi > lo and runs(hi):
            good = hi
            while good * 2 * gas <= cap and runs(good * 2):
                good *= 2
            return lo, good
        good, bad = lo, hi
        while bad - good > 1:
            mid = (good + bad) // 2
            if runs(mid):
                good = mid
            else:
                bad = mid
        return lo, good

    def get_plateau_mbs(self, tuning_space_name):
        """Micro-batch size after which the metric stops improving by more than ``METRIC_PERCENT_DIFF_CONST``."""
        recs = sorted(((exp["ds_config"][TRAIN_MICRO_BATCH_SIZE_PER_GPU], val)
                       for exp, val, _ in self.records.get(tuning_space_name, []) if val), key=lambda t: t[0])
        if not recs:
            return 0
        plateau, prev = recs[0][0], recs[0][1]
        for mbs, val in recs[1:]:
            if val < prev * (1 + METRIC_PERCENT_DIFF_CONST):
                break
            plateau, prev = mbs, val
        return plateau

    # ------------------------------------------------------------------------------------ tuning
    def tune_space(self, tuning_space, prev_max_mbs=0, prev_best_mbs=0, prev_best_metric_val=0):
        stage = tuning_space.get(ZERO_OPTIMIZATION, {}).get(ZERO_OPTIMIZATION_STAGE, 0)
        name = f"{TUNING_MICRO_BATCH_SIZE_PREFIX}{stage}"
        need = self.get_instantiation_memory_required_per_gpu(stage)
        act = self.get_activation_memory_per_gpu() or 0
        if need and need + act > self.gpu_mem:
            logger.info(f"ZeRO stage {stage}: estimated {memory_to_string(need + act)}B per GPU exceeds device memory; skipped")
            return 0, 0, 0
        max_mbs = self.max_train_micro_batch_size_per_gpu()
        if act:
            max_mbs = max(1, min(max_mbs, int((self.gpu_mem - need) // act)))
        min_mbs = max(self.min_train_micro_batch_size_per_gpu(), 1)
        user_mbs = self.user_config.get(TRAIN_MICRO_BATCH_SIZE_PER_GPU)
        if isinstance(user_mbs, int):
            mbs_list = [user_mbs]
        else:
            mbs_list = self.get_tuning_micro_batch_size_list(min_mbs, max_mbs, self.num_tuning_micro_batch_sizes())
        space = copy.deepcopy(tuning_space)
        space[TRAIN_MICRO_BATCH_SIZE_PER_GPU] = mbs_list
        exps = self._generate_experiments(space, None)
        if not exps:
            return 0, 0, 0
        ttype = self.autotuning_config.tuner_type
        if ttype == AUTOTUNING_TUNER_MODELBASED:
            tuner = ModelBasedTuner(exps, self.rm, self.metric(), space)
        elif ttype == AUTOTUNING_TUNER_RANDOM:
            tuner = RandomTuner(exps, self.rm, self.metric())
        else:
            tuner = GridSearchTuner(exps, self.rm, self.metric())
        sample = max(1, len(self.rm.nodes) * self.rm.nodes[0].max_slots // (self.exp_num_gpus * self.exp_num_nodes))
        n = tuner.tune(sample_size=sample, n_trials=self.autotuning_config.tuner_num_trials,
                       early_stopping=self.autotuning_config.tuner_early_stopping)
        if tuner.best_exp:
            self.update_records(name, tuner.best_exp, tuner.best_metric_val, n)
            best_mbs = tuner.best_exp["ds_config"].get(TRAIN_MICRO_BATCH_SIZE_PER_GPU, 0)
            return max(mbs_list), best_mbs, tuner.best_metric_val or 0
        return 0, 0, 0

    def tune(self):
        if self.autotuning_config.enabled is False and not getattr(self.args, "autotuning", ""):
            logger.info("autotuning is not enabled in the config; running the default spaces anyway")
        self.start_time = datetime.datetime.now()
        self.model_info = self.model_info_profile_run()
        if self.model_info:
            logger.info(f"The model has {number_to_string(self.model_info.get('num_params', 0))} parameters.")
        user_zero = self.user_config.get(ZERO_OPTIMIZATION, {})
        user_stage = user_zero.get(ZERO_OPTIMIZATION_STAGE, None) if isinstance(user_zero, dict) else None
        stages = [user_stage] if isinstance(user_stage, int) else [0, 1, 2, 3]
        templates = {0: DEFAULT_TEMPLATE_PATH_ZERO_0, 1: DEFAULT_TEMPLATE_PATH_ZERO_1, 2: DEFAULT_TEMPLATE_PATH_ZERO_2,
                     3: DEFAULT_TEMPLATE_PATH_ZERO_3}
        prev = (0, 0, 0)
        for stage in stages:
            with open(templates[stage]) as f:
                space = hjson.load(f)
            if isinstance(user_zero, dict) and user_stage == stage:
                space = replace_dict(space, {ZERO_OPTIMIZATION: {k: v for k, v in user_zero.items() if k != "stage"}})
            if self.fast_enabled():
                first = get_first_config(space)
                first[ZERO_OPTIMIZATION][ZERO_OPTIMIZATION_STAGE] = stage
                space = first
            cur = self.tune_space(space, *prev)
            if cur[2] and prev[2] and cur[2] < prev[2] * (1 - METRIC_PERCENT_DIFF_CONST) and stage >= 2:
                logger.info(f"ZeRO stage {stage} is slower than the previous stage; stopping the stage sweep")
                break
            if cur[2]:
                prev = cur
        self.end_time = datetime.datetime.now()
        return self.get_best_space_records()

    # ----------------------------------------------------------------------------------- results
    def print_tuning_results(self):
        best = self.get_best_space_records()
        rows = []
        for name, (exp, val, n) in best.items():
            if name != GLOBAL_TUNING_SPACE:
                rows.append(f"| {name:<14} | {n:>10} | {val:>12.2f} | {exp['name']}")
        print("| tuning_space   | num_exps   | best_metric  | best_exp_name")
        print("\n".join(rows))
        if GLOBAL_TUNING_SPACE in best:
            exp, val, n = best[GLOBAL_TUNING_SPACE]
            took = getattr(self, "end_time", datetime.datetime.now()) - getattr(self, "start_time", datetime.datetime.now())
            print(f"{n} experiments took {took}; best: {exp['name']} ({self.metric()} = {val:.2f})")

    def write_optimal_config(self):
        best = self.get_best_space_records()
        if GLOBAL_TUNING_SPACE not in best:
            return
        exp, val, _ = best[GLOBAL_TUNING_SPACE]
        cfg = copy.deepcopy(exp["ds_config"])
        cfg.pop(AUTOTUNING, None)
        path = os.path.join(self.results_dir, "ds_config_optimal.json")
        with open(path, "w") as fd:
            json.dump(cfg, fd, indent=2)
        cmd_path = os.path.join(exp.get("result_dir", self.results_dir), "cmd.txt")
        if os.path.exists(cmd_path):
            with open(cmd_path) as f:
                cmd = f.read().split()
            for flag in ("--deepspeed_config", "--deepspeed"):
                if flag in cmd and cmd.index(flag) + 1 < len(cmd):
                    cmd[cmd.index(flag) + 1] = path
            self.optimal_cmd = cmd
            with open(os.path.join(self.results_dir, "cmd_optimal.txt"), "w") as fd:
                fd.write(" ".join(cmd) + "\n")
        self.optimal_ds_config = cfg
        logger.info(f"Wrote the optimal config to {path}")

    def run_after_tuning(self):
        if self.optimal_cmd:
            import subprocess
            logger.info(f"Running the optimal configuration: {' '.join(self.optimal_cmd)}")
            subprocess.Popen(self.optimal_cmd).wait()
        else:
            logger.info("No optimal DeepSpeed configuration found by autotuning.")
