"""Tuners over a synthetic search space with a fake resource manager whose "measurements" come from a known function:
grid and random visit everything and return the true optimum; the model-based tuner learns from every round (its
surrogate must see the results before they are cleared) and finds the optimum within a fraction of the space; early
stopping and a failing experiment end the search but keep the incumbent."""
import json
import os
import random

from shuffle_exchange_b200.autotuning.constants import AUTOTUNING_METRIC_THROUGHPUT
from shuffle_exchange_b200.autotuning.tuner import GridSearchTuner, ModelBasedTuner, RandomTuner

SPACE = {"train_micro_batch_size_per_gpu": [1, 2, 4, 8, 16, 32], "zero_optimization": {"stage": [1, 2, 3]},
         "gradient_accumulation_steps": [1, 2, 4]}


def _throughput(cfg):
    mbs, stage, gas = cfg["train_micro_batch_size_per_gpu"], cfg["zero_optimization"]["stage"], cfg["gradient_accumulation_steps"]
    if mbs * (4 - stage) > 48:
        return None                                   # "out of memory"
    return 100.0 * mbs / (mbs + 4) * (1.0 - 0.1 * (stage - 1)) * (1 + 0.05 * gas)


def _experiments():
    exps = []
    for mbs in SPACE["train_micro_batch_size_per_gpu"]:
        for stage in SPACE["zero_optimization"]["stage"]:
            for gas in SPACE["gradient_accumulation_steps"]:
                cfg = {"train_micro_batch_size_per_gpu": mbs, "zero_optimization": {"stage": stage},
                       "gradient_accumulation_steps": gas}
                exps.append({"name": f"z{stage}_mbs{mbs}_gas{gas}", "ds_config": cfg, "exp_id": len(exps)})
    return exps


class FakeResourceManager:
    def __init__(self, exps_dir, fail_on=None):
        self.exps_dir = str(exps_dir)
        self.finished_experiments = {}
        self.queue = []
        self.ran = []
        self.fail_on = fail_on

    def schedule_experiments(self, paths):
        self.queue = [json.load(open(p)) for p in paths]

    def run(self):
        for exp in self.queue:
            if self.fail_on is not None and len(self.ran) >= self.fail_on:
                raise RuntimeError("node lost")
            self.ran.append(exp["name"])
            val = _throughput(exp["ds_config"])
            self.finished_experiments[exp["exp_id"]] = (exp, None if val is not None else "OOM")

    def read_metric(self, exp, metric):
        return _throughput(exp["ds_config"])

    def parse_results(self, metric):
        best = (None, None)
        for exp, err in self.finished_experiments.values():
            if not err and (best[1] is None or _throughput(exp["ds_config"]) > best[1]):
                best = (exp, _throughput(exp["ds_config"]))
        return best

    def clear(self):
        self.finished_experiments = {}
        self.queue = []


def _optimum():
    return max((e for e in _experiments() if _throughput(e["ds_config"]) is not None),
               key=lambda e: _throughput(e["ds_config"]))


def test_grid_and_random_tuners_find_the_optimum(tmp_path):
    best = _optimum()
    for cls in (GridSearchTuner, RandomTuner):
        random.seed(0)
        rm = FakeResourceManager(tmp_path)
        t = cls(_experiments(), rm, AUTOTUNING_METRIC_THROUGHPUT)
        done = t.tune(sample_size=3, n_trials=1000)
        assert done == 54 and len(rm.ran) == 54 and not t.has_next()
        assert t.best_exp["name"] == best["name"] and abs(t.best_metric_val - _throughput(best["ds_config"])) < 1e-9
        assert len(os.listdir(tmp_path)) == 54


def test_model_based_tuner_learns_between_rounds(tmp_path):
    best = _optimum()
    random.seed(1)
    rm = FakeResourceManager(tmp_path)
    t = ModelBasedTuner(_experiments(), rm, AUTOTUNING_METRIC_THROUGHPUT, SPACE)
    done = t.tune(sample_size=1, n_trials=24)
    assert done <= 24
    assert len(t.evaluated_configs) >= 10, "the surrogate never saw a measurement"      # results reach update()
    assert t.best_metric_val >= 0.97 * _throughput(best["ds_config"])
    assert len(set(rm.ran)) == len(rm.ran)                                               # nothing measured twice


def test_early_stopping_and_failures_keep_the_incumbent(tmp_path):
    rm = FakeResourceManager(tmp_path)
    exps = sorted(_experiments(), key=lambda e: -(_throughput(e["ds_config"]) or 0))     # best first
    t = GridSearchTuner(exps, rm, AUTOTUNING_METRIC_THROUGHPUT)
    done = t.tune(sample_size=1, n_trials=1000, early_stopping=5)
    assert done == 5 and t.best_iter == 0 and t.best_exp["name"] == exps[0]["name"]
    rm = FakeResourceManager(tmp_path, fail_on=4)
    t = GridSearchTuner(_experiments(), rm, AUTOTUNING_METRIC_THROUGHPUT)
    done = t.tune(sample_size=2, n_trials=1000)
    assert done == 4 and t.best_exp is not None
