"""Model-based tuner: fit the cost model on what was measured so far, evaluate the predicted-best untried
configurations next (with an epsilon of random exploration).  Reference: ``tuner/model_based_tuner.py``."""
import numbers
import random

import numpy as np

from ..constants import AUTOTUNING_METRIC_LATENCY
from ..utils import dict_to_feature, flatten, get_tuning_keys
from .base_tuner import BaseTuner
from .cost_model import XGBoostCostModel

INIT_NUM = 2


class ModelBasedTuner(BaseTuner):
    def __init__(self, exps: list, resource_manager, metric, tuning_space):
        super().__init__(exps, resource_manager, metric)
        self.tuning_space = tuning_space
        self.best_iter = 0
        self.all_configs = [e["ds_config"] for e in exps]
        self.num_all_configs = len(self.all_configs)
        self.dims = None
        self.max_val = self._max_values()
        self.tuning_keys = get_tuning_keys(tuning_space)
        self.evaluated_configs, self.evaluated_perf = [], []
        self.trials, self.trial_pt = [], 0
        seed = [random.randrange(self.num_all_configs) for _ in range(min(INIT_NUM, self.num_all_configs))]
        self.visited = set()
        for i in seed:
            if i not in self.visited:
                self.trials.append(i)
                self.visited.add(i)
        self.evaluated = set()
        self.cost_model = XGBoostCostModel("rank")
        self.random_exploration_ratio = 0.2

    def _max_values(self):
        mx = {}
        for cfg in self.all_configs:
            for k, v in flatten(cfg).items():
                if isinstance(v, numbers.Number) and not isinstance(v, bool):
                    mx[k] = max(mx.get(k, 0), abs(v)) or 1
        return mx

    def _feature(self, cfg):
        flat = flatten(cfg)
        keys = [k for k in flat if any(k.endswith(t) for t in self.tuning_keys) or k in self.tuning_keys]
        return dict_to_feature(flat, keys, self.max_val)

    def find_estimated_top_configs(self):
        feats = [self._feature(c) for c in self.all_configs]
        width = max(len(f) for f in feats)
        feats = np.array([f + [0.0] * (width - len(f)) for f in feats])
        pred = self.cost_model.predict(feats)
        order = np.argsort(pred)
        if self.metric != AUTOTUNING_METRIC_LATENCY:
            order = order[::-1]
        return [int(i) for i in order]

    def next_batch(self, sample_size):
        batch = []
        while len(batch) < sample_size and (self.trial_pt < len(self.trials) or len(self.visited) < self.num_all_configs):
            if self.trial_pt >= len(self.trials):
                rest = [i for i in range(self.num_all_configs) if i not in self.visited]
                if not rest:
                    break
                i = random.choice(rest)
                self.trials.append(i)
                self.visited.add(i)
            idx = self.trials[self.trial_pt]
            self.trial_pt += 1
            batch.append(self.all_exps[idx])
            self.evaluated.add(idx)
        return batch

    def has_next(self):
        return len(self.evaluated) < self.num_all_configs

    def update(self):
        for exp_id, (exp, err) in self.rm.finished_experiments.items():
            if err:
                continue
            cfg = exp["ds_config"]
            if cfg in self.evaluated_configs:
                continue
            perf = self.rm.read_metric(exp, self.metric) if hasattr(self.rm, "read_metric") else None
            if perf is None:
                continue
            self.evaluated_configs.append(cfg)
            self.evaluated_perf.append(perf)
        if len(self.evaluated_configs) >= 2:
            feats = [self._feature(c) for c in self.evaluated_configs]
            width = max(len(f) for f in feats)
            self.cost_model.fit([f + [0.0] * (width - len(f)) for f in feats], self.evaluated_perf)
            for i in self.find_estimated_top_configs():
                if i not in self.visited:
                    if random.random() < self.random_exploration_ratio:
                        rest = [j for j in range(self.num_all_configs) if j not in self.visited]
                        i = random.choice(rest)
                    self.trials.append(i)
                    self.visited.add(i)
                    break
