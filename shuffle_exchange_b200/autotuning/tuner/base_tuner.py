"""Tuner loop: repeatedly pick a batch of experiments, run them through the resource manager, keep the best, stop early
after ``early_stopping`` trials without improvement (reference: ``deepspeed/autotuning/tuner/base_tuner.py``)."""
import sys

from ...utils.logging import logger
from ..constants import AUTOTUNING_METRIC_LATENCY
from ..utils import write_experiments


class BaseTuner:
    def __init__(self, exps, resource_manager, metric):
        self.all_exps = exps
        self.rm = resource_manager
        self.best_iter = 0
        self.best_exp = None
        self.best_metric_val = None
        self.metric = metric if metric else AUTOTUNING_METRIC_LATENCY
        logger.info(f"total number of exps =  {len(self.all_exps)}")

    def has_next(self):
        return len(self.all_exps) > 0

    def next_batch(self, sample_size):
        raise NotImplementedError

    def update(self):
        """Hook for tuners that learn from finished experiments."""

    def tune(self, sample_size=1, n_trials=1000, early_stopping=None):
        i = 0
        try:
            while i < n_trials and self.has_next():
                sampled = self.next_batch(sample_size)
                paths = write_experiments(sampled, self.rm.exps_dir)
                self.rm.schedule_experiments(paths)
                self.rm.run()
                exp, metric_val = self.rm.parse_results(self.metric)
                if self.best_exp is None or (metric_val is not None and
                                             (self.best_metric_val is None or metric_val > self.best_metric_val)):
                    self.best_exp, self.best_metric_val, self.best_iter = exp, metric_val, i
                i += len(sampled)
                self.rm.clear()
                self.update()
                if early_stopping and i >= self.best_iter + early_stopping:
                    logger.info(f"Tuner early stopped at iteration {i}. Best iteration is {self.best_iter}. Early stopping threshold is {early_stopping}")
                    break
            return i
        except Exception:
            logger.info("Tuner Error:", sys.exc_info()[0])
            return i
