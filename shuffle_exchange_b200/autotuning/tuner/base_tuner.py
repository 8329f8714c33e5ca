"""Search loop shared by the grid / random / model-based tuners.

A tuner owns a pool of candidate experiments.  Each round it draws a batch (:meth:`next_batch`, the only thing a
subclass must provide), has the resource manager run it, folds the round's best result into the incumbent and lets the
subclass learn from it (:meth:`update`).  The search ends when the pool is empty, the trial budget is spent, or
``early_stopping`` trials have gone by without a new incumbent.  (Reference behaviour:
``deepspeed/autotuning/tuner/base_tuner.py`` — same public surface ``tune / has_next / next_batch / update`` and
``best_exp / best_metric_val / best_iter``.)"""
from __future__ import annotations

from ...utils.logging import logger
from ..constants import AUTOTUNING_METRIC_LATENCY
from ..utils import write_experiments


class BaseTuner:
    def __init__(self, exps, resource_manager, metric):
        self.all_exps = exps
        self.rm = resource_manager
        self.metric = metric or AUTOTUNING_METRIC_LATENCY
        self.best_exp, self.best_metric_val, self.best_iter = None, None, 0
        self.history = []                       # (trials done before the round, experiment, metric value)
        logger.info(f"tuner: {len(exps)} candidate experiment(s), metric = {self.metric}")

    # ------------------------------------------------------------------ subclass interface
    def has_next(self) -> bool:
        return bool(self.all_exps)

    def next_batch(self, sample_size):
        raise NotImplementedError

    def update(self):
        """Called after every round; model-based tuners refit here."""

    # --------------------------------------------------------------------------- the loop
    def _run_round(self, batch):
        self.rm.schedule_experiments(write_experiments(batch, self.rm.exps_dir))
        self.rm.run()
        return self.rm.parse_results(self.metric)       # finished experiments stay with the manager until update() ran

    def _improves(self, value) -> bool:
        if self.best_exp is None:
            return True
        return value is not None and (self.best_metric_val is None or value > self.best_metric_val)

    def tune(self, sample_size=1, n_trials=1000, early_stopping=None) -> int:
        done = 0
        while done < n_trials and self.has_next():
            batch = self.next_batch(sample_size)
            try:
                exp, value = self._run_round(batch)
            except Exception as err:            # a failing experiment ends the search but keeps the incumbent
                logger.info(f"tuner stopped by {type(err).__name__}: {err}")
                self.rm.clear()
                break
            self.history.append((done, exp, value))
            if self._improves(value):
                self.best_exp, self.best_metric_val, self.best_iter = exp, value, done
            done += len(batch)
            self.update()                       # model-based tuners read the round's measurements here ...
            self.rm.clear()                     # ... so the manager forgets them only afterwards
            if early_stopping and done - self.best_iter >= early_stopping:
                logger.info(f"tuner: no improvement for {done - self.best_iter} trial(s) (limit {early_stopping}); "
                            f"best was found at trial {self.best_iter}")
                break
        return done
