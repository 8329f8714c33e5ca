"""Exhaustive and random tuners (reference: ``deepspeed/autotuning/tuner/index_based_tuner.py``)."""
import random

from .base_tuner import BaseTuner


class RandomTuner(BaseTuner):
    """Sample experiments uniformly without replacement."""

    def next_batch(self, sample_size=1):
        batch = []
        for _ in range(min(sample_size, len(self.all_exps))):
            batch.append(self.all_exps.pop(random.randrange(len(self.all_exps))))
        return batch


class GridSearchTuner(BaseTuner):
    """Walk the space in order."""

    def next_batch(self, sample_size=1):
        batch, self.all_exps = self.all_exps[:sample_size], self.all_exps[sample_size:]
        return batch
