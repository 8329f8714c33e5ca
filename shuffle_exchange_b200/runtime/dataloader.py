"""Training-data plumbing of the engine (``engine.deepspeed_io``; reference surface: ``DeepSpeedDataLoader`` and
``RepeatingLoader`` in ``deepspeed/runtime/dataloader.py``).

``DeepSpeedDataLoader`` decides ONCE, at construction, how samples are drawn — the three cases are small strategy
objects below — and builds the underlying ``torch.utils.data.DataLoader`` lazily at ``iter()`` time, so that a
sampler's epoch / curriculum state set between epochs is honoured.
"""
from __future__ import annotations

import math
from typing import Callable, Iterator, Optional

from torch.utils.data import DataLoader, RandomSampler
from torch.utils.data.distributed import DistributedSampler

from ..device import get_accelerator


class RepeatingLoader:
    """Endless view of a finite loader: on exhaustion the sampler's epoch advances and iteration restarts."""

    def __init__(self, loader):
        self.loader = loader
        self.epochs_completed = 0
        self.data_iter = iter(loader)

    def __iter__(self):
        return self

    def _restart(self) -> None:
        self.epochs_completed += 1
        for owner in (self.loader, getattr(self.loader, "dataloader", None)):
            for attr in ("sampler", "data_sampler", "batch_sampler"):
                s = getattr(owner, attr, None)
                if s is not None and hasattr(s, "set_epoch"):
                    s.set_epoch(self.epochs_completed)
                    self.data_iter = iter(self.loader)
                    return
        self.data_iter = iter(self.loader)

    def __next__(self):
        try:
            return next(self.data_iter)
        except StopIteration:
            self._restart()
            return next(self.data_iter)


# ----------------------------------------------------------------------------- sampling strategies
class _Plan:
    """How batches are formed: ``loader_kwargs`` go to the torch DataLoader, ``sampler`` is kept for ``len()``."""
    sampler = None
    per_step = 1           # samples one ``next()`` returns on this process
    workers = 0
    batched_by_sampler = False

    def loader_kwargs(self, drop_last: bool) -> dict:
        return {"batch_size": self.per_step, "sampler": self.sampler, "drop_last": drop_last}


class _RankSharded(_Plan):
    """One process per GPU: a DistributedSampler (or the caller's) gives each data-parallel rank its shard."""

    def __init__(self, dataset, batch_size, sampler, dp_world, dp_rank):
        self.sampler = sampler or DistributedSampler(dataset=dataset, num_replicas=dp_world, rank=dp_rank)
        self.per_step, self.workers = batch_size, 2


class _SingleProcess(_Plan):
    """No launcher rank (``local_rank < 0``): one process feeds every visible device, batch scaled accordingly."""

    def __init__(self, dataset, batch_size, sampler):
        n = max(1, get_accelerator().device_count())
        self.sampler = sampler or RandomSampler(dataset)
        self.per_step, self.workers, self.devices = batch_size * n, 2 * n, n


class _Curriculum(_Plan):
    """Data-efficiency sampling: the sampler itself yields whole batches (index lists) following the curriculum."""
    batched_by_sampler = True

    def __init__(self, dataset, batch_size, conf, dp_world, dp_rank, drop_last):
        from .data_pipeline.data_sampling.data_sampler import DeepSpeedDataSampler
        self.sampler = DeepSpeedDataSampler(conf["data_efficiency"], len(dataset), batch_size, dp_rank, dp_world,
                                            conf.get("data_parallel_group"), conf.get("gradient_accumulation_steps", 1),
                                            conf.get("global_rank", 0), drop_last=drop_last)
        self.per_step, self.workers = batch_size, conf.get("num_workers", 0)

    def loader_kwargs(self, drop_last: bool) -> dict:
        return {"batch_sampler": self.sampler}


class DeepSpeedDataLoader:
    def __init__(self, dataset, batch_size, pin_memory, local_rank, tput_timer, collate_fn=None, num_local_io_workers=None,
                 data_sampler=None, data_parallel_world_size=None, data_parallel_rank=None, dataloader_drop_last=False,
                 deepspeed_dataloader_config=None):
        conf = deepspeed_dataloader_config or {}
        self.deepspeed_dataloader_config = conf
        self.curriculum_learning_enabled = bool(conf.get("curriculum_learning_enabled"))
        if self.curriculum_learning_enabled:
            plan = _Curriculum(dataset, batch_size, conf, data_parallel_world_size, data_parallel_rank, dataloader_drop_last)
        elif local_rank >= 0:
            plan = _RankSharded(dataset, batch_size, data_sampler, data_parallel_world_size, data_parallel_rank)
        else:
            plan = _SingleProcess(dataset, batch_size, data_sampler)
        self._plan = plan
        self.dataset, self.collate_fn, self.tput_timer = dataset, collate_fn, tput_timer
        self.data_sampler = plan.sampler
        self.batch_size = plan.per_step
        self.device_count = getattr(plan, "devices", 1)
        self.num_local_io_workers = (plan.workers if num_local_io_workers is None or self.curriculum_learning_enabled
                                     else num_local_io_workers)
        self.pin_memory = bool(pin_memory) and get_accelerator().is_available()
        self.dataloader_drop_last = dataloader_drop_last
        self.post_process_func: Optional[Callable] = None      # curriculum hook: f(batch, sampler_state) -> batch
        n = len(self.data_sampler)
        self.len = n // self.batch_size if dataloader_drop_last else math.ceil(n / self.batch_size)
        self.dataloader: Optional[DataLoader] = None
        self._it: Optional[Iterator] = None

    def __len__(self):
        return self.len

    def _create_dataloader(self) -> DataLoader:
        kw = dict(pin_memory=self.pin_memory, num_workers=self.num_local_io_workers)
        if self.collate_fn is not None:
            kw["collate_fn"] = self.collate_fn
        kw.update(self._plan.loader_kwargs(self.dataloader_drop_last))
        self.dataloader = DataLoader(self.dataset, **kw)
        return self.dataloader

    def __iter__(self):
        self._it = iter(self._create_dataloader())
        return self

    def __next__(self):
        if self.tput_timer:
            self.tput_timer.start()
        batch = next(self._it)
        if self._plan.batched_by_sampler and self.post_process_func is not None:
            batch = self.post_process_func(batch, self.data_sampler.state_dict())
        return batch
