"""Token-budget batching with batch-size-dependent learning rate.

Reference: ``data_sampling/variable_batch_size_and_lr.py:23-492`` — pack sequences into micro-batches of at most
``max_tokens`` total length ("Attention is all you need" §5.1 batching), group ``effective_batch_size`` micro-batches
(= dp ranks x gradient-accumulation steps) into one step, and scale the learning rate of each step by its number of
sequences relative to ``base_batch_size`` (linear / sqrt / none).

The packer below is a single greedy pass (O(n)); the reference re-validates every candidate prefix (quadratic), the
produced batches are the same: the longest valid prefix starting at the cursor.
"""
from __future__ import annotations

import random
from typing import List, Sequence

import numpy as np
import torch
from torch.optim.lr_scheduler import LRScheduler
from torch.utils.data import DataLoader, DistributedSampler

from ....utils.logging import logger


def batch_by_seqlens(seqlens: Sequence[int], max_tokens: int, sequence_ids_per_mb=None, min_batch_size: int = 1,
                     max_batch_size=None, sequence_picking_order: str = "dataloader", effective_batch_size: int = 1,
                     required_microbatches_of_same_size: bool = False, verbose: bool = False, seed=None):
    """Returns ``(microbatch_ids, batch_sizes, batch_max_seqlens)``: ``microbatch_ids`` is a list of
    ``(batch_id, [sample ids])``, ``effective_batch_size`` consecutive entries per optimizer step."""
    assert sequence_picking_order in ("random", "seqlen", "dataloader")
    ids = list(range(len(seqlens))) if sequence_ids_per_mb is None else list(sequence_ids_per_mb)
    items = [(int(seqlens[i]), i) for i in ids]
    if sequence_picking_order == "random":
        random.Random(seed).shuffle(items)
    elif sequence_picking_order == "seqlen":
        items.sort()
    too_long = [i for v, i in items if v > max_tokens]
    if too_long:
        logger.warning(f"Data indices {too_long} ignored as metrics exceed {max_tokens}.")
        drop = set(too_long)
        items = [m for m in items if m[1] not in drop]

    lanes = effective_batch_size if required_microbatches_of_same_size else 1
    microbatches: List[list] = []
    pos, n = 0, len(items)
    while pos < n:
        # grow `lanes` interleaved micro-batches in lock step while every lane stays within the limits
        lane_tokens = [0] * lanes
        lane_count = 0
        best = 0                                   # number of full rounds that is valid
        rounds = 0
        while pos + (rounds + 1) * lanes <= n:
            chunk = items[pos + rounds * lanes: pos + (rounds + 1) * lanes]
            for l, (v, _) in enumerate(chunk):
                lane_tokens[l] += v
            lane_count += 1
            rounds += 1
            if max(lane_tokens) > max_tokens or (max_batch_size and lane_count > max_batch_size):
                break
            if lane_count >= (min_batch_size or 1):
                best = rounds
        if best == 0:
            break
        taken = items[pos: pos + best * lanes]
        microbatches += [taken[l::lanes] for l in range(lanes)]
        pos += best * lanes

    assert len(microbatches) >= effective_batch_size, "not enough datapoints to create a single sample per dataloader"
    microbatches = microbatches[: len(microbatches) - len(microbatches) % effective_batch_size]
    microbatch_ids, batch_sizes, batch_max_seqlens = [], [], []
    for start in range(0, len(microbatches), effective_batch_size):
        bid = start // effective_batch_size
        mbs = microbatches[start: start + effective_batch_size]
        batch_sizes.append(sum(len(mb) for mb in mbs))
        batch_max_seqlens.append(max(v for mb in mbs for v, _ in mb))
        microbatch_ids += [(bid, [i for _, i in mb]) for mb in mbs]
        if verbose:
            logger.info(f"Batch id {bid}: {len(mbs)} microbatches, {batch_sizes[-1]} sequences, tokens per microbatch "
                        f"{[sum(v for v, _ in mb) for mb in mbs]}")
    return microbatch_ids, batch_sizes, batch_max_seqlens


def scale_lr(base_batch_size, batch_size, base_lr=1, method="linear"):
    if method == "linear":
        return base_lr * batch_size / base_batch_size       # Goyal et al., "Accurate, Large Minibatch SGD"
    if method == "sqrt":
        return base_lr * (batch_size / base_batch_size) ** 0.5   # Krizhevsky, "One weird trick"
    if method in (None, "none"):
        return base_lr
    raise ValueError("Unknown scaling method: {}".format(method))


def dataloader_for_variable_batch_size(dataset, microbatch_ids, batch_max_seqlens, dataloader_rank=0,
                                       dataloader_batch_size=1, dataloader_num_replicas=1, dataloader_collate_fn=None,
                                       dataloader_num_workers=2, dataloader_pin_memory=False,
                                       required_microbatches_of_same_seqlen=False, sample_padding_fn=None):
    """Dataloader over the pre-computed micro-batches: element ``i`` of the sampler is micro-batch ``i`` (a list of
    sample ids); ``DistributedSampler`` (no shuffle) deals consecutive micro-batches to the dp ranks."""
    sampler = DistributedSampler(dataset=microbatch_ids, num_replicas=dataloader_num_replicas, rank=dataloader_rank,
                                 shuffle=False, drop_last=False)

    def collate(list_of_microbatches):
        assert len(list_of_microbatches) == 1, "a dataloader batch is exactly one pre-packed micro-batch"
        bid, sample_ids = list_of_microbatches[0]
        samples = [dataset[i] for i in sample_ids]
        if required_microbatches_of_same_seqlen:
            assert sample_padding_fn is not None, "padding dataloader_padding_fn must be provided if required_microbatches_of_same_seqlen is True"
            samples = [sample_padding_fn(s, batch_max_seqlens[bid]) for s in samples]
        return dataloader_collate_fn(samples) if dataloader_collate_fn else samples

    return DataLoader(dataset=microbatch_ids, batch_size=dataloader_batch_size, sampler=sampler,
                      num_workers=dataloader_num_workers, collate_fn=collate, pin_memory=dataloader_pin_memory)


class VariableBatchSizeLR(LRScheduler):
    """Wraps any LR scheduler: after every ``step`` the wrapped scheduler's rates are multiplied by the scale of the
    *next* batch's size."""

    @property
    def optimizer(self):
        return self.base_lr_scheduler.optimizer

    def __init__(self, lr_scheduler, base_batch_size, batch_sizes, dataloader, lr_scaling_method="linear",
                 last_epoch=-1, verbose=False):
        self.batch_sizes = batch_sizes
        self.base_batch_size = base_batch_size
        self.lr_scaling_method = lr_scaling_method
        self.dataloader = dataloader
        self.base_lr_scheduler = lr_scheduler
        self.base_lrs = [g["lr"] for g in self.base_lr_scheduler.optimizer.param_groups]
        self.last_epoch = last_epoch
        self.verbose = verbose
        self.step(0)

    def state_dict(self):
        return {"base_lr_scheduler": self.base_lr_scheduler.state_dict(), "base_batch_size": self.base_batch_size,
                "lr_scaling_method": self.lr_scaling_method, "batch_sizes": self.batch_sizes,
                "base_lrs": self.base_lrs, "last_epoch": self.last_epoch, "verbose": self.verbose}

    def load_state_dict(self, state_dict):
        self.base_lr_scheduler.load_state_dict(state_dict["base_lr_scheduler"])
        for k in ("base_batch_size", "lr_scaling_method", "batch_sizes", "base_lrs", "last_epoch", "verbose"):
            setattr(self, k, state_dict[k])

    def get_last_lr(self):
        return [g["lr"] for g in self.base_lr_scheduler.optimizer.param_groups]

    def get_lr(self):
        return [g["lr"] for g in self.base_lr_scheduler.optimizer.param_groups]

    def step(self, epoch=None):
        if self.last_epoch < 0:
            self.last_epoch = 0
        elif epoch is None:
            self.last_epoch += 1
        else:
            self.last_epoch = epoch
        bid = self.last_epoch % len(self.batch_sizes)
        # undo the previous scaling, advance the wrapped schedule, scale for the batch about to be consumed
        for g, lr in zip(self.base_lr_scheduler.optimizer.param_groups, self.base_lrs):
            g["lr"] = lr
        self.base_lr_scheduler.step()
        self.base_lrs = [g["lr"] for g in self.base_lr_scheduler.optimizer.param_groups]
        for g in self.base_lr_scheduler.optimizer.param_groups:
            g["lr"] = scale_lr(self.base_batch_size, self.batch_sizes[bid], g["lr"], self.lr_scaling_method)
            if self.verbose:
                logger.info(f"Batch id {self.last_epoch}, lr {g['lr']:.8f}, batch size {self.batch_sizes[bid]}")


def lr_scheduler_for_variable_batch_size(base_batch_size, batch_sizes, dataloader, lr_scheduler_or_optimizer,
                                         lr_scaling_method="linear"):
    """Accepts an optimizer (constant base LR) or an LR scheduler."""
    if isinstance(lr_scheduler_or_optimizer, torch.optim.Optimizer):
        class _Stub(LRScheduler):
            def __init__(self, optimizer):
                self.optimizer = optimizer
                self.base_lrs = [g["lr"] for g in optimizer.param_groups]
                self.last_epoch = 0

            def get_lr(self):
                return [g["lr"] for g in self.optimizer.param_groups]

            def step(self, epoch=None):
                pass

            def state_dict(self):
                return {}

            def load_state_dict(self, sd):
                pass
        sched = _Stub(lr_scheduler_or_optimizer)
    elif hasattr(lr_scheduler_or_optimizer, "optimizer"):
        sched = lr_scheduler_or_optimizer
    else:
        raise ValueError("Unknown type for lr_scheduler_or_optimizer: {}".format(type(lr_scheduler_or_optimizer)))
    return VariableBatchSizeLR(lr_scheduler=sched, base_batch_size=base_batch_size, batch_sizes=batch_sizes,
                               dataloader=dataloader, lr_scaling_method=lr_scaling_method)


def get_dataloader_and_lr_scheduler_for_variable_batch_size(dataset, dataset_seqlens, max_tokens, effective_batch_size,
                                                            base_batch_size=None, dataset_filter_ids=None,
                                                            lr_scheduler_or_optimizer=None, lr_scaling_method="linear",
                                                            min_batch_size=1, max_batch_size=None,
                                                            sequence_picking_order="dataloader", dataloader_batch_size=1,
                                                            dataloader_rank=0, dataloader_num_replicas=1,
                                                            dataloader_num_workers=0, dataloader_collate_fn=None,
                                                            dataloader_pin_memory=False,
                                                            required_microbatches_of_same_size=False,
                                                            required_microbatches_of_same_seqlen=False,
                                                            sample_padding_fn=None, verbose=False, seed=None):
    mb_ids, batch_sizes, max_seqlens = batch_by_seqlens(
        seqlens=dataset_seqlens, max_tokens=max_tokens, sequence_ids_per_mb=dataset_filter_ids,
        min_batch_size=min_batch_size, max_batch_size=max_batch_size, sequence_picking_order=sequence_picking_order,
        effective_batch_size=effective_batch_size, required_microbatches_of_same_size=required_microbatches_of_same_size,
        verbose=verbose, seed=seed)
    loader = dataloader_for_variable_batch_size(
        dataset=dataset, microbatch_ids=mb_ids, batch_max_seqlens=max_seqlens, dataloader_rank=dataloader_rank,
        dataloader_batch_size=dataloader_batch_size, dataloader_num_replicas=dataloader_num_replicas,
        dataloader_collate_fn=dataloader_collate_fn, dataloader_num_workers=dataloader_num_workers,
        dataloader_pin_memory=dataloader_pin_memory,
        required_microbatches_of_same_seqlen=required_microbatches_of_same_seqlen, sample_padding_fn=sample_padding_fn)
    sched = None
    if lr_scheduler_or_optimizer is not None:
        sched = lr_scheduler_for_variable_batch_size(base_batch_size=base_batch_size, batch_sizes=batch_sizes,
                                                     dataloader=loader, lr_scaling_method=lr_scaling_method,
                                                     lr_scheduler_or_optimizer=lr_scheduler_or_optimizer)
    return loader, sched


def get_dataloader_and_lr_scheduler_for_variable_batch_size_deepspeed(dataset, engine, dataset_seqlens=None,
                                                                      dataset_filter_ids=None, dataloader_collate_fn=None,
                                                                      sample_padding_fn=None, batch_seqlens_fn=None):
    """Engine-aware wrapper: reads ``data_efficiency.data_sampling.dynamic_batching`` from the engine config and the
    dp rank / gas / pipeline requirements from the engine itself."""
    from ..constants import (DATA_SAMPLING, DATA_SAMPLING_NUM_WORKERS, DATA_SAMPLING_PIN_MEMORY, DYNAMIC_BATCHING,
                             DYNAMIC_BATCHING_LR_SCALING_METHOD, DYNAMIC_BATCHING_MAX_BATCH_SIZE,
                             DYNAMIC_BATCHING_MAX_TOKENS, DYNAMIC_BATCHING_METRICS_PATH, DYNAMIC_BATCHING_MIN_BATCH_SIZE,
                             DYNAMIC_BATCHING_SEQUENCE_PICKING_ORDER, DYNAMIC_BATCHING_VERBOSE)
    from .indexed_dataset import MMapIndexedDataset
    ds_cfg = engine.data_efficiency_config()[DATA_SAMPLING]
    cfg = ds_cfg[DYNAMIC_BATCHING]
    if dataset_seqlens is None:
        sample_to_seqlen = MMapIndexedDataset(cfg[DYNAMIC_BATCHING_METRICS_PATH], skip_warmup=True)
        dataset_seqlens = np.array([sample_to_seqlen[i][0] for i in range(len(sample_to_seqlen))])
    from ...pipe.engine import PipelineEngine
    pipelined = isinstance(engine, PipelineEngine)
    return get_dataloader_and_lr_scheduler_for_variable_batch_size(
        dataset=dataset, dataset_filter_ids=dataset_filter_ids, dataset_seqlens=dataset_seqlens,
        effective_batch_size=engine.dp_world_size * engine.gradient_accumulation_steps(),
        max_tokens=cfg[DYNAMIC_BATCHING_MAX_TOKENS], lr_scaling_method=cfg.get(DYNAMIC_BATCHING_LR_SCALING_METHOD, "linear"),
        sequence_picking_order=cfg.get(DYNAMIC_BATCHING_SEQUENCE_PICKING_ORDER, "dataloader"),
        min_batch_size=cfg.get(DYNAMIC_BATCHING_MIN_BATCH_SIZE, 1), max_batch_size=cfg.get(DYNAMIC_BATCHING_MAX_BATCH_SIZE),
        dataloader_batch_size=engine.train_micro_batch_size_per_gpu(), dataloader_rank=engine.data_parallel_group.rank()
        if hasattr(engine.data_parallel_group, "rank") else 0, dataloader_num_replicas=engine.dp_world_size,
        dataloader_num_workers=ds_cfg.get(DATA_SAMPLING_NUM_WORKERS, 0), dataloader_collate_fn=dataloader_collate_fn,
        dataloader_pin_memory=ds_cfg.get(DATA_SAMPLING_PIN_MEMORY, False), sample_padding_fn=sample_padding_fn,
        lr_scheduler_or_optimizer=engine.lr_scheduler or engine.optimizer,
        required_microbatches_of_same_size=pipelined, required_microbatches_of_same_seqlen=pipelined,
        verbose=cfg.get(DYNAMIC_BATCHING_VERBOSE, False), base_batch_size=engine.train_batch_size())
