"""Offline dataset analysis for curriculum learning: compute per-sample difficulty metrics and write the index
files :class:`DeepSpeedDataSampler` consumes.

Reference: ``data_sampling/data_analyzer.py:22-815`` (``DataAnalyzer`` thread/worker map-reduce over files and
``DistributedDataAnalyzer`` over ``torch.distributed``).  Output layout kept (all MMIDIDX datasets):

    <save_path>/<metric>/<metric>_sample_to_metric              one item per sample: [value]
    <save_path>/<metric>/<metric>_index_to_metric               one item per distinct value (ascending): [value]
    <save_path>/<metric>/<metric>_index_to_sample               same order: [sample ids having that value]
    <save_path>/<metric>/<metric>_index_to_sample_percentile_merged   100 items: sample ids of each percentile
    <save_path>/<metric>/<metric>_metric_value                  (accumulate_value_over_samples) the summed vector

The map phase is vectorised per batch; sorting/grouping in the reduce phase is ``np.argsort`` + ``np.unique`` on
the gathered value array instead of the reference's per-value file merging.
"""
from __future__ import annotations

import os
from typing import Callable, List

import numpy as np
import torch
from torch.utils.data import BatchSampler, DataLoader, SequentialSampler, Subset

from .... import comm as dist
from ....utils.logging import logger
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype, split_index

SINGLE = "single_value_per_sample"
ACCUMULATE = "accumulate_value_over_samples"


def _write_items(fname, items, dtype):
    b = create_mmap_dataset_builder(fname, dtype)
    for it in items:
        b.add_item_numpy(np.asarray(it, dtype=dtype).reshape(-1))
    close_mmap_dataset_builder(b, fname)


def write_metric_indexes(save_path: str, metric: str, values: np.ndarray, sample_ids: np.ndarray, total_samples: int):
    """Reduce phase for one ``single_value_per_sample`` metric."""
    folder = f"{save_path}/{metric}"
    os.makedirs(folder, exist_ok=True)
    order = np.argsort(sample_ids, kind="stable")
    values, sample_ids = values[order], sample_ids[order]
    vdtype = find_fit_int_dtype(int(values.min()), int(values.max()))
    sdtype = find_fit_int_dtype(0, max(total_samples - 1, 0))
    _write_items(f"{folder}/{metric}_sample_to_metric", values.reshape(-1, 1), vdtype)
    by_value = np.argsort(values, kind="stable")
    uniq, starts = np.unique(values[by_value], return_index=True)
    groups = np.split(sample_ids[by_value], starts[1:])
    _write_items(f"{folder}/{metric}_index_to_metric", uniq.reshape(-1, 1), vdtype)
    _write_items(f"{folder}/{metric}_index_to_sample", groups, sdtype)
    DataAnalyzer.output_index_to_sample_percentile(f"{folder}/{metric}_index_to_sample",
                                                   f"{folder}/{metric}_index_to_metric", metric, folder,
                                                   total_samples, sdtype)


class DataAnalyzer:
    def __init__(self, dataset, num_workers=1, worker_id=0, num_threads=1, num_threads_reduce=1, specific_threads=[],
                 batch_size=1, metric_names=[], metric_functions=[], metric_types=[], metric_dtypes=[], save_path="./",
                 collate_fn=None, custom_map_init=None, custom_map_update=None, custom_map_finalize=None,
                 custom_reduce=None, sample_indices=None):
        self.dataset = dataset
        self.num_workers, self.worker_id = num_workers, worker_id
        self.num_threads, self.num_threads_reduce, self.specific_threads = num_threads, num_threads_reduce, specific_threads
        self.batch_size = batch_size
        self.metric_names, self.metric_functions = list(metric_names), list(metric_functions)
        self.metric_types, self.metric_dtypes = list(metric_types), list(metric_dtypes)
        self.save_path, self.collate_fn = save_path, collate_fn
        self.custom_map_init, self.custom_map_update = custom_map_init, custom_map_update
        self.custom_map_finalize, self.custom_reduce = custom_map_finalize, custom_reduce
        self.sample_indices = sample_indices

    # ------------------------------------------------------------------------------------ map
    def _shard_file(self, metric, worker_id):
        return f"{self.save_path}/{metric}/worker{worker_id}_map.npz"

    def run_map(self):
        """Every worker processes its contiguous slice of the dataset and stores raw (sample id, value) arrays."""
        lo, hi = split_index(0, len(self.dataset), self.num_workers)[self.worker_id]
        loader = DataLoader(Subset(self.dataset, range(int(lo), int(hi))), batch_size=self.batch_size, shuffle=False,
                            num_workers=0, collate_fn=self.collate_fn)
        values = {m: [] for m in self.metric_names}
        ids = []
        cursor = int(lo)
        for batch in loader:
            n = None
            for m, fn, typ in zip(self.metric_names, self.metric_functions, self.metric_types):
                v = fn(batch)
                v = v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)
                if typ == SINGLE:
                    v = v.reshape(-1)
                    n = len(v)
                values[m].append(v)
            if n is None:
                n = self.batch_size
            if self.sample_indices is not None:
                ids.append(np.asarray(self.sample_indices[cursor: cursor + n]))
            else:
                ids.append(np.arange(cursor, cursor + n))
            cursor += n
        ids = np.concatenate(ids) if ids else np.zeros(0, dtype=np.int64)
        for m, typ in zip(self.metric_names, self.metric_types):
            os.makedirs(f"{self.save_path}/{m}", exist_ok=True)
            if typ == SINGLE:
                v = np.concatenate(values[m]) if values[m] else np.zeros(0)
                np.savez(self._shard_file(m, self.worker_id), ids=ids, values=v)
            else:
                total = np.sum(np.stack(values[m]), axis=0) if values[m] else np.zeros(0)
                np.savez(self._shard_file(m, self.worker_id), values=total)
        logger.info(f"worker {self.worker_id}: map phase finished on samples [{lo}, {hi})")

    # --------------------------------------------------------------------------------- reduce
    def run_reduce(self):
        if self.custom_reduce is not None:
            return self.custom_reduce(self.dataset, self.metric_names, self.metric_types, self.save_path,
                                      self.num_workers, self.num_threads, self.num_threads_reduce)
        for m, typ, dt in zip(self.metric_names, self.metric_types, self.metric_dtypes):
            shards = [np.load(self._shard_file(m, w)) for w in range(self.num_workers)]
            if typ == SINGLE:
                ids = np.concatenate([s["ids"] for s in shards])
                vals = np.concatenate([s["values"] for s in shards])
                write_metric_indexes(self.save_path, m, vals, ids, len(self.dataset))
            elif typ == ACCUMULATE:
                total = np.sum(np.stack([s["values"] for s in shards]), axis=0)
                _write_items(f"{self.save_path}/{m}/{m}_metric_value", [total], dt or np.int64)
            else:
                raise ValueError(f"unknown metric type {typ}")

    def run_map_reduce(self, comm_group=None):
        self.run_map()
        if dist.is_initialized():
            dist.barrier(group=comm_group)
        if self.worker_id == 0:
            self.run_reduce()
        if dist.is_initialized():
            dist.barrier(group=comm_group)

    # -------------------------------------------------------------------------------- helpers
    @staticmethod
    def get_metric_value_percentiles(metric_name, num_sample_per_value, total_num_samples):
        logger.info(f"Checking the value percentiles of metric {metric_name}...")
        processed, current_pct, out = 0, 0, {}
        for value in sorted(num_sample_per_value):
            processed += num_sample_per_value[value]
            pct = int(processed * 100 / total_num_samples) if total_num_samples else 100
            if pct > current_pct:
                out[pct] = value
                current_pct = pct
        return out

    @staticmethod
    def output_index_to_sample_percentile(index_to_sample_fname, index_to_metric_fname, metric_name, metric_save_path,
                                          total_num_samples, sample_idx_dtype):
        """Merge the value-ordered sample lists into 100 equal-population percentile buckets."""
        to_sample = MMapIndexedDataset(index_to_sample_fname, skip_warmup=True)
        flat = np.concatenate([np.array(to_sample[r]) for r in range(len(to_sample))]) if len(to_sample) else np.zeros(0)
        edges = np.linspace(0, len(flat), 101, dtype=np.int64)
        _write_items(f"{metric_save_path}/{metric_name}_index_to_sample_percentile_merged",
                     [flat[edges[i]: edges[i + 1]] for i in range(100)], sample_idx_dtype)


class DistributedDataAnalyzer:
    """Same outputs, but ranks of a process group each analyse a slice and the (id, value) pairs are gathered on
    rank 0 with ``gather`` of padded tensors (reference :455-730 uses a distributed sample sort; the gathered arrays
    here are a few bytes per sample, far below the dataset itself)."""

    def __init__(self, dataset, num_workers=1, num_threads=1, worker_id=0, batch_size=1, metric_names=[],
                 metric_functions=[], metric_types=[], save_path="./", collate_fn=None, device="cpu",
                 comm_group=None, sample_indices=None):
        self.dataset, self.batch_size = dataset, batch_size
        self.metric_names, self.metric_functions, self.metric_types = list(metric_names), list(metric_functions), list(metric_types)
        self.save_path, self.collate_fn, self.device = save_path, collate_fn, device
        self.sample_indices = sample_indices
        if not dist.is_initialized():
            dist.init_distributed()
        self.comm_group = comm_group
        self.num_workers = dist.get_world_size(group=comm_group)
        self.worker_id = dist.get_rank(group=comm_group)

    def run_map_reduce(self):
        lo, hi = split_index(0, len(self.dataset), self.num_workers)[self.worker_id]
        loader = DataLoader(Subset(self.dataset, range(int(lo), int(hi))), batch_size=self.batch_size, shuffle=False,
                            collate_fn=self.collate_fn)
        vals = {m: [] for m in self.metric_names}
        for batch in loader:
            for m, fn in zip(self.metric_names, self.metric_functions):
                v = fn(batch)
                vals[m].append(v.detach().to("cpu").reshape(-1).double() if torch.is_tensor(v)
                               else torch.as_tensor(np.asarray(v)).reshape(-1).double())
        ids = torch.arange(int(lo), int(hi), dtype=torch.double)
        if self.sample_indices is not None:
            ids = torch.as_tensor(np.asarray(self.sample_indices[int(lo): int(hi)])).double()
        for m, typ in zip(self.metric_names, self.metric_types):
            local = torch.cat(vals[m]) if vals[m] else torch.zeros(0, dtype=torch.double)
            if typ == ACCUMULATE:
                total = torch.stack(vals[m]).sum(0).to(self.device) if vals[m] else torch.zeros(1, device=self.device)
                dist.all_reduce(total, group=self.comm_group)
                if self.worker_id == 0:
                    os.makedirs(f"{self.save_path}/{m}", exist_ok=True)
                    _write_items(f"{self.save_path}/{m}/{m}_metric_value", [total.cpu().numpy()], np.int64)
                continue
            pair = torch.stack([ids[: len(local)], local], dim=1).to(self.device)
            gathered = Dist.gather_v(pair, 0, self.comm_group, self.num_workers, self.worker_id)
            if self.worker_id == 0:
                allp = torch.cat(gathered).cpu().numpy()
                write_metric_indexes(self.save_path, m, allp[:, 1].astype(np.int64), allp[:, 0].astype(np.int64),
                                     len(self.dataset))
        dist.barrier(group=self.comm_group)


class Dist:
    """Variable-length collectives built from fixed-size ones."""

    @staticmethod
    def min_max(tensor, comm_group):
        lo, hi = tensor.min().clone(), tensor.max().clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=comm_group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=comm_group)
        return lo, hi

    @staticmethod
    def gather_v(tensor, dst, comm_group, num_workers, worker_id):
        """Gather tensors whose dim-0 differs per rank; returns the list on ``dst`` (None elsewhere)."""
        n = torch.tensor([tensor.shape[0]], dtype=torch.int64, device=tensor.device)
        sizes = [torch.zeros_like(n) for _ in range(num_workers)]
        dist.all_gather(sizes, n, group=comm_group)
        sizes = [int(s.item()) for s in sizes]
        mx = max(sizes) if sizes else 0
        padded = tensor.new_zeros((mx,) + tuple(tensor.shape[1:]))
        padded[: tensor.shape[0]] = tensor
        outs = [torch.empty_like(padded) for _ in range(num_workers)]
        dist.all_gather(outs, padded, group=comm_group)    # all_gather keeps gloo/nccl parity; payload is tiny
        if worker_id != dst:
            return None
        return [o[:s] for o, s in zip(outs, sizes)]
