"""Curriculum-aware batch sampler of the data-efficiency pipeline.

Reference: ``data_sampling/data_sampler.py:36-349``.  Behaviour kept:

* one global batch (``micro_batch * dp * gas`` sample ids) is drawn at a time on global rank 0 and broadcast over
  the data-parallel group; each ``__iter__`` step yields this rank's slice of one micro-batch-times-dp chunk;
* with curriculum learning every metric has its own :class:`CurriculumScheduler`; whenever a difficulty changes, the
  samples that just became eligible (value-based: metric in (previous, current]; percentile-based: that slice of the
  difficulty-sorted sample list; several clustered metrics: intersection minus what is already clustered) form a new
  shuffled *cluster* persisted as an mmap dataset under ``data_cluster_path``; batches draw from all clusters with
  probability proportional to cluster size, each cluster being consumed sequentially and reshuffled when exhausted;
* ``state_dict`` carries batch remainder, consumed samples, curriculum step, difficulties, cluster file names and
  read positions, and the RNG state.

Without curriculum learning the sampler degenerates to a seeded shuffle of ``range(one_epoch_total_samples)`` per
epoch (the reference leaves that case to the plain dataloader).
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .... import comm as dist
from ....utils.logging import logger
from ..constants import *  # noqa: F401,F403
from ..curriculum_scheduler import CurriculumScheduler
from .indexed_dataset import MMapIndexedDataset
from .utils import close_mmap_dataset_builder, create_mmap_dataset_builder, find_fit_int_dtype


class _Cluster:
    """A shuffled list of sample ids on disk + a read cursor."""

    def __init__(self, root: str, fname: str, position: int = 0):
        self.root, self.fname, self.position = root, fname, position
        self.ids = np.array(MMapIndexedDataset(self.path, skip_warmup=True)[0])

    @property
    def path(self):
        return f"{self.root}/{self.fname}"

    @staticmethod
    def create(root, fname, ids, dtype):
        path = f"{root}/{fname}"
        builder = create_mmap_dataset_builder(path, dtype)
        builder.add_item_numpy(ids)
        close_mmap_dataset_builder(builder, path)
        return _Cluster(root, fname)

    def reshuffle(self, rng, dtype):
        ids = self.ids.copy()
        rng.shuffle(ids)
        builder = create_mmap_dataset_builder(self.path, dtype)
        builder.add_item_numpy(ids)
        close_mmap_dataset_builder(builder, self.path)
        self.ids = ids

    def take(self, n, rng, dtype):
        out = list(self.ids[self.position: self.position + n])
        self.position += n
        if len(out) < n:
            rest = n - len(out)
            logger.info(f"reshuffling cluster {self.fname}.")
            self.reshuffle(rng, dtype)
            out += list(self.ids[:rest])
            self.position = rest
        return out

    def __len__(self):
        return len(self.ids)


class DeepSpeedDataSampler:
    def __init__(self, data_efficiency_config, one_epoch_total_samples, micro_batch_size, data_parallel_rank,
                 data_parallel_size, data_parallel_group, gradient_accumulation_steps, global_rank, drop_last=True):
        self.data_efficiency_config = data_efficiency_config
        self.one_epoch_total_samples = one_epoch_total_samples
        self.index_dtype = find_fit_int_dtype(0, one_epoch_total_samples)
        ds_cfg = data_efficiency_config[DATA_SAMPLING]
        self.total_samples = one_epoch_total_samples * ds_cfg.get(DATA_SAMPLING_NUM_EPOCHS, DATA_SAMPLING_NUM_EPOCHS_DEFAULT)
        self.micro_batch_size = micro_batch_size
        self.data_parallel_rank, self.data_parallel_size = data_parallel_rank, data_parallel_size
        self.data_parallel_group = data_parallel_group
        self.micro_batch_times_data_parallel_size = micro_batch_size * data_parallel_size
        self.gradient_accumulation_steps = gradient_accumulation_steps
        self.global_batch_size = self.micro_batch_times_data_parallel_size * gradient_accumulation_steps
        self.global_rank, self.drop_last = global_rank, drop_last
        self.np_rng = np.random.default_rng(data_efficiency_config.get(DATA_EFFICIENCY_SEED, DATA_EFFICIENCY_SEED_DEFAULT))
        self.batch, self.consumed_samples = [], 0
        self._epoch_perm, self._epoch_pos = None, 0
        cl = ds_cfg.get(CURRICULUM_LEARNING, {})
        self.curriculum_enabled = bool(cl.get(CURRICULUM_LEARNING_ENABLED, False))
        self.curriculum_step = 0
        self.current_difficulties = {}
        self.clusters = []                       # rank 0 only
        self.data_cluster_paths, self.data_cluster_current_position = [], []
        self.curriculum_schedulers, self.difficulty_type, self.clustering_type = {}, {}, {}
        self.curriculum_index_to_sample, self.curriculum_index_to_metric = {}, {}
        self.data_1epoch_size = None
        if self.curriculum_enabled:
            self.cluster_root = cl[CURRICULUM_LEARNING_CLUSTER_PATH]
            if global_rank == 0:
                os.makedirs(self.cluster_root, exist_ok=True)
            for metric, mcfg in cl[CURRICULUM_LEARNING_METRICS].items():
                self.curriculum_schedulers[metric] = CurriculumScheduler(mcfg)
                self.difficulty_type[metric] = mcfg[CURRICULUM_LEARNING_DIFFICULTY_TYPE]
                self.clustering_type[metric] = mcfg[CURRICULUM_LEARNING_CLUSTERING_TYPE]
                if global_rank == 0 and self.clustering_type[metric] != CURRICULUM_LEARNING_SINGLE_CLUSTER:
                    self.curriculum_index_to_sample[metric] = MMapIndexedDataset(mcfg[CURRICULUM_LEARNING_SAMPLE_PATH],
                                                                                 skip_warmup=True)
                    if self.difficulty_type[metric] == CURRICULUM_LEARNING_VALUE_BASED:
                        self.curriculum_index_to_metric[metric] = MMapIndexedDataset(
                            mcfg[CURRICULUM_LEARNING_METRIC_PATH], skip_warmup=True)
        assert self.total_samples > 0, f"no sample to consume: {self.total_samples}"
        assert micro_batch_size > 0 and data_parallel_size > 0
        assert data_parallel_rank < data_parallel_size, \
            f"data_parallel_rank should be smaller than data size: {data_parallel_rank}, {data_parallel_size}"

    def __len__(self):
        return self.total_samples

    def set_custom_curriculum_learning_schedule(self, schedule_func_dict):
        for metric, sched in self.curriculum_schedulers.items():
            if metric in schedule_func_dict:
                sched.set_custom_get_difficulty(schedule_func_dict[metric])

    def get_start_end_idx(self, batch_len=None):
        """Slice of a chunk of ``batch_len`` samples that belongs to this data-parallel rank."""
        batch_len = batch_len or self.micro_batch_times_data_parallel_size
        edge = lambda r: round(r * batch_len / self.data_parallel_size)
        return edge(self.data_parallel_rank), edge(self.data_parallel_rank + 1)

    # ------------------------------------------------------------------ eligibility queries (rank 0)
    def get_sample_based_on_metric_value(self, metric, value_start, value_end):
        to_sample, to_metric = self.curriculum_index_to_sample[metric], self.curriculum_index_to_metric[metric]
        picked = [np.array(to_sample[row]) for row in range(len(to_sample))
                  if value_start < to_metric[row] <= value_end]
        return np.concatenate(picked, axis=None) if picked else None

    def get_sample_based_on_metric_percentile(self, metric, percentile_start, percentile_end):
        to_sample = self.curriculum_index_to_sample[metric]
        if self.data_1epoch_size is None:
            self.data_1epoch_size = int(sum(len(to_sample[r]) for r in range(len(to_sample))))
        max_pct = self.data_efficiency_config[DATA_SAMPLING][CURRICULUM_LEARNING][CURRICULUM_LEARNING_METRICS][metric][
            CURRICULUM_LEARNING_MAX_DIFFICULTY]
        per_pct = self.data_1epoch_size // max_pct
        lo = per_pct * percentile_start
        hi = self.data_1epoch_size if percentile_end == max_pct else per_pct * percentile_end
        picked, seen = [], 0
        for row in range(len(to_sample)):
            n = len(to_sample[row])
            if seen + n > lo:
                a = max(0, lo - seen)
                b = n if seen + n <= hi else hi - seen
                picked.append(np.array(to_sample[row][a:b]))
            seen += n
            if seen >= hi:
                break
        return np.concatenate(picked, axis=None) if picked else None

    def _eligible(self, metric, lo, hi):
        if self.difficulty_type[metric] == CURRICULUM_LEARNING_VALUE_BASED:
            return self.get_sample_based_on_metric_value(metric, lo, hi)
        return self.get_sample_based_on_metric_percentile(metric, lo, hi)

    def get_new_cluster(self, previous_difficulties):
        fname = CURRICULUM_LEARNING_CLUSTER_PREFIX + "".join(
            f"_{m}{self.current_difficulties[m]}" for m in self.curriculum_schedulers)
        if self.global_rank == 0:
            clustered = [m for m, t in self.clustering_type.items() if t != CURRICULUM_LEARNING_SINGLE_CLUSTER]
            new = None
            if len(clustered) > 1:
                for m in self.curriculum_schedulers:
                    if self.clustering_type[m] == CURRICULUM_LEARNING_SINGLE_CLUSTER:
                        part = np.arange(self.one_epoch_total_samples, dtype=self.index_dtype)
                    else:
                        floor = float("-inf") if self.difficulty_type[m] == CURRICULUM_LEARNING_VALUE_BASED else 0
                        part = self._eligible(m, floor, self.current_difficulties[m])
                    new = part if new is None else np.intersect1d(new, part, assume_unique=True)
                for c in self.clusters:
                    new = np.setdiff1d(new, c.ids, assume_unique=True)
            else:
                if not self.clusters:
                    new = np.arange(self.one_epoch_total_samples, dtype=self.index_dtype)
                for m in clustered:
                    new = self._eligible(m, previous_difficulties[m], self.current_difficulties[m])
            if new is not None and len(new) > 0:
                logger.info(f"new data cluster (previous_difficulties {previous_difficulties}, current_difficulties "
                            f"{self.current_difficulties}) with size {len(new)} generated.")
                new = np.array(new, dtype=self.index_dtype)
                self.np_rng.shuffle(new)
                self.clusters.append(_Cluster.create(self.cluster_root, fname, new, self.index_dtype))
            else:
                logger.info(f"new data cluster (previous_difficulties {previous_difficulties}, current_difficulties "
                            f"{self.current_difficulties}) has no matched data thus skipped.")
        if dist.is_initialized():
            dist.barrier(group=self.data_parallel_group)
        if os.path.isfile(f"{self.cluster_root}/{fname}.bin") and fname not in self.data_cluster_paths:
            self.data_cluster_paths.append(fname)
            self.data_cluster_current_position.append(0)

    def sample_from_clusters(self):
        sizes = np.array([len(c) for c in self.clusters], dtype=np.float64)
        draws = self.np_rng.choice(len(self.clusters), self.global_batch_size, replace=True, p=sizes / sizes.sum())
        return np.bincount(draws, minlength=len(self.clusters))

    # --------------------------------------------------------------------------- batches
    def _plain_batch(self):
        out = []
        while len(out) < self.global_batch_size:
            if self._epoch_perm is None or self._epoch_pos >= len(self._epoch_perm):
                self._epoch_perm = self.np_rng.permutation(self.one_epoch_total_samples)
                self._epoch_pos = 0
            take = min(self.global_batch_size - len(out), len(self._epoch_perm) - self._epoch_pos)
            out += self._epoch_perm[self._epoch_pos: self._epoch_pos + take].tolist()
            self._epoch_pos += take
        return out

    def get_next_global_batch(self):
        if not self.curriculum_enabled:
            self.batch = self._plain_batch()      # same seed on every rank -> identical without communication
            return
        self.curriculum_step += 1
        changed, previous = False, {}
        for m, sched in self.curriculum_schedulers.items():
            nxt = sched.update_difficulty(self.curriculum_step)
            if m not in self.current_difficulties or nxt != self.current_difficulties[m]:
                changed = True
            if m in self.current_difficulties:
                previous[m] = self.current_difficulties[m]
            else:
                previous[m] = float("-inf") if self.difficulty_type[m] == CURRICULUM_LEARNING_VALUE_BASED else 0
            self.current_difficulties[m] = nxt
        if changed:
            self.get_new_cluster(previous)
        multi = dist.is_initialized() and self.data_parallel_size > 1
        device = torch.device("cuda", torch.cuda.current_device()) if (multi and dist.get_backend() == "nccl") else "cpu"
        if self.global_rank == 0:
            batch = []
            for cidx, n in enumerate(self.sample_from_clusters()):
                batch += self.clusters[cidx].take(int(n), self.np_rng, self.index_dtype)
                self.data_cluster_current_position[cidx] = self.clusters[cidx].position
            self.np_rng.shuffle(batch)
            assert len(batch) <= self.global_batch_size
            batch = batch + [-1] * (self.global_batch_size - len(batch))   # fixed shape for the broadcast
            t = torch.tensor(batch, device=device, dtype=torch.long).view(-1)
        else:
            t = torch.empty(self.global_batch_size, device=device, dtype=torch.long)
        if multi:
            dist.broadcast(t, 0, group=self.data_parallel_group)
        self.batch = t[t != -1].tolist()

    def __iter__(self):
        chunk = self.micro_batch_times_data_parallel_size
        while self.consumed_samples <= self.total_samples:
            if len(self.batch) == 0:
                self.get_next_global_batch()
            current, self.batch = self.batch[:chunk], self.batch[chunk:]
            if len(current) == chunk or (len(current) > 0 and not self.drop_last):
                a, b = self.get_start_end_idx(len(current))
                yield current[a:b]
                self.consumed_samples += len(current)

    # -------------------------------------------------------------------------- checkpoint
    def state_dict(self):
        return {
            CURRICULUM_LEARNING_BATCH: self.batch,
            CURRICULUM_LEARNING_CONSUMED_SAMPLES: self.consumed_samples,
            CURRICULUM_LEARNING_STEP: self.curriculum_step,
            CURRICULUM_LEARNING_CURRENT_DIFFICULTIES: self.current_difficulties,
            CURRICULUM_LEARNING_DATA_CLUSTER_PATHS: self.data_cluster_paths,
            CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION: self.data_cluster_current_position,
            CURRICULUM_LEARNING_NP_RNG_STATE: self.np_rng.bit_generator.state,
        }

    def load_state_dict(self, state_dict):
        self.batch = state_dict[CURRICULUM_LEARNING_BATCH]
        self.consumed_samples = state_dict[CURRICULUM_LEARNING_CONSUMED_SAMPLES]
        self.curriculum_step = state_dict[CURRICULUM_LEARNING_STEP]
        self.current_difficulties = state_dict[CURRICULUM_LEARNING_CURRENT_DIFFICULTIES]
        # older checkpoints stored absolute paths; keep only the file name so moved cluster folders still load
        self.data_cluster_paths = [p.split("/")[-1] for p in state_dict[CURRICULUM_LEARNING_DATA_CLUSTER_PATHS]]
        self.data_cluster_current_position = list(state_dict[CURRICULUM_LEARNING_DATA_CLUSTER_CURRENT_POSITION])
        rng_state = state_dict[CURRICULUM_LEARNING_NP_RNG_STATE]
        if isinstance(rng_state, dict):
            self.np_rng.bit_generator.state = rng_state
        for m, sched in self.curriculum_schedulers.items():
            if m in self.current_difficulties:
                sched.set_current_difficulty(self.current_difficulties[m])
        if self.global_rank == 0 and self.curriculum_enabled:
            self.clusters = [_Cluster(self.cluster_root, f, pos)
                             for f, pos in zip(self.data_cluster_paths, self.data_cluster_current_position)]
