"""On-disk dataset formats of the data-efficiency pipeline: mmap and legacy indexed datasets (build, merge, read back,
slices, pickling) — the formats are shared with Megatron-DeepSpeed tooling, so the header fields are checked too."""
import pickle
import struct

import numpy as np
import pytest
import torch

from shuffle_exchange_b200.runtime.data_pipeline.data_sampling import indexed_dataset as ids


def _docs(seed, n):
    g = np.random.default_rng(seed)
    return [g.integers(0, 30000, size=int(g.integers(1, 40))).astype(np.int32) for _ in range(n)]


def test_mmap_dataset_build_merge_and_read(tmp_path):
    a, b = _docs(0, 7), _docs(1, 5)
    pa, pb = str(tmp_path / "a"), str(tmp_path / "b")
    for docs, prefix in ((a, pa), (b, pb)):
        bld = ids.make_builder(ids.data_file_path(prefix), impl="mmap", vocab_size=30000)
        assert bld._dtype == np.uint16 if hasattr(bld, "_dtype") else True
        for d in docs:
            bld.add_item(torch.from_numpy(d.astype(np.int64)))
            bld.end_document()
        bld.finalize(ids.index_file_path(prefix))
    # merge b into a fresh copy of a
    pm = str(tmp_path / "m")
    bld = ids.make_builder(ids.data_file_path(pm), impl="mmap", vocab_size=30000)
    bld.merge_file_(pa)
    bld.merge_file_(pb)
    bld.finalize(ids.index_file_path(pm))
    assert ids.infer_dataset_impl(pm) == "mmap" and ids.dataset_exists(pm, "mmap")
    ds = ids.make_dataset(pm, impl="mmap", skip_warmup=True)
    want = a + b
    assert len(ds) == len(want) and list(ds.sizes) == [len(d) for d in want]
    for i in (0, 6, 7, 11):
        assert np.array_equal(np.asarray(ds[i]).astype(np.int32), want[i])
    sl = ds[2:5]
    assert len(sl) == 3 and np.array_equal(np.asarray(sl[1]).astype(np.int32), want[3])
    part = ds.get(7, offset=1, length=max(1, len(want[7]) - 1) if len(want[7]) > 1 else None)
    assert np.array_equal(np.asarray(part).astype(np.int32), want[7][1:1 + len(part)])
    assert list(ds.doc_idx) == list(range(len(want) + 1))
    ds2 = pickle.loads(pickle.dumps(ds))                               # DataLoader workers receive a pickled dataset
    assert np.array_equal(np.asarray(ds2[4]), np.asarray(ds[4]))
    # header: magic, version 1, dtype code, counts
    raw = open(ids.index_file_path(pm), "rb").read(34)
    assert raw[:9] == b"MMIDIDX\x00\x00" and struct.unpack("<Q", raw[9:17])[0] == 1
    assert struct.unpack("<B", raw[17:18])[0] == ids.code(np.uint16)
    assert struct.unpack("<QQ", raw[18:34]) == (len(want), len(want) + 1)


def test_legacy_indexed_dataset_round_trip(tmp_path):
    docs = _docs(2, 6)
    prefix = str(tmp_path / "legacy")
    bld = ids.IndexedDatasetBuilder(ids.data_file_path(prefix), dtype=np.int32)
    for d in docs:
        bld.add_item(torch.from_numpy(d))
        bld.end_document()
    bld.finalize(ids.index_file_path(prefix))
    assert ids.infer_dataset_impl(prefix) == "cached"
    ds = ids.make_dataset(prefix, impl="lazy")
    assert isinstance(ds, ids.IndexedDataset) and isinstance(ids.make_dataset(prefix, impl="infer"), ids.IndexedCachedDataset)
    assert isinstance(ids.make_builder(str(tmp_path / "x.bin"), impl="cached"), ids.IndexedDatasetBuilder)
    assert ids.make_dataset(str(tmp_path / "absent"), impl="infer") is None
    assert len(ds) == len(docs)
    for i, d in enumerate(docs):
        assert np.array_equal(np.asarray(ds[i]), d)
    sl = ds[1:4]
    assert len(sl) == 3 and all(np.array_equal(a, d) for a, d in zip(sl, docs[1:4])) and ds[3:3] == []
    assert ds.num_tokens(2) == ds.size(2) == len(docs[2])
    with pytest.raises(ValueError):
        ds[0:4:2]
    cached = ids.IndexedCachedDataset(prefix)
    cached.prefetch([1, 3])
    assert np.array_equal(np.asarray(cached[3]), docs[3])
    with pytest.raises(IndexError):
        ds[len(docs)]


class _SeqDataset(torch.utils.data.Dataset):
    """Sample i is a vector of length 1 + (7 i mod 13): its 'difficulty' is its length."""

    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return torch.full((1 + (7 * i) % 13,), i, dtype=torch.int64)


def _len_metric(batch):
    return torch.tensor([len(x) for x in batch])


def _token_histogram(batch):
    return torch.tensor([sum(len(x) for x in batch), len(batch)])


def _check_analyzer_output(root, n):
    ds = _SeqDataset(n)
    lens = np.array([len(ds[i]) for i in range(n)])
    folder = f"{root}/seqlen"
    s2m = ids.MMapIndexedDataset(f"{folder}/seqlen_sample_to_metric", skip_warmup=True)
    assert [int(s2m[i][0]) for i in range(n)] == lens.tolist()
    i2m = ids.MMapIndexedDataset(f"{folder}/seqlen_index_to_metric", skip_warmup=True)
    i2s = ids.MMapIndexedDataset(f"{folder}/seqlen_index_to_sample", skip_warmup=True)
    values = [int(i2m[k][0]) for k in range(len(i2m))]
    assert values == sorted(set(lens.tolist()))
    for k, v in enumerate(values):
        assert sorted(np.asarray(i2s[k]).tolist()) == np.nonzero(lens == v)[0].tolist()
    pct = ids.MMapIndexedDataset(f"{folder}/seqlen_index_to_sample_percentile_merged", skip_warmup=True)
    merged = np.concatenate([np.asarray(pct[k]) for k in range(len(pct))])
    assert len(pct) == 100 and sorted(merged.tolist()) == list(range(n))
    assert (np.diff(lens[merged]) >= 0).all()                      # percentiles are ordered by difficulty


def test_data_analyzer_two_workers_map_then_reduce(tmp_path):
    from shuffle_exchange_b200.runtime.data_pipeline.data_sampling.data_analyzer import ACCUMULATE, SINGLE, DataAnalyzer
    n, root = 103, str(tmp_path)
    common = dict(batch_size=8, metric_names=["seqlen", "totals"], metric_functions=[_len_metric, _token_histogram],
                  metric_types=[SINGLE, ACCUMULATE], metric_dtypes=[np.int64, np.int64], save_path=root,
                  collate_fn=lambda xs: xs, num_workers=2)
    for w in range(2):
        DataAnalyzer(_SeqDataset(n), worker_id=w, **common).run_map()
    DataAnalyzer(_SeqDataset(n), worker_id=0, **common).run_reduce()
    _check_analyzer_output(root, n)
    tot = ids.MMapIndexedDataset(f"{root}/totals/totals_metric_value", skip_warmup=True)
    assert np.asarray(tot[0]).tolist() == [sum(1 + (7 * i) % 13 for i in range(n)), n]
    pcts = DataAnalyzer.get_metric_value_percentiles("seqlen", {1: 10, 5: 40, 9: 50}, 100)
    assert pcts == {10: 1, 50: 5, 100: 9}


def _dist_analyzer_worker(rank, world, root):
    from shuffle_exchange_b200.runtime.data_pipeline.data_sampling.data_analyzer import SINGLE, DistributedDataAnalyzer
    DistributedDataAnalyzer(_SeqDataset(57), batch_size=4, metric_names=["seqlen"], metric_functions=[_len_metric],
                            metric_types=[SINGLE], save_path=root, collate_fn=lambda xs: xs).run_map_reduce()
    if rank == 0:
        _check_analyzer_output(root, 57)


def test_distributed_data_analyzer_three_ranks(tmp_path):
    from conftest import run_dist
    run_dist(_dist_analyzer_worker, 3, (str(tmp_path),))


def test_variable_batch_size_packing_and_lr_scaling():
    from shuffle_exchange_b200.runtime.data_pipeline.data_sampling import variable_batch_size_and_lr as vb
    seqlens = [5, 9, 3, 12, 7, 2, 8, 4, 6, 30, 1, 10]
    mb_ids, batch_sizes, max_seqlens = vb.batch_by_seqlens(seqlens, max_tokens=16, effective_batch_size=2)
    flat = [i for _, mb in mb_ids for i in mb]
    assert 9 not in flat                                              # 30 tokens can never fit
    assert len(flat) == len(set(flat))                                # nobody is used twice
    for bid, mb in mb_ids:
        assert sum(seqlens[i] for i in mb) <= 16
    assert len(mb_ids) == 2 * len(batch_sizes)                        # two micro-batches per optimizer step
    for b, (size, mx) in enumerate(zip(batch_sizes, max_seqlens)):
        members = [i for bid, mb in mb_ids if bid == b for i in mb]
        assert size == len(members) and mx == max(seqlens[i] for i in members)
    # same-size micro-batches for pipelines
    mb2, _, _ = vb.batch_by_seqlens(seqlens, max_tokens=16, effective_batch_size=2, required_microbatches_of_same_size=True,
                                    sequence_picking_order="seqlen")
    for b in range(len(mb2) // 2):
        assert len(mb2[2 * b][1]) == len(mb2[2 * b + 1][1])
    assert vb.scale_lr(8, 16, 1.0, "linear") == 2.0 and abs(vb.scale_lr(8, 32, 1.0, "sqrt") - 2.0) < 1e-12
    assert vb.scale_lr(8, 16, 0.5, None) == 0.5

    ds = _SeqDataset(len(seqlens))
    w = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.SGD([w], lr=0.1)
    loader, sched = vb.get_dataloader_and_lr_scheduler_for_variable_batch_size(
        dataset=ds, dataset_seqlens=seqlens, max_tokens=16, effective_batch_size=2, base_batch_size=4,
        lr_scheduler_or_optimizer=opt, dataloader_num_workers=0)
    batches = list(loader)
    assert [len(b) for b in batches] == [len(mb) for _, mb in mb_ids]
    assert abs(opt.param_groups[0]["lr"] - 0.1 * batch_sizes[0] / 4) < 1e-12       # scaled for the first batch
    sched.step()
    assert abs(opt.param_groups[0]["lr"] - 0.1 * batch_sizes[1 % len(batch_sizes)] / 4) < 1e-12
    sd = sched.state_dict()
    sched.step()
    sched.load_state_dict(sd)
    assert sched.last_epoch == 1


def test_curriculum_sampler_consumes_the_analyzer_output(tmp_path):
    """Analyzer -> sampler: with a value-based ``seqlen`` curriculum rising from 3 to 13, early batches hold only short
    samples, the eligible pool grows cluster by cluster, ranks get disjoint slices of each global batch, and a resumed
    sampler continues exactly where the saved one would."""
    from shuffle_exchange_b200.runtime.data_pipeline.data_sampling.data_analyzer import SINGLE, DataAnalyzer
    from shuffle_exchange_b200.runtime.data_pipeline.data_sampling.data_sampler import DeepSpeedDataSampler
    n, root = 130, str(tmp_path)
    ds = _SeqDataset(n)
    DataAnalyzer(ds, batch_size=8, metric_names=["seqlen"], metric_functions=[_len_metric], metric_types=[SINGLE],
                 metric_dtypes=[np.int64], save_path=root, collate_fn=lambda xs: xs).run_map_reduce()
    lens = np.array([len(ds[i]) for i in range(n)])
    cfg = {"seed": 7, "data_sampling": {"num_epochs": 50, "curriculum_learning": {
        "enabled": True, "data_cluster_path": f"{root}/clusters", "curriculum_metrics": {"seqlen": {
            "index_to_sample_path": f"{root}/seqlen/seqlen_index_to_sample",
            "index_to_metric_path": f"{root}/seqlen/seqlen_index_to_metric",
            "difficulty_type": "value", "clustering_type": "schedule_based", "min_difficulty": 3, "max_difficulty": 13,
            "schedule_type": "fixed_linear", "schedule_config": {"total_curriculum_step": 10, "difficulty_step": 2}}}}}}

    def sampler(rank):
        return DeepSpeedDataSampler(cfg, n, micro_batch_size=4, data_parallel_rank=rank, data_parallel_size=2,
                                    data_parallel_group=None, gradient_accumulation_steps=1, global_rank=0)

    s0 = sampler(0)
    it = iter(s0)
    seen_max = []
    for step in range(14):
        mine = next(it)
        assert len(mine) == 4 and all(0 <= i < n for i in mine)
        limit = s0.current_difficulties["seqlen"]
        assert lens[mine].max() <= limit, (step, lens[mine].max(), limit)     # nothing harder than the schedule allows
        seen_max.append(int(limit))
    assert seen_max[0] == 3 or seen_max[0] == 4 and seen_max[-1] == 13 and seen_max == sorted(seen_max)
    assert len(s0.clusters) >= 4 and sum(len(c) for c in s0.clusters) == int((lens <= 13).sum())
    union = np.concatenate([c.ids for c in s0.clusters])
    assert len(np.unique(union)) == len(union)                                  # clusters never overlap
    # the two data-parallel ranks split every global batch without overlap
    a, b = sampler(0), sampler(1)
    assert a.get_start_end_idx() == (0, 4) and b.get_start_end_idx() == (4, 8)
    # resume: the restored sampler yields what the original yields next
    sd = s0.state_dict()
    import copy
    sd = copy.deepcopy(sd)
    want = [next(it) for _ in range(3)]
    s1 = sampler(0)
    s1.load_state_dict(sd)
    it1 = iter(s1)
    got = [next(it1) for _ in range(3)]
    assert got == want


def _engine_curriculum_worker(rank, world, root):
    """``initialize(training_data=..., config={"data_efficiency": ...})`` on two ranks: rank 0 draws every global batch
    from the curriculum clusters and broadcasts it, each rank receives its own slice, difficulty grows with the steps,
    the post-process hook sees every batch."""
    import shuffle_exchange_b200 as ds
    from shuffle_exchange_b200.runtime.data_pipeline.data_sampling.data_analyzer import SINGLE, DataAnalyzer
    n = 96
    data = _SeqDataset(n)
    if rank == 0:
        DataAnalyzer(data, batch_size=8, metric_names=["seqlen"], metric_functions=[_len_metric], metric_types=[SINGLE],
                     metric_dtypes=[np.int64], save_path=root, collate_fn=lambda xs: xs).run_map()
        DataAnalyzer(data, batch_size=8, metric_names=["seqlen"], metric_functions=[_len_metric], metric_types=[SINGLE],
                     metric_dtypes=[np.int64], save_path=root, collate_fn=lambda xs: xs).run_reduce()
    torch.distributed.barrier()
    net = torch.nn.Linear(4, 4)
    cfg = {"train_micro_batch_size_per_gpu": 3, "steps_per_print": 10 ** 9,
           "optimizer": {"type": "SGD", "params": {"lr": 0.1}},
           "data_efficiency": {"enabled": True, "seed": 5, "data_sampling": {"enabled": True, "num_epochs": 100,
               "curriculum_learning": {"enabled": True, "data_cluster_path": f"{root}/clusters", "curriculum_metrics": {
                   "seqlen": {"index_to_sample_path": f"{root}/seqlen/seqlen_index_to_sample",
                              "index_to_metric_path": f"{root}/seqlen/seqlen_index_to_metric",
                              "difficulty_type": "value", "clustering_type": "schedule_based", "min_difficulty": 2,
                              "max_difficulty": 13, "schedule_type": "fixed_linear",
                              "schedule_config": {"total_curriculum_step": 6, "difficulty_step": 1}}}}}}}
    engine, _, loader, _ = ds.initialize(model=net, model_parameters=net.parameters(), training_data=data, config=cfg,
                                         collate_fn=lambda xs: xs)
    assert engine.curriculum_learning_enabled() and loader.curriculum_learning_enabled
    hooked = []
    engine.set_data_post_process_func(lambda batch, state: (hooked.append(len(batch)), batch)[1])
    it = iter(loader)
    longest = []
    for step in range(8):
        batch = next(it)
        assert len(batch) == 3                                           # this rank's micro-batch
        ids = torch.tensor([int(x[0]) for x in batch])                   # sample i is filled with the value i
        gathered = [torch.zeros_like(ids) for _ in range(world)]
        torch.distributed.all_gather(gathered, ids)
        assert len(set(torch.cat(gathered).tolist())) == 3 * world      # ranks hold disjoint slices of one global batch
        longest.append(max(len(x) for x in batch))
    bounds = [min(13, int(2 + 11 * k / 6)) for k in range(1, 9)]        # fixed_linear: 3, 5, 7, 9, 11, 13, 13, 13
    assert all(l <= b for l, b in zip(longest, bounds)), (longest, bounds)   # short samples first
    assert len(hooked) == 8
    sampler = loader.data_sampler
    assert sampler.curriculum_step == 8 and sampler.current_difficulties["seqlen"] == 13


def test_engine_level_curriculum_data_loading_two_ranks(tmp_path):
    from conftest import run_dist
    run_dist(_engine_curriculum_worker, 2, (str(tmp_path),))
