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
    cached = ids.IndexedCachedDataset(prefix)
    cached.prefetch([1, 3])
    assert np.array_equal(np.asarray(cached[3]), docs[3])
    with pytest.raises(IndexError):
        ds[len(docs)]
