"""Memory-mapped indexed dataset (Megatron ``MMIDIDX`` on-disk format) used by the data-efficiency pipeline for
both token data and the per-sample metric / cluster index files.

Reference: ``data_sampling/indexed_dataset.py:369-627`` (``MMapIndexedDataset`` + ``MMapIndexedDatasetBuilder``).
File layout kept byte-compatible so existing ``.bin/.idx`` pairs load:

    .idx : b"MMIDIDX\\0\\0" | <Q version=1> | <B dtype code> | <Q n_items> | <Q n_docs>
           | int32 sizes[n_items] | int64 byte pointers[n_items] | int64 doc_idx[n_docs]
    .bin : the items back to back

The legacy (non-mmap, ``TNTIDX``) format is kept as ``IndexedDataset`` / ``IndexedCachedDataset`` /
``IndexedDatasetBuilder`` further down for files written that way; ``make_dataset`` / ``make_builder`` produce the
``mmap`` form (``infer_dataset_impl`` reports anything else as unsupported).
"""
from __future__ import annotations

import os
import shutil
import struct

import numpy as np
import torch

_MAGIC = b"MMIDIDX\x00\x00"
_DTYPES = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float64, 7: np.float32,
           8: np.uint16, 9: np.uint32, 10: np.uint64}


def code(dtype) -> int:
    for k, v in _DTYPES.items():
        if v == dtype:
            return k
    raise ValueError(dtype)


def index_file_path(prefix_path: str) -> str:
    return prefix_path + ".idx"


def data_file_path(prefix_path: str) -> str:
    return prefix_path + ".bin"


def get_available_dataset_impl():
    return ["lazy", "cached", "mmap"]


def infer_dataset_impl(path):
    """Which reader the files at ``path`` (+ ``.idx`` / ``.bin``) need: the index magic tells the two formats apart."""
    if not IndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        print("Path should be a basename that both .idx and .bin can be appended to get full filenames.")
        return None
    with open(index_file_path(path), "rb") as f:
        magic = f.read(9)
    if magic[:8] == _LEGACY_MAGIC:
        return "cached"
    return "mmap" if magic == _MAGIC else None


def dataset_exists(path, impl="mmap"):
    return MMapIndexedDataset.exists(path) if impl == "mmap" else IndexedDataset.exists(path)


def make_builder(out_file, impl="mmap", vocab_size=None):
    if impl == "mmap":
        dtype = np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32
        return MMapIndexedDatasetBuilder(out_file, dtype=dtype)
    return IndexedDatasetBuilder(out_file)


def make_dataset(path, impl="mmap", skip_warmup=False):
    if not IndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        print("Path should be a basename that both .idx and .bin can be appended to get full filenames.")
        return None
    if impl == "infer":
        impl = infer_dataset_impl(path)
    if impl == "lazy":
        return IndexedDataset(path)
    if impl == "cached":
        return IndexedCachedDataset(path)
    if impl == "mmap":
        return MMapIndexedDataset(path, skip_warmup)
    print(f"Unknown dataset implementation: {impl}")
    return None


class _Index:
    def __init__(self, path: str):
        with open(path, "rb") as f:
            if f.read(9) != _MAGIC:
                raise ValueError(f"{path} is not an MMIDIDX index file")
            (version,) = struct.unpack("<Q", f.read(8))
            assert version == 1
            (dcode,) = struct.unpack("<B", f.read(1))
            self.dtype = _DTYPES[dcode]
            (self.n,) = struct.unpack("<Q", f.read(8))
            (self.n_docs,) = struct.unpack("<Q", f.read(8))
            offset = f.tell()
        self._mm = np.memmap(path, mode="r", order="C")
        buf = memoryview(self._mm)
        self.sizes = np.frombuffer(buf, dtype=np.int32, count=self.n, offset=offset)
        self.pointers = np.frombuffer(buf, dtype=np.int64, count=self.n, offset=offset + self.sizes.nbytes)
        self.doc_idx = np.frombuffer(buf, dtype=np.int64, count=self.n_docs,
                                     offset=offset + self.sizes.nbytes + self.pointers.nbytes)

    @staticmethod
    def write(path: str, dtype, sizes, doc_idx):
        sizes = np.asarray(sizes, dtype=np.int32)
        itemsize = np.dtype(dtype).itemsize
        pointers = np.zeros(len(sizes), dtype=np.int64)
        if len(sizes) > 1:
            np.cumsum(sizes[:-1].astype(np.int64) * itemsize, out=pointers[1:])
        with open(path, "wb") as f:
            f.write(_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", code(dtype)))
            f.write(struct.pack("<Q", len(sizes)))
            f.write(struct.pack("<Q", len(doc_idx)))
            f.write(sizes.tobytes(order="C"))
            f.write(pointers.tobytes(order="C"))
            f.write(np.asarray(doc_idx, dtype=np.int64).tobytes(order="C"))


class MMapIndexedDataset(torch.utils.data.Dataset):
    def __init__(self, path, skip_warmup=False):
        super().__init__()
        self._path = None
        self._do_init(path, skip_warmup)

    def __getstate__(self):
        return self._path

    def __setstate__(self, state):
        self._do_init(state, skip_warmup=True)

    def _do_init(self, path, skip_warmup):
        self._path = path
        self._index = _Index(index_file_path(path))
        if not skip_warmup:
            with open(data_file_path(path), "rb") as f:      # touch the pages once
                while f.read(100 * 1024 * 1024):
                    pass
        self._bin = np.memmap(data_file_path(path), mode="r", order="C")
        self._buf = memoryview(self._bin)

    def __len__(self):
        return self._index.n

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            ptr, size = self._index.pointers[idx], self._index.sizes[idx]
            return np.frombuffer(self._buf, dtype=self._index.dtype, count=size, offset=ptr)
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            ptr = self._index.pointers[start]
            sizes = self._index.sizes[idx]
            flat = np.frombuffer(self._buf, dtype=self._index.dtype, count=int(sum(sizes)), offset=ptr)
            return np.split(flat, np.cumsum(sizes)[:-1])
        raise TypeError(f"bad index {idx!r}")

    def get(self, idx, offset=0, length=None):
        ptr, size = self._index.pointers[idx], self._index.sizes[idx]
        length = size - offset if length is None else length
        ptr += offset * np.dtype(self._index.dtype).itemsize
        return np.frombuffer(self._buf, dtype=self._index.dtype, count=length, offset=ptr)

    @property
    def sizes(self):
        return self._index.sizes

    def size(self, index):
        return self._index.sizes[index]

    @property
    def doc_idx(self):
        return self._index.doc_idx

    def get_doc_idx(self):
        return self._index.doc_idx

    def set_doc_idx(self, doc_idx_):
        self._index.doc_idx = doc_idx_

    @property
    def supports_prefetch(self):
        return False

    @property
    def dtype(self):
        return self._index.dtype

    @staticmethod
    def exists(path):
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))


class MMapIndexedDatasetBuilder:
    def __init__(self, out_file, dtype=np.int64):
        self._data_file = open(out_file, "wb")
        self._dtype = dtype
        self._sizes = []
        self._doc_idx = [0]

    def add_item(self, tensor):
        self.add_item_numpy(np.asarray(tensor.numpy() if torch.is_tensor(tensor) else tensor))

    def add_items(self, arr_list):
        for a in arr_list:
            self.add_item_numpy(np.asarray(a))

    def add_item_numpy(self, np_array):
        arr = np_array.astype(self._dtype, copy=False)
        self._data_file.write(arr.tobytes(order="C"))
        self._sizes.append(arr.size)

    def end_document(self):
        self._doc_idx.append(len(self._sizes))

    def merge_file_(self, another_file):
        index = _Index(index_file_path(another_file))
        assert index.dtype == self._dtype
        total = len(self._sizes)
        self._sizes.extend(int(s) for s in index.sizes)
        self._doc_idx.extend(int(d) + total for d in index.doc_idx[1:])
        with open(data_file_path(another_file), "rb") as f:
            shutil.copyfileobj(f, self._data_file)

    def finalize(self, index_file):
        self._data_file.close()
        _Index.write(index_file, self._dtype, self._sizes, self._doc_idx)


# ================================================================ helpers of the legacy ("lazy" / "cached") format
# Reference ``indexed_dataset.py:60-330``: the pre-mmap fairseq format — ``TNTIDX`` header, explicit ``sizes`` /
# ``data_offsets`` / ``dim_offsets`` arrays in the index, items read with ``file.readinto``.  Kept so datasets written
# by older tooling can still be read and produced; new data should use the mmap builder above.
_LEGACY_MAGIC = b"TNTIDX\x00\x00"


def read_longs(f, n):
    a = np.empty(n, dtype=np.int64)
    f.readinto(a)
    return a


def write_longs(f, a):
    f.write(np.array(a, dtype=np.int64))


def create_doc_idx(sizes):
    """Item indices at which documents start: an item of size 0 terminates a document."""
    doc_idx = [0]
    for i, s in enumerate(sizes):
        if s == 0:
            doc_idx.append(i + 1)
    return doc_idx


def exscan_from_cumsum_(arr):
    """In place: an inclusive prefix sum becomes the exclusive one (shift right, first element 0)."""
    if arr.size > 1:
        arr[1:] = arr[:-1]
    if arr.size > 0:
        arr[0] = 0


def get_pointers_with_total(sizes, elemsize, dtype):
    """Byte offset of every item given item lengths (elements) and the element size; also the total byte count."""
    pointers = np.array(sizes, dtype=dtype)
    pointers *= elemsize
    np.cumsum(pointers, axis=0, out=pointers)
    total = int(pointers[-1]) if len(sizes) > 0 else 0
    exscan_from_cumsum_(pointers)
    return pointers, total


class IndexedDatasetBuilder:
    """Writer of the legacy format: ``add_item`` appends a tensor, an item of size 0 (``end_document``) closes a
    document, ``finalize(index_file)`` writes the index."""
    element_sizes = {np.uint8: 1, np.int8: 1, np.int16: 2, np.int32: 4, np.int64: 8, np.float32: 4, np.float64: 8,
                     np.uint16: 2}

    def __init__(self, out_file, dtype=np.int32):
        self.out_file = open(out_file, "wb")
        self.dtype = dtype
        self.data_offsets, self.dim_offsets, self.sizes, self.doc_idx = [0], [0], [], [0]
        self.element_size = self.element_sizes[self.dtype]

    def add_item(self, tensor):
        arr = np.array(tensor.numpy() if torch.is_tensor(tensor) else tensor, dtype=self.dtype)
        nbytes = self.out_file.write(arr)
        self.data_offsets.append(self.data_offsets[-1] + nbytes // self.element_size)
        for s in arr.shape:
            self.sizes.append(s)
        self.dim_offsets.append(self.dim_offsets[-1] + arr.ndim)

    def end_document(self):
        self.doc_idx.append(len(self.sizes))

    def merge_file_(self, another_file):
        other = IndexedDataset(another_file)
        assert other.dtype == self.dtype
        base = self.data_offsets[-1]
        self.data_offsets.extend(base + o for o in other.data_offsets[1:])
        self.sizes.extend(other.sizes)
        dbase = self.dim_offsets[-1]
        self.dim_offsets.extend(dbase + o for o in other.dim_offsets[1:])
        ibase = len(self.sizes) - len(other.sizes)
        self.doc_idx.extend(ibase + d for d in other.doc_idx[1:])
        with open(data_file_path(another_file), "rb") as f:
            while True:
                chunk = f.read(1 << 20)
                if not chunk:
                    break
                self.out_file.write(chunk)

    def finalize(self, index_file):
        self.out_file.close()
        import struct
        with open(index_file, "wb") as f:
            f.write(_LEGACY_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<QQ", code(self.dtype), self.element_size))
            f.write(struct.pack("<QQ", len(self.data_offsets) - 1, len(self.sizes)))
            f.write(struct.pack("<Q", len(self.doc_idx)))
            write_longs(f, self.dim_offsets)
            write_longs(f, self.data_offsets)
            write_longs(f, self.sizes)
            write_longs(f, self.doc_idx)


class IndexedDataset(torch.utils.data.Dataset):
    """Reader of the legacy format (items read from the ``.bin`` file on demand)."""

    def __init__(self, path):
        super().__init__()
        self.path = path
        import struct
        with open(index_file_path(path), "rb") as f:
            magic = f.read(8)
            assert magic == _LEGACY_MAGIC, "Index file doesn't match expected format. Make sure that --dataset-impl is configured properly."
            assert struct.unpack("<Q", f.read(8)) == (1,)
            dcode, self.element_size = struct.unpack("<QQ", f.read(16))
            self.dtype = _DTYPES[dcode]
            self._len, n_sizes = struct.unpack("<QQ", f.read(16))
            n_docs = struct.unpack("<Q", f.read(8))[0]
            self.dim_offsets = read_longs(f, self._len + 1)
            self.data_offsets = read_longs(f, self._len + 1)
            self.sizes = read_longs(f, n_sizes)
            self.doc_idx = read_longs(f, n_docs)
        self.data_file = None

    def _open(self):
        if self.data_file is None:
            self.data_file = open(data_file_path(self.path), "rb", buffering=0)

    def __len__(self):
        return self._len

    def check_index(self, i):
        if not 0 <= i < self._len:
            raise IndexError("index out of range")

    def read_data(self, path=None):
        """Open the data file (kept lazy so that the object pickles into DataLoader workers)."""
        self._open()

    def __getitem__(self, i):
        if isinstance(i, slice):                        # contiguous items: one read, then cut (reference :205-217)
            start, stop, step = i.indices(self._len)
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            if stop <= start:
                return []
            self._open()
            counts = [int(self.data_offsets[k + 1] - self.data_offsets[k]) for k in range(start, stop)]
            flat = np.empty(sum(counts), dtype=self.dtype)
            self.data_file.seek(int(self.data_offsets[start]) * self.element_size)
            self.data_file.readinto(flat)
            return np.split(flat, np.cumsum(counts)[:-1])
        self.check_index(i)
        self._open()
        shape = tuple(int(s) for s in self.sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]])
        a = np.empty(shape, dtype=self.dtype)
        self.data_file.seek(int(self.data_offsets[i]) * self.element_size)
        self.data_file.readinto(a)
        return a

    def num_tokens(self, index):
        return int(self.sizes[index])

    def size(self, index):
        return int(self.sizes[index])

    def __del__(self):
        if getattr(self, "data_file", None):
            self.data_file.close()

    @staticmethod
    def exists(path):
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))

    @property
    def supports_prefetch(self):
        return False


class IndexedCachedDataset(IndexedDataset):
    """Legacy reader with an in-memory cache filled by ``prefetch(indices)`` (one contiguous buffer)."""

    def __init__(self, path):
        super().__init__(path)
        self.cache = None
        self.cache_index = {}

    @property
    def supports_prefetch(self):
        return True

    def prefetch(self, indices):
        if all(i in self.cache_index for i in indices):
            return
        self._open()
        indices = sorted(set(indices))
        total = sum(int(self.data_offsets[i + 1] - self.data_offsets[i]) for i in indices)
        self.cache = np.empty(total, dtype=self.dtype)
        self.cache_index.clear()
        ptx = 0
        for i in indices:
            self.cache_index[i] = ptx
            size = int(self.data_offsets[i + 1] - self.data_offsets[i])
            a = self.cache[ptx: ptx + size]
            self.data_file.seek(int(self.data_offsets[i]) * self.element_size)
            self.data_file.readinto(a)
            ptx += size
        self.data_file.close()
        self.data_file = None

    def __getitem__(self, i):
        if i not in self.cache_index:
            return super().__getitem__(i)
        shape = tuple(int(s) for s in self.sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]])
        ptx = self.cache_index[i]
        n = int(np.prod(shape)) if shape else 1
        return self.cache[ptx: ptx + n].reshape(shape).copy()
