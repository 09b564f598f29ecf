"""On-disk token datasets, byte-compatible with the reference's ``.bin`` / ``.idx`` pairs
(megatron/data/indexed_dataset.py:24-585 surface).

Two index formats:
* ``MMIDIDX`` (mmap): magic 9 B, version u64 = 1, dtype code u8, number of sequences u64,
  number of documents u64, then sizes int32[n], byte pointers int64[n], doc_idx int64[d].
* ``TNTIDX`` (legacy lazy / cached): magic 8 B, version u64, dtype code u64, element size
  u64, n u64, s u64, doc_count u64, dim_offsets int64[n+1], data_offsets int64[n+1],
  sizes int64[s], doc_idx int64[doc_count].
"""
from __future__ import annotations

import os
import shutil
import struct
from functools import lru_cache
from itertools import accumulate

import numpy as np
import torch

from ..utils import print_rank_0

DTYPES = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float32,
          7: np.float64, 8: np.uint16}
_MMAP_MAGIC = b"MMIDIDX\x00\x00"
_LEGACY_MAGIC = b"TNTIDX\x00\x00"


def best_fitting_dtype(vocab_size=None):
    return np.uint16 if vocab_size is not None and vocab_size < 65500 else np.int32


__best_fitting_dtype = best_fitting_dtype


def code(dtype):
    for k, v in DTYPES.items():
        if v == dtype:
            return k
    raise ValueError(dtype)


def index_file_path(prefix):
    return prefix + ".idx"


def data_file_path(prefix):
    return prefix + ".bin"


def get_available_dataset_impl():
    return ["lazy", "cached", "mmap"]


def infer_dataset_impl(path):
    if not IndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        print("Path should be a basename that both .idx and .bin can be appended to get full filenames.")
        return None
    with open(index_file_path(path), "rb") as f:
        magic = f.read(8)
    if magic == _LEGACY_MAGIC:
        return "cached"
    if magic == _MMAP_MAGIC[:8]:
        return "mmap"
    return None


def make_builder(out_file, impl, vocab_size=None):
    if impl == "mmap":
        return MMapIndexedDatasetBuilder(out_file, dtype=best_fitting_dtype(vocab_size))
    return IndexedDatasetBuilder(out_file)


def make_dataset(path, impl, skip_warmup=False):
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
    if impl == "mmap" and MMapIndexedDataset.exists(path):
        return MMapIndexedDataset(path, skip_warmup)
    print(f"Unknown dataset implementation: {impl}")
    return None


def dataset_exists(path, impl):
    return MMapIndexedDataset.exists(path) if impl == "mmap" else IndexedDataset.exists(path)


def _read_longs(f, n):
    a = np.empty(n, dtype=np.int64)
    f.readinto(a)
    return a


class IndexedDataset(torch.utils.data.Dataset):
    """Legacy format, read lazily with file seeks."""

    def __init__(self, path):
        super().__init__()
        self.path = path
        self.data_file = None
        with open(index_file_path(path), "rb") as f:
            assert f.read(8) == _LEGACY_MAGIC, "Index file doesn't match expected format."
            assert struct.unpack("<Q", f.read(8)) == (1,)
            c, self.element_size = struct.unpack("<QQ", f.read(16))
            self.dtype = DTYPES[c]
            self._len, self.s = struct.unpack("<QQ", f.read(16))
            self.doc_count = struct.unpack("<Q", f.read(8))[0]
            self.dim_offsets = _read_longs(f, self._len + 1)
            self.data_offsets = _read_longs(f, self._len + 1)
            self.sizes = _read_longs(f, self.s)
            self.doc_idx = _read_longs(f, self.doc_count)

    def read_data(self, path):
        self.data_file = open(data_file_path(path), "rb", buffering=0)

    def check_index(self, i):
        if i < 0 or i >= self._len:
            raise IndexError("index out of range")

    def __del__(self):
        if self.data_file:
            self.data_file.close()

    def _read(self, i):
        if not self.data_file:
            self.read_data(self.path)
        self.check_index(i)
        shape = self.sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]]
        a = np.empty(shape, dtype=self.dtype)
        self.data_file.seek(self.data_offsets[i] * self.element_size)
        self.data_file.readinto(a)
        return a

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            return self._read(int(idx))
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            return [self._read(i) for i in range(start, stop)]
        raise TypeError(type(idx))

    def __len__(self):
        return self._len

    def num_tokens(self, index):
        return self.sizes[index]

    def size(self, index):
        return self.sizes[index]

    @staticmethod
    def exists(path):
        return os.path.exists(index_file_path(path)) and os.path.exists(data_file_path(path))

    @property
    def supports_prefetch(self):
        return False


class IndexedCachedDataset(IndexedDataset):
    """Legacy format with an in-memory cache filled by ``prefetch``."""

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
        if not self.data_file:
            self.read_data(self.path)
        indices = sorted(set(indices))
        total = sum(self.data_offsets[i + 1] - self.data_offsets[i] for i in indices)
        self.cache = np.empty(total, dtype=self.dtype)
        self.cache_index.clear()
        ptx = 0
        for i in indices:
            self.cache_index[i] = ptx
            n = self.data_offsets[i + 1] - self.data_offsets[i]
            self.data_file.seek(self.data_offsets[i] * self.element_size)
            self.data_file.readinto(self.cache[ptx:ptx + n])
            ptx += n
        self.data_file.close()
        self.data_file = None

    def _read(self, i):
        if self.cache is None or i not in self.cache_index:
            return super()._read(i)
        shape = self.sizes[self.dim_offsets[i]:self.dim_offsets[i + 1]]
        n = int(np.prod(shape))
        p = self.cache_index[i]
        return self.cache[p:p + n].reshape(shape).copy()


class IndexedDatasetBuilder:
    element_sizes = {np.uint8: 1, np.int8: 1, np.int16: 2, np.int32: 4, np.int64: 8, np.float32: 4,
                     np.float64: 8}

    def __init__(self, out_file, dtype=np.int32):
        self.out_file = open(out_file, "wb")
        self.dtype = dtype
        self.data_offsets, self.dim_offsets, self.sizes = [0], [0], []
        self.element_size = self.element_sizes[self.dtype]
        self.doc_idx = [0]

    def add_item(self, tensor):
        a = np.array(tensor.numpy() if hasattr(tensor, "numpy") else tensor, dtype=self.dtype)
        self.out_file.write(a.tobytes(order="C"))
        self.data_offsets.append(self.data_offsets[-1] + a.size)
        self.sizes.extend(a.shape)
        self.dim_offsets.append(self.dim_offsets[-1] + len(a.shape))

    def end_document(self):
        self.doc_idx.append(len(self.sizes))

    def merge_file_(self, another_file):
        index = IndexedDataset(another_file)
        assert index.dtype == self.dtype
        begin = self.data_offsets[-1]
        self.data_offsets.extend(begin + o for o in index.data_offsets[1:])
        self.sizes.extend(index.sizes)
        begin = self.dim_offsets[-1]
        self.dim_offsets.extend(begin + o for o in index.dim_offsets[1:])
        with open(data_file_path(another_file), "rb") as f:
            shutil.copyfileobj(f, self.out_file)

    def finalize(self, index_file):
        self.out_file.close()
        with open(index_file, "wb") as idx:
            idx.write(_LEGACY_MAGIC)
            idx.write(struct.pack("<Q", 1))
            idx.write(struct.pack("<QQ", code(self.dtype), self.element_size))
            idx.write(struct.pack("<QQ", len(self.data_offsets) - 1, len(self.sizes)))
            idx.write(struct.pack("<Q", len(self.doc_idx)))
            for arr in (self.dim_offsets, self.data_offsets, self.sizes, self.doc_idx):
                idx.write(np.array(arr, dtype=np.int64).tobytes(order="C"))


def _warmup_mmap_file(path):
    with open(path, "rb") as stream:
        while stream.read(100 * 1024 * 1024):
            pass


class MMapIndexedDataset(torch.utils.data.Dataset):
    class Index:
        _HDR_MAGIC = _MMAP_MAGIC

        @classmethod
        def writer(cls, path, dtype):
            class _Writer:
                def __enter__(self):
                    self._file = open(path, "wb")
                    self._file.write(cls._HDR_MAGIC)
                    self._file.write(struct.pack("<Q", 1))
                    self._file.write(struct.pack("<B", code(dtype)))
                    return self

                @staticmethod
                def _get_pointers(sizes):
                    itemsize = np.dtype(dtype).itemsize
                    ptrs = np.zeros(len(sizes), dtype=np.int64)
                    if len(sizes) > 1:
                        np.cumsum(np.asarray(sizes[:-1], dtype=np.int64) * itemsize, out=ptrs[1:])
                    return ptrs

                def write(self, sizes, doc_idx):
                    pointers = self._get_pointers(sizes)
                    self._file.write(struct.pack("<Q", len(sizes)))
                    self._file.write(struct.pack("<Q", len(doc_idx)))
                    self._file.write(np.array(sizes, dtype=np.int32).tobytes(order="C"))
                    self._file.write(pointers.tobytes(order="C"))
                    self._file.write(np.array(doc_idx, dtype=np.int64).tobytes(order="C"))

                def __exit__(self, exc_type, exc_val, exc_tb):
                    self._file.close()
            return _Writer()

        def __init__(self, path, skip_warmup=False):
            with open(path, "rb") as stream:
                assert stream.read(9) == self._HDR_MAGIC, "Index file doesn't match expected format."
                assert struct.unpack("<Q", stream.read(8)) == (1,)
                (dtype_code,) = struct.unpack("<B", stream.read(1))
                self._dtype = DTYPES[dtype_code]
                self._dtype_size = self._dtype().itemsize
                self._len = struct.unpack("<Q", stream.read(8))[0]
                self._doc_count = struct.unpack("<Q", stream.read(8))[0]
                offset = stream.tell()
            if not skip_warmup:
                print_rank_0("    warming up index mmap file...")
                _warmup_mmap_file(path)
            self._bin_buffer_mmap = np.memmap(path, mode="r", order="C")
            self._bin_buffer = memoryview(self._bin_buffer_mmap)
            print_rank_0("    reading sizes...")
            self._sizes = np.frombuffer(self._bin_buffer, dtype=np.int32, count=self._len, offset=offset)
            print_rank_0("    reading pointers...")
            self._pointers = np.frombuffer(self._bin_buffer, dtype=np.int64, count=self._len,
                                           offset=offset + self._sizes.nbytes)
            print_rank_0("    reading document index...")
            self._doc_idx = np.frombuffer(self._bin_buffer, dtype=np.int64, count=self._doc_count,
                                          offset=offset + self._sizes.nbytes + self._pointers.nbytes)

        def __del__(self):
            if hasattr(self, "_bin_buffer_mmap"):
                self._bin_buffer_mmap._mmap.close()
                del self._bin_buffer_mmap

        dtype = property(lambda self: self._dtype)
        sizes = property(lambda self: self._sizes)
        doc_idx = property(lambda self: self._doc_idx)

        @lru_cache(maxsize=8)
        def __getitem__(self, i):
            return self._pointers[i], self._sizes[i]

        def __len__(self):
            return self._len

    def __init__(self, path, skip_warmup=False):
        super().__init__()
        self._path = self._index = self._bin_buffer = None
        self._do_init(path, skip_warmup)

    def __getstate__(self):
        return self._path

    def __setstate__(self, state):
        self._do_init(state, skip_warmup=True)

    def _do_init(self, path, skip_warmup):
        self._path = path
        self._index = self.Index(index_file_path(path), skip_warmup)
        if not skip_warmup:
            print_rank_0("    warming up data mmap file...")
            _warmup_mmap_file(data_file_path(path))
        print_rank_0("    creating numpy buffer of mmap...")
        self._bin_buffer_mmap = np.memmap(data_file_path(path), mode="r", order="C")
        print_rank_0("    creating memory view of numpy buffer...")
        self._bin_buffer = memoryview(self._bin_buffer_mmap)

    def __del__(self):
        if getattr(self, "_bin_buffer_mmap", None) is not None:
            self._bin_buffer_mmap._mmap.close()
            del self._bin_buffer_mmap
        self._index = None

    def __len__(self):
        return len(self._index)

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            ptr, size = self._index[int(idx)]
            return np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=size, offset=ptr)
        if isinstance(idx, slice):
            start, stop, step = idx.indices(len(self))
            if step != 1:
                raise ValueError("Slices into indexed_dataset must be contiguous")
            ptr = self._index._pointers[start]
            sizes = self._index._sizes[idx]
            offsets = list(accumulate(sizes))
            total = sum(sizes)
            arr = np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=total, offset=ptr)
            return np.split(arr, offsets[:-1])
        raise TypeError(type(idx))

    def get(self, idx, offset=0, length=None):
        """Zero-copy slice of one sequence."""
        ptr, size = self._index[int(idx)]
        if length is None:
            length = size - offset
        ptr += offset * np.dtype(self._index.dtype).itemsize
        return np.frombuffer(self._bin_buffer, dtype=self._index.dtype, count=length, offset=ptr)

    sizes = property(lambda self: self._index.sizes)
    doc_idx = property(lambda self: self._index.doc_idx)

    def size(self, index):
        return self._index.sizes[index]

    def get_doc_idx(self):
        return self._index._doc_idx

    def set_doc_idx(self, doc_idx_):
        self._index._doc_idx = doc_idx_

    @property
    def supports_prefetch(self):
        return False

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
        a = np.array(tensor.numpy() if hasattr(tensor, "numpy") else tensor, dtype=self._dtype)
        self._data_file.write(a.tobytes(order="C"))
        self._sizes.append(a.size)

    def end_document(self):
        self._doc_idx.append(len(self._sizes))

    def merge_file_(self, another_file):
        index = MMapIndexedDataset.Index(index_file_path(another_file), skip_warmup=True)
        assert index.dtype == self._dtype
        offset = len(self._sizes)
        self._sizes.extend(index.sizes)
        self._doc_idx.extend((offset + np.asarray(index.doc_idx))[1:].tolist())
        with open(data_file_path(another_file), "rb") as f:
            shutil.copyfileobj(f, self._data_file)

    def finalize(self, index_file):
        self._data_file.close()
        with MMapIndexedDataset.Index.writer(index_file, self._dtype) as index:
            index.write(self._sizes, self._doc_idx)
