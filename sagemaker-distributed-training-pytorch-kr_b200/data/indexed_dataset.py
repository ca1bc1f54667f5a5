"""Memory-mapped token store in Megatron's ``.bin`` / ``.idx`` format (MMapIndexedDataset):
``.idx`` = magic 'MMIDIDX\\0\\0', version u64, dtype code u8, n_sequences u64, n_docs u64, then
sizes int32[n], pointers int64[n], doc_idx int64[n_docs]; ``.bin`` = the concatenated tokens."""
from __future__ import annotations

import os
import struct

import numpy as np

_MAGIC = b"MMIDIDX\x00\x00"
_DTYPES = {1: np.uint8, 2: np.int8, 3: np.int16, 4: np.int32, 5: np.int64, 6: np.float32, 7: np.float64, 8: np.uint16}
_CODES = {v: k for k, v in _DTYPES.items()}


def index_file_path(prefix): return prefix + ".idx"
def data_file_path(prefix): return prefix + ".bin"


class MMapIndexedDatasetBuilder:
    def __init__(self, out_file, dtype=np.int32):
        self._f = open(out_file, "wb")
        self._dtype = dtype
        self._sizes, self._doc_idx = [], [0]

    def add_item(self, tokens):
        arr = np.asarray(tokens, dtype=self._dtype)
        self._f.write(arr.tobytes(order="C"))
        self._sizes.append(arr.size)

    def end_document(self):
        self._doc_idx.append(len(self._sizes))

    def finalize(self, index_file):
        self._f.close()
        sizes = np.asarray(self._sizes, dtype=np.int32)
        item = np.dtype(self._dtype).itemsize
        ptrs = np.zeros(len(sizes), dtype=np.int64)
        if len(sizes) > 1:
            np.cumsum(sizes[:-1].astype(np.int64) * item, out=ptrs[1:])
        with open(index_file, "wb") as f:
            f.write(_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<B", _CODES[np.dtype(self._dtype).type]))
            f.write(struct.pack("<Q", len(sizes)))
            f.write(struct.pack("<Q", len(self._doc_idx)))
            f.write(sizes.tobytes(order="C"))
            f.write(ptrs.tobytes(order="C"))
            f.write(np.asarray(self._doc_idx, dtype=np.int64).tobytes(order="C"))


class MMapIndexedDataset:
    def __init__(self, path_prefix, skip_warmup=True):
        self._path = path_prefix
        with open(index_file_path(path_prefix), "rb") as f:
            assert f.read(9) == _MAGIC, "index file does not match the expected mmap format"
            assert struct.unpack("<Q", f.read(8))[0] == 1
            self._dtype = _DTYPES[struct.unpack("<B", f.read(1))[0]]
            self._len = struct.unpack("<Q", f.read(8))[0]
            self._ndoc = struct.unpack("<Q", f.read(8))[0]
            off = f.tell()
        self._idx_mm = np.memmap(index_file_path(path_prefix), mode="r", order="C")
        buf = memoryview(self._idx_mm)
        self.sizes = np.frombuffer(buf, dtype=np.int32, count=self._len, offset=off)
        self._ptrs = np.frombuffer(buf, dtype=np.int64, count=self._len, offset=off + self.sizes.nbytes)
        self.doc_idx = np.frombuffer(buf, dtype=np.int64, count=self._ndoc, offset=off + self.sizes.nbytes + self._ptrs.nbytes)
        self._bin_mm = np.memmap(data_file_path(path_prefix), mode="r", order="C")
        self._bin = memoryview(self._bin_mm)

    def __len__(self):
        return self._len

    def __getitem__(self, i):
        return np.frombuffer(self._bin, dtype=self._dtype, count=int(self.sizes[i]), offset=int(self._ptrs[i]))

    def get(self, idx, offset=0, length=None):
        size = int(self.sizes[idx])
        if length is None:
            length = size - offset
        ptr = int(self._ptrs[idx]) + offset * np.dtype(self._dtype).itemsize
        return np.frombuffer(self._bin, dtype=self._dtype, count=length, offset=ptr)

    @staticmethod
    def exists(prefix):
        return os.path.exists(index_file_path(prefix)) and os.path.exists(data_file_path(prefix))


# ------------------------------------------------------------------------------------------
# legacy (non-mmap) format: `--data-impl lazy | cached` (megatron/data/indexed_dataset.py IndexedDataset / IndexedCachedDataset,
# inherited from fairseq).  .idx = magic 'TNTIDX\0\0', version u64 (1), dtype code u64, element size u64, n_items u64,
# n_sizes u64, n_docs u64, dim_offsets i64[n_items + 1], data_offsets i64[n_items + 1] (in elements), sizes i64[n_sizes],
# doc_idx i64[n_docs]; .bin = the concatenated items.
# ------------------------------------------------------------------------------------------
_LEGACY_MAGIC = b"TNTIDX\x00\x00"


class IndexedDatasetBuilder:
    def __init__(self, out_file, dtype=np.int32):
        self._f, self._dtype = open(out_file, "wb"), dtype
        self._data_offsets, self._dim_offsets, self._sizes, self._doc_idx = [0], [0], [], [0]
        self._item = np.dtype(dtype).itemsize

    def add_item(self, tokens):
        arr = np.asarray(tokens, dtype=self._dtype)
        self._f.write(arr.tobytes(order="C"))
        self._data_offsets.append(self._data_offsets[-1] + arr.size)
        self._sizes.extend(arr.shape)
        self._dim_offsets.append(self._dim_offsets[-1] + arr.ndim)

    def end_document(self):
        self._doc_idx.append(len(self._data_offsets) - 1)

    def finalize(self, index_file):
        self._f.close()
        with open(index_file, "wb") as f:
            f.write(_LEGACY_MAGIC)
            f.write(struct.pack("<Q", 1))
            f.write(struct.pack("<QQ", _CODES[np.dtype(self._dtype).type], self._item))
            f.write(struct.pack("<QQ", len(self._data_offsets) - 1, len(self._sizes)))
            f.write(struct.pack("<Q", len(self._doc_idx)))
            for arr in (self._dim_offsets, self._data_offsets, self._sizes, self._doc_idx):
                f.write(np.asarray(arr, dtype=np.int64).tobytes(order="C"))


class IndexedDataset:
    """`--data-impl lazy`: every access seeks and reads the .bin file"""

    def __init__(self, path_prefix):
        self._path = path_prefix
        with open(index_file_path(path_prefix), "rb") as f:
            assert f.read(8) == _LEGACY_MAGIC, "index file does not match the legacy (lazy / cached) format"
            assert struct.unpack("<Q", f.read(8))[0] == 1
            code, self._item = struct.unpack("<QQ", f.read(16))
            self._dtype = _DTYPES[code]
            self._len, n_sizes = struct.unpack("<QQ", f.read(16))
            n_docs = struct.unpack("<Q", f.read(8))[0]
            rd = lambda n: np.frombuffer(f.read(8 * n), dtype=np.int64).copy()      # noqa: E731
            self.dim_offsets, self.data_offsets = rd(self._len + 1), rd(self._len + 1)
            self._shapes, self.doc_idx = rd(n_sizes), rd(n_docs)
        self.sizes = np.asarray([int(np.prod(self._shapes[self.dim_offsets[i]:self.dim_offsets[i + 1]])) for i in range(self._len)], dtype=np.int32)
        self._data = None

    def _file(self):
        if self._data is None:
            self._data = open(data_file_path(self._path), "rb", buffering=0)
        return self._data

    def __len__(self):
        return self._len

    def __getitem__(self, i):
        return self.get(i)

    def get(self, idx, offset=0, length=None):
        size = int(self.sizes[idx])
        if length is None:
            length = size - offset
        f = self._file()
        f.seek((int(self.data_offsets[idx]) + offset) * self._item)
        return np.frombuffer(f.read(length * self._item), dtype=self._dtype)

    @staticmethod
    def exists(prefix):
        return os.path.exists(index_file_path(prefix)) and os.path.exists(data_file_path(prefix))


class IndexedCachedDataset(IndexedDataset):
    """`--data-impl cached`: `prefetch(indices)` loads the requested items into one array; reads hit the cache"""

    def __init__(self, path_prefix):
        super().__init__(path_prefix)
        self._cache, self._cache_index = None, {}

    @property
    def supports_prefetch(self):
        return True

    def prefetch(self, indices):
        indices = sorted(set(int(i) for i in indices))
        if all(i in self._cache_index for i in indices):
            return
        total = int(sum(self.sizes[i] for i in indices))
        self._cache, self._cache_index, ptr = np.empty(total, dtype=self._dtype), {}, 0
        for i in indices:
            n = int(self.sizes[i])
            self._cache[ptr:ptr + n] = IndexedDataset.get(self, i)
            self._cache_index[i] = ptr
            ptr += n

    def get(self, idx, offset=0, length=None):
        if idx not in self._cache_index:
            return super().get(idx, offset, length)
        size = int(self.sizes[idx])
        length = size - offset if length is None else length
        p = self._cache_index[idx] + offset
        return self._cache[p:p + length]


def infer_dataset_impl(path):
    """'mmap' / 'cached' (legacy format) from the index file's magic; None if there is no such dataset"""
    if not os.path.exists(index_file_path(path)):
        return None
    with open(index_file_path(path), "rb") as f:
        magic = f.read(9)
    if magic == _MAGIC:
        return "mmap"
    if magic[:8] == _LEGACY_MAGIC:
        return "cached"
    return None


def make_builder(out_file, impl="mmap", dtype=np.int32):
    return MMapIndexedDatasetBuilder(out_file, dtype) if impl == "mmap" else IndexedDatasetBuilder(out_file, dtype)


def make_dataset(path, impl="mmap", skip_warmup=True):
    """`--data-impl {mmap, lazy, cached, infer}` (reference: arguments.py:1156)"""
    if not MMapIndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        return None
    if impl == "infer":
        impl = infer_dataset_impl(path)
    if impl == "lazy":
        return IndexedDataset(path)
    if impl == "cached":
        return IndexedCachedDataset(path)
    return MMapIndexedDataset(path, skip_warmup)
