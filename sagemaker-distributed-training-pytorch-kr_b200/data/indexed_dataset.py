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


def make_dataset(path, impl="mmap", skip_warmup=True):
    if not MMapIndexedDataset.exists(path):
        print(f"Dataset does not exist: {path}")
        return None
    return MMapIndexedDataset(path, skip_warmup)
