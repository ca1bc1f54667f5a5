"""Sample-mapping builders for sentence-level datasets (role of megatron/data/dataset_utils.py::get_samples_mapping and
realm_dataset_utils.py::get_block_samples_mapping; the C++ side is Megatron's helpers.cpp).

`build_mapping` / `build_blocks_mapping` call csrc/host/data_helpers.cpp; the `*_py` functions are pure-Python twins used as
test oracles (they reproduce the rows before the final shuffle: the C++ shuffle draws from a 64-bit Mersenne twister).
The GPT workload of the reference uses `build_sample_idx` / `build_blending_indices` (data/gpt_dataset.py); these two
complete the helper set for BERT-style and retrieval-style datasets stored in the same `.bin/.idx` format.
"""
from __future__ import annotations

import ctypes

import numpy as np

from ..utils import hostlib

LONG_SENTENCE_LEN = 512


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def build_mapping(docs, sizes, num_epochs: int, max_num_samples: int, max_seq_length: int, short_seq_prob: float, seed: int,
                  min_num_sent: int = 2, shuffle: bool = True) -> np.ndarray:
    """docs: int64 [n_docs + 1] sentence boundaries; sizes: int32 tokens per sentence -> int64 [n, 3] rows
    (first sentence, one-past-last sentence, target sequence length)."""
    docs = np.ascontiguousarray(docs, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    L = hostlib.lib()
    L.b2h_build_mapping.restype = ctypes.c_int64
    args = (_ptr(docs), ctypes.c_int64(len(docs) - 1), _ptr(sizes), ctypes.c_int32(num_epochs), ctypes.c_int64(max_num_samples),
            ctypes.c_int32(max_seq_length), ctypes.c_double(short_seq_prob), ctypes.c_int32(seed), ctypes.c_int32(min_num_sent))
    n = L.b2h_build_mapping(*args, None, ctypes.c_int64(0), ctypes.c_int32(0))
    out = np.zeros((n, 3), dtype=np.int64)
    got = L.b2h_build_mapping(*args, _ptr(out), ctypes.c_int64(n), ctypes.c_int32(int(shuffle)))
    assert got == n
    return out


def build_blocks_mapping(docs, sizes, titles_sizes, num_epochs: int, max_num_samples: int, max_seq_length: int, seed: int,
                         use_one_sent_blocks: bool = False, shuffle: bool = True) -> np.ndarray:
    """-> int64 [n, 4] rows (first sentence, one-past-last sentence, document, block id)."""
    docs = np.ascontiguousarray(docs, dtype=np.int64)
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    titles_sizes = np.ascontiguousarray(titles_sizes, dtype=np.int32)
    assert len(titles_sizes) == len(docs) - 1
    L = hostlib.lib()
    L.b2h_build_blocks_mapping.restype = ctypes.c_int64
    args = (_ptr(docs), ctypes.c_int64(len(docs) - 1), _ptr(sizes), _ptr(titles_sizes), ctypes.c_int32(num_epochs),
            ctypes.c_int64(max_num_samples), ctypes.c_int32(max_seq_length), ctypes.c_int32(seed), ctypes.c_int32(int(use_one_sent_blocks)))
    n = L.b2h_build_blocks_mapping(*args, None, ctypes.c_int64(0), ctypes.c_int32(0))
    out = np.zeros((n, 4), dtype=np.int64)
    got = L.b2h_build_blocks_mapping(*args, _ptr(out), ctypes.c_int64(n), ctypes.c_int32(int(shuffle)))
    assert got == n
    return out


class _MT32:
    """raw 32-bit draws of the classic MT19937 seeding (what std::mt19937(seed) produces)"""

    def __init__(self, seed: int):
        self.rs = np.random.RandomState(seed & 0xFFFFFFFF)

    def __call__(self) -> int:
        return int(self.rs.randint(0, 2 ** 32, dtype=np.uint64))


def _target_len(ratio, max_len, rng):
    if ratio == 0:
        return max_len
    r = rng()
    return 2 + r % (max_len - 1) if r % ratio == 0 else max_len


def build_mapping_py(docs, sizes, num_epochs, max_num_samples, max_seq_length, short_seq_prob, seed, min_num_sent=2):
    ratio = int(round(1.0 / short_seq_prob)) if short_seq_prob > 0 else 0
    rng = _MT32(seed)
    rows = []
    for _ in range(num_epochs):
        if len(rows) >= max_num_samples:
            break
        for d in range(len(docs) - 1):
            first, last = int(docs[d]), int(docs[d + 1])
            remain = last - first
            if remain < max(min_num_sent, 2) or any(sizes[s] > LONG_SENTENCE_LEN for s in range(first, last)):
                continue
            prev, seq_len, num_sent = first, 0, 0
            target = _target_len(ratio, max_seq_length, rng)
            for s in range(first, last):
                seq_len += int(sizes[s]); num_sent += 1; remain -= 1
                if (seq_len >= target and remain > 1 and num_sent >= min_num_sent) or remain == 0:
                    rows.append((prev, s + 1, target))
                    prev, seq_len, num_sent = s + 1, 0, 0
                    target = _target_len(ratio, max_seq_length, rng)
    return np.asarray(rows, dtype=np.int64).reshape(-1, 3)


def build_blocks_mapping_py(docs, sizes, titles_sizes, num_epochs, max_num_samples, max_seq_length, seed, use_one_sent_blocks=False):
    mn = 1 if use_one_sent_blocks else 2
    rows = []
    for _ in range(num_epochs):
        block = 0
        if len(rows) >= max_num_samples:
            break
        for d in range(len(docs) - 1):
            first, last = int(docs[d]), int(docs[d + 1])
            target = max_seq_length - int(titles_sizes[d])
            remain = last - first
            if remain < mn or any(sizes[s] > LONG_SENTENCE_LEN for s in range(first, last)):
                continue
            prev, seq_len, num_sent = first, 0, 0
            for s in range(first, last):
                seq_len += int(sizes[s]); num_sent += 1; remain -= 1
                if (seq_len >= target and remain >= mn and num_sent >= mn) or remain == 0:
                    rows.append((prev, s + 1, d, block))
                    block += 1
                    prev, seq_len, num_sent = s + 1, 0, 0
    return np.asarray(rows, dtype=np.int64).reshape(-1, 4)
