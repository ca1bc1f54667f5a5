"""GPT pre-training datasets.

* ``GPTDataset`` - samples of ``seq_length + 1`` tokens stitched across documents through three index
  arrays (doc_idx / sample_idx / shuffle_idx) cached under ``<data dir>/index-cache`` with md5-derived
  names; single / blended / per-split data paths; document-range splits.  Behavioural counterpart of the
  reference's 3_training_megatron-lm/megatron/data/gpt_dataset.py (build_train_valid_test_datasets :20-113,
  GPTDataset :248-311, _build_index_mappings :314-494, numpy helpers :497-601).  `sample_idx` comes from
  the C++ builder (csrc/host/data_helpers.cpp); ``build_sample_idx_py`` is the Python twin / oracle.
* ``MockGPTDataset`` - synthetic token stream of the configured shape (BASELINE.json: synthetic data).

SageMaker edit kept: the index cache is built by LOCAL_RANK 0 of every node (gpt_dataset.py:371-372) and
the build is confirmed collectively over the DP then PP groups (:462-469).
"""
from __future__ import annotations

import ctypes
import hashlib
import os
import time
from typing import List, Optional

import numpy as np
import torch
import torch.distributed as dist

from ..parallel import state as mpu
from ..utils import hostlib
from .indexed_dataset import make_dataset


def _p0(msg):
    if not dist.is_initialized() or dist.get_rank() == 0:
        print(msg, flush=True)


# ------------------------------------------------------------------------------------------ splits
def get_train_valid_test_split_(splits_string: str, size: int) -> List[int]:
    """'969,30,1' -> document index boundaries [0, a, b, size]."""
    if "," in splits_string:
        splits = [float(s) for s in splits_string.split(",")]
    elif "/" in splits_string:
        splits = [float(s) for s in splits_string.split("/")]
    else:
        splits = [float(splits_string)]
    while len(splits) < 3:
        splits.append(0.0)
    splits = splits[:3]
    tot = sum(splits)
    assert tot > 0.0
    splits = [s / tot for s in splits]
    idx = [0]
    for s in splits:
        idx.append(idx[-1] + int(round(s * float(size))))
    diff = idx[-1] - size
    for i in range(1, len(idx)):
        idx[i] -= diff
    assert len(idx) == 4 and idx[-1] == size
    return idx


def get_datasets_weights_and_num_samples(data_prefix, train_valid_test_num_samples):
    assert len(data_prefix) % 2 == 0
    n = len(data_prefix) // 2
    weights = [float(data_prefix[2 * i]) for i in range(n)]
    prefixes = [data_prefix[2 * i + 1].strip() for i in range(n)]
    s = sum(weights)
    assert s > 0.0
    weights = [w / s for w in weights]
    if isinstance(train_valid_test_num_samples, (list, tuple)):
        per = [[int(np.ceil(v * w * 1.005)) for v in train_valid_test_num_samples] for w in weights]
    else:
        per = [int(np.ceil(train_valid_test_num_samples * w * 1.005)) for w in weights]
    return prefixes, weights, per


# ------------------------------------------------------------------------------------------ index builders
def num_tokens_per_epoch(documents, sizes) -> int:
    return int(np.sum(sizes[documents]))


def num_epochs_for(tokens_per_epoch: int, seq_length: int, num_samples: int) -> int:
    epochs, total = 0, 0
    while True:
        epochs += 1
        total += tokens_per_epoch
        if (total - 1) // seq_length >= num_samples:
            return epochs


def build_doc_idx(documents, num_epochs, np_rng, separate_last_epoch):
    if not separate_last_epoch or num_epochs == 1:
        d = np.mgrid[0:num_epochs, 0:len(documents)][1]
        d[:] = documents
        d = d.reshape(-1).astype(np.int32)
        np_rng.shuffle(d)
        return d
    first = build_doc_idx(documents, num_epochs - 1, np_rng, False)
    last = build_doc_idx(documents, 1, np_rng, False)
    return np.concatenate((first, last))


def build_sample_idx_py(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch):
    """Python twin of b2h_build_sample_idx (oracle; same contract as gpt_dataset.py:533-579)."""
    num_samples = (num_epochs * tokens_per_epoch - 1) // seq_length
    out = np.zeros([num_samples + 1, 2], dtype=np.int32)
    si, di, off = 1, 0, 0
    while si <= num_samples:
        remaining = seq_length + 1
        while remaining != 0:
            doc_len = sizes[doc_idx[di]] - off
            remaining -= doc_len
            if remaining <= 0:
                off += remaining + doc_len - 1
                remaining = 0
            else:
                di += 1
                off = 0
        out[si, 0], out[si, 1] = di, off
        si += 1
    return out


def build_sample_idx(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch):
    sizes = np.ascontiguousarray(sizes, dtype=np.int32)
    doc_idx = np.ascontiguousarray(doc_idx, dtype=np.int32)
    num_samples = (num_epochs * tokens_per_epoch - 1) // seq_length
    out = np.zeros([num_samples + 1, 2], dtype=np.int32)
    L = hostlib.lib()
    L.b2h_build_sample_idx.restype = ctypes.c_int64
    rows = L.b2h_build_sample_idx(sizes.ctypes.data_as(ctypes.c_void_p), doc_idx.ctypes.data_as(ctypes.c_void_p),
                                  ctypes.c_int64(doc_idx.size), ctypes.c_int32(seq_length), ctypes.c_int32(num_epochs),
                                  ctypes.c_int64(tokens_per_epoch), out.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(out.shape[0]))
    assert rows == num_samples + 1, "sample index build failed (token stream shorter than requested)"
    return out


def build_shuffle_idx(num_samples, total_size, np_rng):
    dtype = np.uint32 if total_size < (np.iinfo(np.uint32).max - 1) else np.int64
    first = np.arange(0, num_samples, dtype=dtype)
    np_rng.shuffle(first)
    if num_samples == total_size:
        return first
    last = np.arange(num_samples, total_size, dtype=dtype)
    np_rng.shuffle(last)
    return np.concatenate((first, last))


def build_blending_indices(weights, size):
    w = np.ascontiguousarray(weights, dtype=np.float64)
    di = np.zeros(size, dtype=np.uint8)
    dsi = np.zeros(size, dtype=np.int64)
    hostlib.lib().b2h_build_blending_indices(di.ctypes.data_as(ctypes.c_void_p), dsi.ctypes.data_as(ctypes.c_void_p),
                                            w.ctypes.data_as(ctypes.c_void_p), ctypes.c_int32(len(weights)), ctypes.c_int64(size))
    return di, dsi


def _build_index_mappings(name, data_prefix, documents, sizes, splits_string, num_samples, seq_length, seed, data_cache_path=None):
    tokens_per_epoch = num_tokens_per_epoch(documents, sizes)
    num_epochs = num_epochs_for(tokens_per_epoch, seq_length, num_samples)
    np_rng = np.random.RandomState(seed=seed)
    desc = f"GPT Dataset\n\nData prefix {data_prefix}\nDataset name {name}\nNumber of samples {num_samples}\nSequence length {seq_length}\nRandom seed {seed}\nSplit {splits_string}\n"
    h = hashlib.md5(desc.encode("utf-8")).hexdigest()
    fns = {k: f"{h}_{k}" for k in ("desc", "doc", "sample", "shuffle")}
    fns["desc"] += ".dsc"
    for k in ("doc", "sample", "shuffle"):
        fns[k] += "_idx.npy"
    prefixes = [os.path.join(os.path.dirname(data_prefix), "index-cache")]
    if data_cache_path is not None:
        prefixes.append(data_cache_path)
    build, found_prefix = True, None
    for pref in prefixes:
        if all(os.path.isfile(os.path.join(pref, f)) for f in fns.values()):
            build, found_prefix = False, pref
            break
    # no cache found: write to the LAST place looked in - the --data-cache-path when one is given (a read-only data channel
    # keeps working), else <data dir>/index-cache (reference gpt_dataset.py:345-366)
    cache_dir = prefixes[-1] if found_prefix is None else found_prefix
    data_cache_success = True
    # every node builds its own cache on its local rank 0 (SageMaker edit, node-local storage)
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if build and local_rank == 0:
        _p0(" > WARNING: could not find index map files, building the indices on local rank 0 ...")
        if num_epochs == 1:
            separate_last_epoch = False
        else:
            from_prev = ((num_epochs - 1) * tokens_per_epoch - 1) // seq_length
            last_epoch_samples = num_samples - from_prev
            per_epoch = (tokens_per_epoch - 1) // seq_length
            assert 0 <= last_epoch_samples <= per_epoch + 1
            separate_last_epoch = last_epoch_samples < int(0.80 * per_epoch)
        try:
            os.makedirs(cache_dir, exist_ok=True)
            with open(os.path.join(cache_dir, fns["desc"]), "wt") as f:
                f.write(desc)
            t0 = time.time()
            doc_idx = build_doc_idx(documents, num_epochs, np_rng, separate_last_epoch)
            np.save(os.path.join(cache_dir, fns["doc"]), doc_idx, allow_pickle=True)
            _p0(" > elasped time to build and save doc-idx mapping (seconds): {:4f}".format(time.time() - t0))
            t0 = time.time()
            assert doc_idx.dtype == np.int32 and sizes.dtype == np.int32
            sample_idx = build_sample_idx(sizes, doc_idx, seq_length, num_epochs, tokens_per_epoch)
            np.save(os.path.join(cache_dir, fns["sample"]), sample_idx, allow_pickle=True)
            _p0(" > elasped time to build and save sample-idx mapping (seconds): {:4f}".format(time.time() - t0))
            t0 = time.time()
            n_ = (((num_epochs - 1) * tokens_per_epoch - 1) // seq_length) if separate_last_epoch else sample_idx.shape[0] - 1
            shuffle_idx = build_shuffle_idx(n_, sample_idx.shape[0] - 1, np_rng)
            np.save(os.path.join(cache_dir, fns["shuffle"]), shuffle_idx, allow_pickle=True)
            _p0(" > elasped time to build and save shuffle-idx mapping (seconds): {:4f}".format(time.time() - t0))
        except OSError:
            print(f"There was an error trying to create the data cache directory ({cache_dir}); set --data-cache-path to a writable location.")
            data_cache_success = False
    if dist.is_initialized() and dist.get_world_size() > 1:
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
        counts = torch.tensor([int(data_cache_success)], dtype=torch.long, device=dev)
        dist.all_reduce(counts, group=mpu.get_data_parallel_group())
        dist.all_reduce(counts, group=mpu.get_pipeline_model_parallel_group())
        expect = dist.get_world_size() // max(1, mpu.get_tensor_model_parallel_world_size())
        if counts[0].item() != expect:
            _p0("Data index creation unsuccessful, exiting.")
            raise SystemExit(1)
    elif not data_cache_success:
        _p0("Data index creation unsuccessful, exiting.")
        raise SystemExit(1)
    t0 = time.time()
    doc_idx = np.load(os.path.join(cache_dir, fns["doc"]), allow_pickle=True, mmap_mode="r")
    sample_idx = np.load(os.path.join(cache_dir, fns["sample"]), allow_pickle=True, mmap_mode="r")
    shuffle_idx = np.load(os.path.join(cache_dir, fns["shuffle"]), allow_pickle=True, mmap_mode="r")
    _p0("    loaded indexed file in {:3.3f} seconds".format(time.time() - t0))
    _p0("    total number of samples: {}".format(sample_idx.shape[0]))
    _p0("    total number of epochs: {}".format(num_epochs))
    return doc_idx, sample_idx, shuffle_idx, desc, h


class GPTDataset(torch.utils.data.Dataset):
    def __init__(self, name, data_prefix, documents, indexed_dataset, splits_string, num_samples, seq_length, seed,
                 return_doc_ids=False, data_cache_path=None):
        self.name, self.indexed_dataset, self.return_doc_ids = name, indexed_dataset, return_doc_ids
        assert np.min(documents) >= 0 and np.max(documents) < indexed_dataset.sizes.shape[0]
        self.doc_idx, self.sample_idx, self.shuffle_idx, self.desc, self.desc_hash = _build_index_mappings(
            name, data_prefix, documents, np.asarray(indexed_dataset.sizes), splits_string, num_samples, seq_length, seed, data_cache_path)

    def __len__(self):
        return self.sample_idx.shape[0] - 1

    def __getitem__(self, idx):
        idx = self.shuffle_idx[idx]
        f, l = self.sample_idx[idx], self.sample_idx[idx + 1]
        doc_f, doc_l, off_f, off_l = int(f[0]), int(l[0]), int(f[1]), int(l[1])
        ids = []
        if doc_f == doc_l:
            ids.append(self.doc_idx[doc_f])
            sample = self.indexed_dataset.get(self.doc_idx[doc_f], offset=off_f, length=off_l - off_f + 1)
        else:
            ids.append(self.doc_idx[doc_f])
            parts = [self.indexed_dataset.get(self.doc_idx[doc_f], offset=off_f)]
            for i in range(doc_f + 1, doc_l):
                ids.append(self.doc_idx[i])
                parts.append(self.indexed_dataset.get(self.doc_idx[i]))
            ids.append(self.doc_idx[doc_l])
            parts.append(self.indexed_dataset.get(self.doc_idx[doc_l], length=off_l + 1))
            sample = np.concatenate(parts)
        out = {"text": np.array(sample, dtype=np.int64)}
        if self.return_doc_ids:
            out["doc_ids"] = np.array(ids, dtype=np.int64)
        return out


class BlendableDataset(torch.utils.data.Dataset):
    def __init__(self, datasets, weights, size):
        self.datasets, self.size = datasets, size
        w = np.array(weights, dtype=np.float64)
        w /= w.sum()
        self.dataset_index, self.dataset_sample_index = build_blending_indices(w, size)

    def __len__(self):
        return self.size

    def __getitem__(self, idx):
        d, s = int(self.dataset_index[idx]), int(self.dataset_sample_index[idx])
        return self.datasets[d][s % len(self.datasets[d])]


class MockGPTDataset(torch.utils.data.Dataset):
    """Deterministic synthetic samples: {'text': int64[seq_length + 1]} with ids in [0, vocab)."""
    in_memory = True

    def __init__(self, num_samples: int, seq_length: int, vocab_size: int, seed: int = 1234):
        self.n, self.s, self.v, self.seed = int(num_samples), seq_length, vocab_size, seed

    def __len__(self):
        return self.n

    def __getitem__(self, idx):
        g = np.random.RandomState(self.seed + int(idx))
        return {"text": g.randint(0, self.v, size=self.s + 1).astype(np.int64)}


def _build_one(data_prefix, data_impl, splits_string, nums, seq_length, seed, skip_warmup, return_doc_ids=False, data_cache_path=None):
    indexed = make_dataset(data_prefix, data_impl, skip_warmup)
    assert indexed is not None, f"dataset {data_prefix} not found"
    total_docs = indexed.sizes.shape[0]
    splits = get_train_valid_test_split_(splits_string, total_docs)
    _p0(" > dataset split:")
    for i, nm in enumerate(("train", "validation", "test")):
        _p0("    {}:\n     document indices in [{}, {}) total of {} documents".format(nm, splits[i], splits[i + 1], splits[i + 1] - splits[i]))

    def build(i, nm):
        if splits[i + 1] <= splits[i]:
            return None
        docs = np.arange(splits[i], splits[i + 1], dtype=np.int32)
        return GPTDataset(nm, data_prefix, docs, indexed, splits_string, nums[i], seq_length, seed, return_doc_ids, data_cache_path)
    return build(0, "train"), build(1, "valid"), build(2, "test")


def build_train_valid_test_datasets(data_prefix, data_impl, splits_string, train_valid_test_num_samples, seq_length, seed, skip_warmup,
                                    train_data_prefix=None, valid_data_prefix=None, test_data_prefix=None, return_doc_ids=False,
                                    data_cache_path=None, mock: bool = False, vocab_size: int = 32768):
    """single / blended (`w1 p1 w2 p2 ...`) / separate per-split prefixes (gpt_dataset.py:20-113) or mock data."""
    nums = train_valid_test_num_samples
    if mock or (not data_prefix and not train_data_prefix):
        _p0(" > using synthetic (mock) GPT data of shape [seq_length + 1]")
        return tuple(MockGPTDataset(max(n, 1), seq_length, vocab_size, seed + i) for i, n in enumerate(nums))
    if data_prefix:
        if len(data_prefix) == 1:
            return _build_one(data_prefix[0], data_impl, splits_string, nums, seq_length, seed, skip_warmup, return_doc_ids, data_cache_path)
        prefixes, weights, per = get_datasets_weights_and_num_samples(data_prefix, nums)
        tot = [sum(p[i] for p in per) for i in range(3)]
        tr, va, te = [], [], []
        for i, pref in enumerate(prefixes):
            a, b, c = _build_one(pref, data_impl, splits_string, per[i], seq_length, seed, skip_warmup, return_doc_ids, data_cache_path)
            for lst, d in ((tr, a), (va, b), (te, c)):
                if d:
                    lst.append(d)
        mk = lambda ds, n: BlendableDataset(ds, weights, n) if ds else None
        return mk(tr, tot[0]), mk(va, tot[1]), mk(te, tot[2])
    out = []
    for nm, pref, n in (("train", train_data_prefix, nums[0]), ("valid", valid_data_prefix, nums[1]), ("test", test_data_prefix, nums[2])):
        if pref is None:
            out.append(None)
            continue
        if len(pref) == 1:
            indexed = make_dataset(pref[0], data_impl, skip_warmup)
            docs = np.arange(0, indexed.sizes.shape[0], dtype=np.int32)
            out.append(GPTDataset(nm, pref[0], docs, indexed, splits_string, n, seq_length, seed, data_cache_path=data_cache_path))
        else:
            prefixes, weights, per = get_datasets_weights_and_num_samples(pref, n)
            ds = []
            for i, pf in enumerate(prefixes):
                indexed = make_dataset(pf, data_impl, skip_warmup)
                docs = np.arange(0, indexed.sizes.shape[0], dtype=np.int32)
                ds.append(GPTDataset(nm, pf, docs, indexed, splits_string, per[i], seq_length, seed, data_cache_path=data_cache_path))
            out.append(BlendableDataset(ds, weights, sum(per)))
    return tuple(out)
