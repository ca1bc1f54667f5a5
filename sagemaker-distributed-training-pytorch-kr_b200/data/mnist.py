"""MNIST input pipeline: raw IDX reader (the four ubyte files torchvision's MNIST stores under
``<root>/MNIST/raw``), a synthetic stand-in of the same shape for boxes without the dataset, the
distributed sampler and a minimal batch loader (reference: pytorch_mnist_ddp.py:216-255)."""
from __future__ import annotations

import gzip
import math
import os
import struct
from typing import Iterator, Optional, Tuple

import numpy as np
import torch

MEAN, STD = 0.1307, 0.3081


def _read_idx(path: str) -> np.ndarray:
    op = gzip.open if path.endswith(".gz") else open
    with op(path, "rb") as f:
        magic = struct.unpack(">I", f.read(4))[0]
        nd = magic & 0xFF
        dims = struct.unpack(">" + "I" * nd, f.read(4 * nd))
        return np.frombuffer(f.read(), dtype=np.uint8).reshape(dims)


def write_idx(path: str, arr: np.ndarray) -> None:
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "wb") as f:
        f.write(struct.pack(">I", 0x0800 | arr.ndim))
        f.write(struct.pack(">" + "I" * arr.ndim, *arr.shape))
        f.write(arr.astype(np.uint8).tobytes())


def find_raw(root: str, train: bool) -> Optional[Tuple[str, str]]:
    stem = "train" if train else "t10k"
    for base in (os.path.join(root, "MNIST", "raw"), root):
        for ext in ("", ".gz"):
            a = os.path.join(base, f"{stem}-images-idx3-ubyte{ext}")
            b = os.path.join(base, f"{stem}-labels-idx1-ubyte{ext}")
            if os.path.exists(a) and os.path.exists(b):
                return a, b
    return None


class MNISTTensors(torch.utils.data.Dataset):
    """Whole split resident as normalised float tensors (60k x 1 x 28 x 28 = 188 MB fp32)."""

    def __init__(self, root: str, train: bool = True, synthetic: Optional[bool] = None, synthetic_size: Optional[int] = None, seed: int = 0):
        files = find_raw(root, train) if root else None
        if synthetic is None:
            synthetic = files is None
        if synthetic:
            n = synthetic_size or (60000 if train else 10000)
            g = torch.Generator().manual_seed(seed + (0 if train else 1))
            self.labels = torch.randint(0, 10, (n,), generator=g)
            # class-dependent blobs so that a model can actually fit the synthetic set
            base = torch.rand(10, 1, 28, 28, generator=torch.Generator().manual_seed(1234))
            noise = torch.rand(n, 1, 28, 28, generator=g) * 0.5
            self.images = ((base[self.labels] + noise).clamp(0, 1) - MEAN) / STD
            self.synthetic = True
        else:
            img, lab = _read_idx(files[0]), _read_idx(files[1])
            self.images = ((torch.from_numpy(img.copy()).float() / 255.0) - MEAN).div_(STD).unsqueeze(1)
            self.labels = torch.from_numpy(lab.copy()).long()
            self.synthetic = False

    def __len__(self):
        return self.labels.numel()

    def __getitem__(self, i):
        return self.images[i], self.labels[i]


class DistributedSampler:
    """Same contract as torch.utils.data.DistributedSampler: pad to ceil(N/world)*world,
    epoch-seeded permutation, strided shard."""

    def __init__(self, n: int, num_replicas: int, rank: int, shuffle: bool = True, seed: int = 0, drop_last: bool = False):
        self.n, self.world, self.rank, self.shuffle, self.seed, self.epoch = n, num_replicas, rank, shuffle, seed, 0
        self.num_samples = n // num_replicas if drop_last else math.ceil(n / num_replicas)
        self.total = self.num_samples * num_replicas
        self.drop_last = drop_last

    def set_epoch(self, epoch: int):
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if self.drop_last:
            idx = idx[: self.total]
        elif self.total > self.n:
            pad = self.total - self.n
            idx = torch.cat([idx, idx.repeat(math.ceil(pad / self.n))[:pad]])
        return idx[self.rank: self.total: self.world]

    def __iter__(self):
        return iter(self.indices().tolist())

    def __len__(self):
        return self.num_samples


class TensorBatchLoader:
    """Batches straight out of resident tensors; pinned staging + non_blocking H2D when a device is given."""

    def __init__(self, ds: MNISTTensors, batch_size: int, sampler: Optional[DistributedSampler] = None, shuffle: bool = False,
                 device: Optional[torch.device] = None, pin: bool = False):
        self.dataset, self.bs, self.sampler, self.shuffle, self.device = ds, batch_size, sampler, shuffle, device
        self.pin = pin and torch.cuda.is_available()

    def __len__(self):
        n = len(self.sampler) if self.sampler is not None else len(self.dataset)
        return math.ceil(n / self.bs)

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        if self.sampler is not None:
            idx = self.sampler.indices()
        elif self.shuffle:
            idx = torch.randperm(len(self.dataset))
        else:
            idx = torch.arange(len(self.dataset))
        for s in range(0, idx.numel(), self.bs):
            sel = idx[s: s + self.bs]
            x, y = self.dataset.images[sel], self.dataset.labels[sel]
            if self.pin:
                x, y = x.pin_memory(), y.pin_memory()
            if self.device is not None:
                x, y = x.to(self.device, non_blocking=True), y.to(self.device, non_blocking=True)
            yield x, y
