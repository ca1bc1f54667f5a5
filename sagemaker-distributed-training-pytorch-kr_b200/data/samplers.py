"""Batch samplers for pre-training (megatron/data/data_samplers.py): a DP-sharded,
resumable-from-`consumed_samples` sequential sampler and its cyclic/shuffling variant, plus the loader
factory used by the training loop."""
from __future__ import annotations

import torch

from ..parallel import state as mpu


class MegatronPretrainingSampler:
    def __init__(self, total_samples, consumed_samples, micro_batch_size, data_parallel_rank, data_parallel_size, drop_last=True):
        assert total_samples > 0, f"no sample to consume: {total_samples}"
        assert consumed_samples < total_samples, f"no samples left to consume: {consumed_samples}, {total_samples}"
        assert micro_batch_size > 0 and data_parallel_size > 0 and data_parallel_rank < data_parallel_size
        self.total_samples, self.consumed_samples = total_samples, consumed_samples
        self.mbs, self.dp_rank, self.dp_size, self.drop_last = micro_batch_size, data_parallel_rank, data_parallel_size, drop_last
        self.step = micro_batch_size * data_parallel_size

    def __len__(self):
        return self.total_samples

    def __iter__(self):
        batch = []
        lo, hi = self.dp_rank * self.mbs, (self.dp_rank + 1) * self.mbs
        for idx in range(self.consumed_samples, self.total_samples):
            batch.append(idx)
            if len(batch) == self.step:
                yield batch[lo:hi]
                batch = []
        if batch and not self.drop_last:
            yield batch[lo:hi]


class MegatronPretrainingRandomSampler:
    def __init__(self, total_samples, consumed_samples, micro_batch_size, data_parallel_rank, data_parallel_size, seed=1234):
        self.total_samples, self.consumed_samples = total_samples, consumed_samples
        self.mbs, self.dp_rank, self.dp_size, self.seed = micro_batch_size, data_parallel_rank, data_parallel_size, seed
        self.step = micro_batch_size * data_parallel_size
        self.last_batch_size = total_samples % self.step

    def __len__(self):
        return self.total_samples

    def __iter__(self):
        active = self.total_samples - self.last_batch_size
        epoch = self.consumed_samples // active
        cur = self.consumed_samples % active
        assert cur % self.step == 0
        bucket = (self.total_samples // self.step) * self.mbs
        offset = cur // self.dp_size
        start = self.dp_rank * bucket
        g = torch.Generator().manual_seed(self.seed + epoch)
        order = torch.randperm(bucket, generator=g).tolist()
        idxs = [start + x for x in order[offset:]]
        batch = []
        for i in idxs:
            batch.append(i)
            if len(batch) == self.mbs:
                self.consumed_samples += self.step
                yield batch
                batch = []


def build_pretraining_data_loader(dataset, consumed_samples, args):
    if dataset is None:
        return None
    dp_rank, dp_size = mpu.get_data_parallel_rank(), mpu.get_data_parallel_world_size()
    if args.dataloader_type == "single":
        sampler = MegatronPretrainingSampler(len(dataset), consumed_samples, args.micro_batch_size, dp_rank, dp_size)
    elif args.dataloader_type == "cyclic":
        sampler = MegatronPretrainingRandomSampler(len(dataset), consumed_samples, args.micro_batch_size, dp_rank, dp_size, args.seed)
    else:
        raise Exception(f"{args.dataloader_type} dataloader type is not supported.")
    return torch.utils.data.DataLoader(dataset, batch_sampler=sampler, num_workers=args.num_workers if not getattr(dataset, "in_memory", False) else 0,
                                       pin_memory=torch.cuda.is_available())
