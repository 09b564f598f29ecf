"""Batch samplers and the DataLoader factory (megatron/data/data_samplers.py:14-186
surface). Each data-parallel rank takes its ``micro_batch_size`` slice of every
``micro_batch_size * dp`` chunk; sampling resumes from ``consumed_samples``."""
import random

import torch

from .. import global_vars
from ..parallel import state as mpu


def build_pretraining_data_loader(dataset, consumed_samples, args=None):
    if dataset is None:
        return None
    args = args or global_vars.get_args()
    if args.dataloader_type == "single":
        sampler = MegatronPretrainingSampler(
            total_samples=len(dataset), consumed_samples=consumed_samples,
            micro_batch_size=args.micro_batch_size, data_parallel_rank=mpu.get_data_parallel_rank(),
            data_parallel_size=mpu.get_data_parallel_world_size())
    elif args.dataloader_type == "cyclic":
        sampler = MegatronPretrainingRandomSampler(
            dataset, total_samples=len(dataset), consumed_samples=consumed_samples,
            micro_batch_size=args.micro_batch_size, data_parallel_rank=mpu.get_data_parallel_rank(),
            data_parallel_size=mpu.get_data_parallel_world_size(), data_sharding=args.data_sharding)
    else:
        raise Exception(f"{args.dataloader_type} dataloader type is not supported.")
    return torch.utils.data.DataLoader(dataset, batch_sampler=sampler, num_workers=args.num_workers,
                                       pin_memory=True)


class MegatronPretrainingSampler:
    """Sequential: sample ids consumed_samples, consumed_samples + 1, ... are cut into chunks of
    micro_batch_size * dp; rank r yields rows [r * mbs, (r + 1) * mbs) of every chunk. An incomplete last
    chunk is dropped unless ``drop_last`` is False (then every rank yields its, possibly short or empty, share)."""

    def __init__(self, total_samples, consumed_samples, micro_batch_size, data_parallel_rank,
                 data_parallel_size, drop_last=True):
        if total_samples <= 0:
            raise AssertionError(f"no sample to consume: {total_samples}")
        if consumed_samples >= total_samples:
            raise AssertionError(f"no samples left to consume: {consumed_samples}, {total_samples}")
        if micro_batch_size <= 0 or not 0 <= data_parallel_rank < data_parallel_size:
            raise AssertionError((micro_batch_size, data_parallel_rank, data_parallel_size))
        self.total_samples, self.consumed_samples = total_samples, consumed_samples
        self.micro_batch_size, self.data_parallel_rank = micro_batch_size, data_parallel_rank
        self.chunk = micro_batch_size * data_parallel_size
        self.drop_last = drop_last

    def __len__(self):
        return self.total_samples

    def __iter__(self):
        mine = self.data_parallel_rank * self.micro_batch_size
        for first in range(self.consumed_samples, self.total_samples, self.chunk):
            last = min(first + self.chunk, self.total_samples)
            if last - first < self.chunk and self.drop_last:
                return
            yield list(range(min(first + mine, last), min(first + mine + self.micro_batch_size, last)))


class RandomSeedDataset(torch.utils.data.Dataset):
    def __init__(self, dataset, seed=None):
        self.base_seed = seed if seed is not None else global_vars.get_args().seed
        self.curr_seed = self.base_seed
        self.dataset = dataset

    def __len__(self):
        return len(self.dataset)

    def set_epoch(self, epoch):
        self.curr_seed = self.base_seed + epoch

    def __getitem__(self, idx):
        import numpy as np
        seed = idx + self.curr_seed
        torch.manual_seed(seed)
        random.seed(seed)
        np.random.seed(seed)
        return self.dataset[idx]


class MegatronPretrainingRandomSampler:
    def __init__(self, dataset, total_samples, consumed_samples, micro_batch_size, data_parallel_rank,
                 data_parallel_size, data_sharding):
        self.dataset = dataset
        self.total_samples, self.consumed_samples = total_samples, consumed_samples
        self.micro_batch_size = micro_batch_size
        self.data_parallel_rank, self.data_parallel_size = data_parallel_rank, data_parallel_size
        self.data_sharding = data_sharding
        self.chunk = micro_batch_size * data_parallel_size
        self.last_batch_size = self.total_samples % self.chunk
        assert self.total_samples > 0 and self.micro_batch_size > 0 and data_parallel_size > 0
        assert self.data_parallel_rank < data_parallel_size

    def __len__(self):
        return self.total_samples

    def __iter__(self):
        active = self.total_samples - self.last_batch_size
        self.epoch = self.consumed_samples // active
        current = self.consumed_samples % active
        assert current % self.chunk == 0
        if isinstance(self.dataset, RandomSeedDataset):
            self.dataset.set_epoch(self.epoch)
        g = torch.Generator()
        g.manual_seed(self.epoch)
        if self.data_sharding:
            bucket = (self.total_samples // self.chunk) * self.micro_batch_size
            offset = current // self.data_parallel_size
            start = self.data_parallel_rank * bucket
            order = torch.randperm(bucket, generator=g).tolist()
            idx_range = [start + x for x in order[offset:]]
        else:
            full = (self.total_samples // self.micro_batch_size) * self.micro_batch_size
            order = torch.randperm(full, generator=g).tolist()
            idx_range = order[current:][self.data_parallel_rank::self.data_parallel_size]
        batch = []
        for idx in idx_range:
            batch.append(idx)
            if len(batch) == self.micro_batch_size:
                self.consumed_samples += self.chunk
                yield batch
                batch = []
