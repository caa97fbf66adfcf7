"""Seeded random sampling, rank sharding, batching and resume-skipping.

HF Trainer semantics: a fresh random permutation per epoch from `seed` (42), each rank takes a
disjoint strided shard (padded by wrap-around so all ranks see the same number of batches), and on
resume the already-consumed batches of the current epoch are skipped.
"""
from __future__ import annotations

import torch


class DistributedBatchSampler:
    def __init__(self, dataset_len: int, batch_size: int, rank: int = 0, world_size: int = 1,
                 seed: int = 42, shuffle: bool = True, drop_last: bool = False):
        self.n, self.bs, self.rank, self.world = dataset_len, batch_size, rank, world_size
        self.seed, self.shuffle, self.drop_last = seed, shuffle, drop_last
        per_rank = -(-self.n // self.world)
        self.per_rank = per_rank
        self.batches_per_epoch = per_rank // self.bs if drop_last else -(-per_rank // self.bs)

    def __len__(self) -> int:
        return self.batches_per_epoch

    def epoch_indices(self, epoch: int) -> list[int]:
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed + epoch)
            perm = torch.randperm(self.n, generator=g).tolist()
        else:
            perm = list(range(self.n))
        total = self.per_rank * self.world
        perm = perm + perm[: total - len(perm)]
        return perm[self.rank:total:self.world]

    def iter_epoch(self, epoch: int, skip_batches: int = 0):
        idx = self.epoch_indices(epoch)
        for b in range(skip_batches, self.batches_per_epoch):
            chunk = idx[b * self.bs:(b + 1) * self.bs]
            if chunk:
                yield chunk


class DataLoader:
    """Minimal loader: sampler -> dataset -> collator (runs on the host; batches are tiny)."""

    def __init__(self, dataset, sampler: DistributedBatchSampler, collator):
        self.ds, self.sampler, self.collate = dataset, sampler, collator

    def __len__(self) -> int:
        return len(self.sampler)

    def iter_epoch(self, epoch: int, skip_batches: int = 0):
        for chunk in self.sampler.iter_epoch(epoch, skip_batches):
            yield self.collate([self.ds[i] for i in chunk])
