"""DistributedSampler with index math identical to ``torch.utils.data.DistributedSampler``
(used as ``DistributedSampler(dataset)`` at ref:pytorch/resnet/main.py:94 and ref:pytorch/unet/train.py:96;
SURVEY.md N13): ``randperm(len, seed + epoch)``, pad to a multiple of the world size by wrapping around (or drop the
tail), take ``indices[rank::world]``.  The reference never calls ``set_epoch`` so every epoch sees the same
permutation; ``set_epoch`` is offered all the same."""
from __future__ import annotations

import math
from typing import Iterator, Optional

import torch


class DistributedSampler:
    def __init__(self, dataset, num_replicas: Optional[int] = None, rank: Optional[int] = None, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        if num_replicas is None or rank is None:
            from . import dist as D

            if D.is_initialized():
                num_replicas = D.get_world_size() if num_replicas is None else num_replicas
                rank = D.get_rank() if rank is None else rank
            else:
                import torch.distributed as tdist

                if not tdist.is_available() or not tdist.is_initialized():
                    raise RuntimeError("Requires distributed package to be available")
                num_replicas = tdist.get_world_size() if num_replicas is None else num_replicas
                rank = tdist.get_rank() if rank is None else rank
        if rank >= num_replicas or rank < 0:
            raise ValueError(f"Invalid rank {rank}, rank should be in the interval [0, {num_replicas - 1}]")
        self.dataset = dataset
        self.num_replicas = num_replicas
        self.rank = rank
        self.epoch = 0
        self.drop_last = drop_last
        n = len(self.dataset)
        if self.drop_last and n % self.num_replicas != 0:
            self.num_samples = math.ceil((n - self.num_replicas) / self.num_replicas)
        else:
            self.num_samples = math.ceil(n / self.num_replicas)
        self.total_size = self.num_samples * self.num_replicas
        self.shuffle = shuffle
        self.seed = seed

    def __iter__(self) -> Iterator[int]:
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            indices = torch.randperm(n, generator=g).tolist()
        else:
            indices = list(range(n))
        if not self.drop_last:
            padding = self.total_size - len(indices)
            if padding <= len(indices):
                indices += indices[:padding]
            else:
                indices += (indices * math.ceil(padding / len(indices)))[:padding]
        else:
            indices = indices[: self.total_size]
        assert len(indices) == self.total_size
        indices = indices[self.rank: self.total_size: self.num_replicas]
        assert len(indices) == self.num_samples
        return iter(indices)

    def __len__(self) -> int:
        return self.num_samples

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch
