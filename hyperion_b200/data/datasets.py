"""Runtime dataset adapters with the reference's names and item contract
(02_development/distributed_utils.py:43-67), plus tensor-backed variants that avoid the
reference's per-item ``torch.tensor(list)`` conversion (a CPU hot spot)."""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


class WikiText2TorchDataset(Dataset):
    """``__getitem__`` -> ``(LongTensor[seq] input_ids, LongTensor[seq] attention_mask)``.

    Accepts an HF ``DatasetDict`` (as the reference does) or a dict of ``{split: (ids, mask)}``
    arrays.  The split is converted to two tensors once, not per item.
    """

    def __init__(self, hf_dataset, split: str = "train"):
        data = hf_dataset[split]
        if isinstance(data, tuple):
            ids, mask = data
            self.ids, self.mask = torch.as_tensor(ids, dtype=torch.long), torch.as_tensor(mask, dtype=torch.long)
        else:
            # straight from the arrow table (fixed-length lists -> one reshape), no datasets formatter: both its torch and
            # numpy formatters import torchvision.io.VideoReader whenever torchvision is loaded, which this image lacks
            import numpy as np

            def column(name):
                if getattr(data, "_indices", None) is None:
                    col = data.data.column(name).combine_chunks()
                    return col.values.to_numpy(zero_copy_only=False).reshape(len(col), -1).astype(np.int64)
                return np.asarray(data.with_format(None)[name], dtype=np.int64)

            self.ids = torch.from_numpy(column("input_ids"))
            self.mask = torch.from_numpy(column("attention_mask"))

    def __len__(self):
        return self.ids.shape[0]

    def __getitem__(self, idx):
        return self.ids[idx], self.mask[idx]


class CIFAR10TorchDataset(Dataset):
    """Pass-through over ``list[(img, label)]`` exactly like the reference."""

    def __init__(self, data):
        self.data = data

    def __len__(self):
        return len(self.data)

    def __getitem__(self, idx):
        img, label = self.data[idx]
        return img, label


class DeviceResidentLoader:
    """Whole dataset on the device; yields shuffled rank-local batches with no host work per step.

    Replaces ``DataLoader(DistributedSampler(...), num_workers=2)`` (distributed_utils.py:149-152)
    for synthetic runs: same sharding rule (rank ``r`` takes indices ``r, r+W, ...`` of the epoch
    permutation, padded to equal length), same ``set_epoch`` contract.
    """

    def __init__(self, tensors, batch_size: int, rank: int = 0, world_size: int = 1, shuffle: bool = True,
                 seed: int = 0, drop_last: bool = False):
        self.tensors = tensors
        self.n = tensors[0].shape[0]
        self.bs, self.rank, self.world, self.shuffle, self.seed, self.drop_last = batch_size, rank, world_size, shuffle, seed, drop_last
        self.epoch = 0
        self.per_rank = (self.n + world_size - 1) // world_size

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def __len__(self):
        return self.per_rank // self.bs if self.drop_last else (self.per_rank + self.bs - 1) // self.bs

    def __iter__(self):
        g = torch.Generator().manual_seed(self.seed + self.epoch)
        perm = torch.randperm(self.n, generator=g) if self.shuffle else torch.arange(self.n)
        total = self.per_rank * self.world
        if total > self.n:
            perm = torch.cat([perm, perm[: total - self.n]])
        idx = perm[self.rank:total:self.world].to(self.tensors[0].device)
        for i in range(len(self)):
            sel = idx[i * self.bs:(i + 1) * self.bs]
            yield tuple(t[sel] for t in self.tensors)
