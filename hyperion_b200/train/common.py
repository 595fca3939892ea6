"""Helpers shared by the trainers: data loading (real files if present, synthetic otherwise), device-side
loss accumulation, per-epoch device timing with max over ranks."""
from __future__ import annotations

import os
import time
from pathlib import Path

import torch
import torch.distributed as dist

from ..data import DeviceResidentLoader, synth
from ..parallel.runtime import max_over_ranks, world


def load_wikitext_tensors(base_dir, device, synthetic: bool, rows: int | None = None, seq_len: int = 128,
                          vocab: int = synth.GPT2_VOCAB, pad_id: int = synth.GPT2_PAD, seed: int = 0):
    """(input_ids, attention_mask) on ``device``: from ``data/processed/wikitext2_tokenized`` if it exists and
    ``synthetic`` is False, else WikiText-2-shaped synthetic tokens (23767 x 128 by default)."""
    path = Path(base_dir) / "data/processed/wikitext2_tokenized"
    if not synthetic and path.exists():
        from datasets import load_from_disk

        from ..data import WikiText2TorchDataset

        ds = WikiText2TorchDataset(load_from_disk(str(path)), "train")
        return ds.ids.to(device), ds.mask.to(device)
    n = rows or synth.WIKITEXT2_ROWS["train"]
    ids, mask = synth.wikitext2_like_arrays(n, seq_len, vocab, pad_id, seed)
    return torch.from_numpy(ids).to(device), torch.from_numpy(mask).to(device)


def load_cifar_tensors(base_dir, device, synthetic: bool, rows: int | None = None, seed: int = 0):
    path = Path(base_dir) / "data/processed/cifar10_train.pt"
    if not synthetic and path.exists():
        data = torch.load(path, weights_only=False)
        x = torch.stack([d[0] for d in data]).to(device)
        y = torch.tensor([d[1] for d in data], device=device)
        return x, y
    n = rows or synth.CIFAR10_ROWS["train"]
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, 10, (n,), generator=g)
    protos = torch.randn(10, 3, 32, 32, generator=torch.Generator().manual_seed(999)) * 0.5
    x = (protos[labels] + 0.5 * torch.randn(n, 3, 32, 32, generator=g)).clamp_(-1, 1)
    return x.to(device), labels.to(device)


class EpochTimer:
    """Wall seconds (the reference's ``duration`` column) plus device milliseconds via CUDA events."""

    def __init__(self, device):
        self.cuda = torch.device(device).type == "cuda"
        self.device = device

    def start(self):
        if self.cuda:
            torch.cuda.synchronize(self.device)
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        self.t0 = time.time()

    def stop(self) -> tuple[float, float]:
        if self.cuda:
            self.e1.record()
            torch.cuda.synchronize(self.device)
            dev_ms = self.e0.elapsed_time(self.e1)
        else:
            dev_ms = (time.time() - self.t0) * 1e3
        return time.time() - self.t0, max_over_ranks(dev_ms, self.device if self.cuda else None)


def mean_over_ranks(t: torch.Tensor) -> float:
    """Scalar all-reduce of a device tensor -> host float (one sync per epoch, K-C4)."""
    t = t.detach().clone().float().reshape(1)
    rank, W = world()
    if W > 1:
        dist.all_reduce(t)
        t /= W
    return float(t.item())
