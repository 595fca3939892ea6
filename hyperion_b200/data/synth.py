"""Synthetic datasets in the reference's on-disk formats (no network on the GPU box).

* WikiText-2-shaped tokens: the reference tokenises WikiText-2-raw-v1 with the GPT-2 tokenizer,
  ``padding="max_length", max_length=128``, pad = eos = 50256, and ``save_to_disk``s a
  ``DatasetDict`` with columns ``input_ids`` / ``attention_mask`` and 23767 / 2461 / 2891 rows
  (02_development/dataset_preparation.ipynb c3 L13-49).  ``write_wikitext2_like`` writes the same
  thing from a seeded generator: Zipf-distributed ids, a realistic length distribution, right-padded.
* CIFAR-10-shaped images: ``torch.save(list[(FloatTensor[3,32,32] in [-1,1], int)])`` with 50000 /
  10000 items (dataset_preparation.ipynb c5 L16-51).

``token_batch`` / ``image_batch`` create device-resident batches directly for the benchmarks.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

GPT2_VOCAB = 50257
GPT2_PAD = 50256
WIKITEXT2_ROWS = {"train": 23767, "validation": 2461, "test": 2891}
CIFAR10_ROWS = {"train": 50000, "test": 10000}


def _zipf_tokens(rng: np.random.Generator, n: int, vocab: int) -> np.ndarray:
    # rank-frequency ~ 1/r like natural text; ids are a fixed permutation of ranks
    ranks = np.minimum(rng.zipf(1.2, size=n), vocab - 1) - 1
    perm = np.random.default_rng(1234).permutation(vocab - 1)
    return perm[ranks].astype(np.int64)


def wikitext2_like_arrays(rows: int, seq_len: int = 128, vocab: int = GPT2_VOCAB, pad_id: int = GPT2_PAD, seed: int = 0):
    """(input_ids, attention_mask) int64 arrays of shape [rows, seq_len], right-padded with ``pad_id``."""
    rng = np.random.default_rng(seed)
    # WikiText-2 lines: many short headings, a long tail truncated at seq_len (mean ~ 75 real tokens)
    lengths = np.clip(rng.gamma(shape=1.6, scale=70.0, size=rows).astype(np.int64) + 3, 3, seq_len)
    ids = np.full((rows, seq_len), pad_id, dtype=np.int64)
    mask = np.zeros((rows, seq_len), dtype=np.int64)
    toks = _zipf_tokens(rng, int(lengths.sum()), vocab)
    off = 0
    for i, L in enumerate(lengths):
        ids[i, :L] = toks[off:off + L]
        mask[i, :L] = 1
        off += L
    return ids, mask


def write_wikitext2_like(out_dir, seq_len: int = 128, seed: int = 0, rows: dict | None = None,
                         vocab: int = GPT2_VOCAB, pad_id: int = GPT2_PAD):
    """Write an HF ``DatasetDict`` (arrow) like ``data/processed/wikitext2_tokenized``."""
    from datasets import Dataset, DatasetDict

    rows = rows or WIKITEXT2_ROWS
    dd = {}
    for k, (split, n) in enumerate(rows.items()):
        ids, mask = wikitext2_like_arrays(n, seq_len, vocab, pad_id, seed + k)
        dd[split] = Dataset.from_dict({"input_ids": ids.tolist(), "attention_mask": mask.tolist()})
    ds = DatasetDict(dd)
    Path(out_dir).parent.mkdir(parents=True, exist_ok=True)
    ds.save_to_disk(str(out_dir))
    return ds


def cifar10_like_list(n: int, seed: int = 0, num_classes: int = 10):
    """``list[(FloatTensor[3,32,32] in [-1,1], int)]``: class-dependent blobs + noise, so it is learnable."""
    g = torch.Generator().manual_seed(seed)
    labels = torch.randint(0, num_classes, (n,), generator=g)
    protos = torch.randn(num_classes, 3, 32, 32, generator=torch.Generator().manual_seed(999)) * 0.5
    imgs = (protos[labels] + 0.5 * torch.randn(n, 3, 32, 32, generator=g)).clamp_(-1, 1)
    return [(imgs[i].clone(), int(labels[i])) for i in range(n)]


def write_cifar10_like(out_dir, seed: int = 0, rows: dict | None = None):
    rows = rows or CIFAR10_ROWS
    out_dir = Path(out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    for k, (split, n) in enumerate(rows.items()):
        torch.save(cifar10_like_list(n, seed + k), out_dir / f"cifar10_{split}.pt")
    return out_dir


def token_batch(batch: int, seq_len: int, vocab: int, device="cpu", seed: int = 0, pad_id: int | None = None,
                pad_fraction: float = 0.0):
    """Random token batch ``[batch, seq_len]`` (+ attention mask) created on ``device``."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    ids = torch.randint(0, vocab - 1, (batch, seq_len), generator=g)
    mask = torch.ones(batch, seq_len, dtype=torch.long)
    if pad_id is not None and pad_fraction > 0:
        lens = torch.randint(int(seq_len * (1 - pad_fraction)), seq_len + 1, (batch,), generator=g)
        ar = torch.arange(seq_len)[None]
        mask = (ar < lens[:, None]).long()
        ids = torch.where(mask.bool(), ids, torch.full_like(ids, pad_id))
    return ids.to(device), mask.to(device)


def image_batch(batch: int, size: int = 224, num_classes: int = 1000, device="cpu", seed: int = 0, dtype=torch.float32):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(batch, 3, size, size, generator=g).to(dtype)
    y = torch.randint(0, num_classes, (batch,), generator=g)
    return x.to(device), y.to(device)
