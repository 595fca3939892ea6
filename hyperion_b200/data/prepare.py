"""Dataset preparation from the real corpora: the reference's ``02_development/dataset_preparation.ipynb``.

* ``make_dirs``          -- ``data/``, ``data/raw``, ``data/processed`` under the project directory (c0 L4-11)
* ``prepare_wikitext2``  -- WikiText-2-raw-v1, blank / whitespace-only lines dropped (36718 -> 23767 train rows, 4358 -> 2891 test,
  3760 -> 2461 validation), GPT-2 tokenizer with ``pad_token = eos_token``, truncated and padded to 128, columns ``input_ids`` and
  ``attention_mask``, saved with ``save_to_disk`` to ``data/processed/wikitext2_tokenized`` (c3 L13-49)
* ``prepare_cifar10``    -- CIFAR-10, ``ToTensor`` + ``Normalize(0.5, 0.5)``, images of the wrong shape or all zero dropped, saved as a
  python ``list[(FloatTensor[3,32,32], int)]`` to ``data/processed/cifar10_{train,test}.pt`` (c5 L16-51)
* ``verify``             -- reloads what was written and prints the notebook's checks (c3 L53-61, c5 L55-57)

Where the text / images come from, in this order: the ``source`` argument (any mapping split -> rows; what the tests use), files already
under ``data/raw`` (``wikitext-2-raw/wiki.<split>.raw`` or an HF cache; ``cifar-10-batches-py``), and only then a download
(``datasets.load_dataset`` / ``torchvision.datasets.CIFAR10(download=True)``) -- a machine without network gets a clear error that names
``python -m hyperion_b200 data synth``, which writes the same on-disk formats from a seeded generator.

The whole split is tokenised / normalised as arrays (one tokenizer call per 1000 lines, one tensor expression per image split) instead of
per example, and the CIFAR filter is one reduction over the split.
"""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

from . import synth

WIKITEXT_FILES = {"train": "wiki.train.raw", "validation": "wiki.valid.raw", "test": "wiki.test.raw"}


class DataUnavailable(RuntimeError):
    pass


def make_dirs(base_dir=".") -> dict:
    base = Path(base_dir)
    d = {"base": base, "data": base / "data", "raw": base / "data" / "raw", "processed": base / "data" / "processed"}
    for p in d.values():
        p.mkdir(parents=True, exist_ok=True)
    return d


# ------------------------------------------------------------------------------------------------------------- WikiText-2
def _wikitext_lines(raw_dir: Path, source, download: bool) -> dict:
    """{split: list[str]} of the raw lines, blank ones still included."""
    if source is not None:
        out = {}
        for split in source:
            rows = source[split]
            out[split] = list(rows["text"]) if hasattr(rows, "column_names") or isinstance(rows, dict) else list(rows)
        return out
    for sub in ("wikitext-2-raw", "wikitext-2-raw-v1", "."):
        if all((raw_dir / sub / f).exists() for f in WIKITEXT_FILES.values()):
            # the archive's layout: one line of text per row, like the hub's dataset rows (which keep the trailing newline)
            return {s: (raw_dir / sub / f).read_text(encoding="utf-8").splitlines(keepends=True) for s, f in WIKITEXT_FILES.items()}
    try:
        if not download:
            raise FileNotFoundError("download=False")
        from datasets import load_dataset

        ds = load_dataset("wikitext", "wikitext-2-raw-v1", cache_dir=str(raw_dir))
        return {split: list(ds[split]["text"]) for split in ds}
    except Exception as e:  # noqa: BLE001  (no network, no cache)
        raise DataUnavailable(f"WikiText-2 is neither under {raw_dir} nor downloadable ({type(e).__name__}: {e}); "
                              "`python -m hyperion_b200 data synth` writes the same format from a seeded generator") from e


def load_tokenizer(tokenizer="gpt2"):
    """A tokenizer object is used as it is; a name / path goes through ``AutoTokenizer``.  ``pad_token`` falls back to ``eos_token``."""
    if isinstance(tokenizer, (str, Path)):
        from transformers import AutoTokenizer

        try:
            tokenizer = AutoTokenizer.from_pretrained(str(tokenizer), use_fast=True)
        except Exception as e:  # noqa: BLE001
            raise DataUnavailable(f"tokenizer {tokenizer!r} is not available locally and cannot be downloaded ({type(e).__name__})") from e
    if tokenizer.pad_token is None:
        tokenizer.pad_token = tokenizer.eos_token
    return tokenizer


def tokenize_lines(lines: list, tokenizer, max_length: int = 128, chunk: int = 1000):
    """(ids int64 [n, max_length], mask int64 [n, max_length]) -- truncation + padding to ``max_length`` like the notebook's map."""
    ids = np.empty((len(lines), max_length), dtype=np.int64)
    mask = np.empty((len(lines), max_length), dtype=np.int64)
    for s in range(0, len(lines), chunk):
        enc = tokenizer(lines[s:s + chunk], truncation=True, padding="max_length", max_length=max_length, return_attention_mask=True)
        ids[s:s + chunk] = np.asarray(enc["input_ids"], dtype=np.int64)
        mask[s:s + chunk] = np.asarray(enc["attention_mask"], dtype=np.int64)
    return ids, mask


def prepare_wikitext2(base_dir=".", *, source=None, tokenizer="gpt2", max_length: int = 128, download: bool = True, verbose: bool = True):
    """Returns the saved ``DatasetDict``; prints the per-split counts the notebook prints."""
    from datasets import Dataset, DatasetDict

    d = make_dirs(base_dir)
    raw = _wikitext_lines(d["raw"], source, download)
    tok = load_tokenizer(tokenizer)
    out = {}
    for split, lines in raw.items():
        kept = [t for t in lines if t.strip()]
        if verbose:
            print(f"{split} split: {len(lines)} lines, {len(kept)} non-empty samples")
        ids, mask = tokenize_lines(kept, tok, max_length)
        out[split] = Dataset.from_dict({"input_ids": ids.tolist(), "attention_mask": mask.tolist()})
    dd = DatasetDict(out)
    dst = d["processed"] / "wikitext2_tokenized"
    dd.save_to_disk(str(dst))
    if verbose:
        print(f"Tokenized WikiText-2 (no empty lines) saved to: {dst}")
    return dd


# ---------------------------------------------------------------------------------------------------------------- CIFAR-10
def _as_uint8_nhwc(images) -> np.ndarray:
    if isinstance(images, np.ndarray) and images.ndim == 4:
        return images
    return np.stack([np.asarray(im) for im in images]) if len(images) else np.zeros((0, 32, 32, 3), np.uint8)


def _cifar_arrays(raw_dir: Path, source, download: bool) -> dict:
    """{split: (uint8 [n, 32, 32, 3] or a list of per-image arrays, labels)}."""
    if source is not None:
        return {s: (v[0], list(v[1])) for s, v in source.items()}
    try:
        import torchvision

        out = {}
        have = (raw_dir / "cifar-10-batches-py").exists()
        for split, train in (("train", True), ("test", False)):
            ds = torchvision.datasets.CIFAR10(root=str(raw_dir), train=train, download=download and not have)
            out[split] = (ds.data, list(ds.targets))
        return out
    except Exception as e:  # noqa: BLE001
        raise DataUnavailable(f"CIFAR-10 is neither under {raw_dir} nor downloadable ({type(e).__name__}: {e}); "
                              "`python -m hyperion_b200 data synth` writes the same format from a seeded generator") from e


def normalize_and_filter(images, labels):
    """``ToTensor`` + ``Normalize((.5,.5,.5), (.5,.5,.5))`` on the whole split, then the notebook's validity filter: shape (3, 32, 32)
    and not all zero *after* normalisation.  Returns ``list[(FloatTensor[3,32,32], int)]``."""
    ragged = not (isinstance(images, np.ndarray) and images.ndim == 4)
    if ragged:                                     # per-image shapes may differ: drop the wrong ones first
        keep = [i for i, im in enumerate(images) if np.asarray(im).shape == (32, 32, 3)]
        images, labels = [np.asarray(images[i]) for i in keep], [labels[i] for i in keep]
    arr = _as_uint8_nhwc(images)
    if arr.shape[0] == 0 or arr.shape[1:] != (32, 32, 3):
        return []
    x = torch.from_numpy(np.ascontiguousarray(arr)).permute(0, 3, 1, 2).float().div_(255.0).sub_(0.5).div_(0.5)
    ok = (x != 0).flatten(1).any(1)
    return [(x[i].clone(), int(labels[i])) for i in torch.nonzero(ok).flatten().tolist()]


def save_ppm(img: torch.Tensor, path) -> Path:
    """The notebook shows one un-normalised sample with matplotlib; here it is written as a binary PPM (no image library needed)."""
    a = ((img.float() / 2 + 0.5).clamp(0, 1) * 255).round().to(torch.uint8).permute(1, 2, 0).contiguous().numpy()
    path = Path(path)
    path.write_bytes(b"P6\n%d %d\n255\n" % (a.shape[1], a.shape[0]) + a.tobytes())
    return path


def prepare_cifar10(base_dir=".", *, source=None, download: bool = True, verbose: bool = True) -> dict:
    d = make_dirs(base_dir)
    out = {}
    for split, (images, labels) in _cifar_arrays(d["raw"], source, download).items():
        rows = normalize_and_filter(images, labels)
        if verbose:
            print(f"CIFAR-10 {split}: {len(labels)} images, {len(rows)} after the shape / all-zero filter")
        torch.save(rows, d["processed"] / f"cifar10_{split}.pt")
        out[split] = rows
    if out.get("train"):
        save_ppm(out["train"][0][0], d["processed"] / "cifar10_sample.ppm")
    if verbose:
        print("Saved filtered CIFAR-10 train and test datasets to", d["processed"])
    return out


# ------------------------------------------------------------------------------------------------------------------ verify
def verify(base_dir=".", verbose: bool = True) -> dict:
    """Reload what ``prepare_*`` (or ``data synth``) wrote; returns row counts, the first sample with a non-empty mask, the image shape."""
    from datasets import load_from_disk

    from .datasets import WikiText2TorchDataset

    proc = Path(base_dir) / "data" / "processed"
    rep: dict = {}
    wt = proc / "wikitext2_tokenized"
    if wt.exists():
        dd = load_from_disk(str(wt))
        rep["wikitext2_rows"] = {s: len(dd[s]) for s in dd}
        ds = WikiText2TorchDataset(dd, "train")
        nz = torch.nonzero(ds.mask.sum(1) > 0).flatten()
        if len(nz):
            i = int(nz[0])
            rep["wikitext2_sample"] = {"index": i, "input_ids": ds.ids[i, :20].tolist(), "attention_mask": ds.mask[i, :20].tolist()}
        rep["wikitext2_seq_len"] = int(ds.ids.shape[1])
        rep["wikitext2_pad_fraction"] = float(1.0 - ds.mask.float().mean())
        rep["matches_reference_rows"] = rep["wikitext2_rows"] == synth.WIKITEXT2_ROWS
    for split in ("train", "test"):
        f = proc / f"cifar10_{split}.pt"
        if f.exists():
            rows = torch.load(f, weights_only=False)
            img, lbl = rows[0]
            rep[f"cifar10_{split}"] = {"rows": len(rows), "shape": tuple(img.shape), "dtype": str(img.dtype), "label": int(lbl),
                                       "min": float(img.min()), "max": float(img.max())}
    if verbose:
        for k, v in rep.items():
            print(f"{k}: {v}")
    return rep
