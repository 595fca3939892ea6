"""Helpers shared by the trainers: data loading (real files if present, synthetic otherwise), device-side
loss accumulation, per-epoch device timing with max over ranks."""
from __future__ import annotations

import os
import time
from pathlib import Path

import torch
import torch.distributed as dist

from ..data import synth
from ..parallel.runtime import max_over_ranks, world


def _warn_missing(path, what: str) -> None:
    """``--real-data`` was asked for and the prepared files are not there: say so (once per process and dataset) before falling back."""
    if str(path) not in _warned and world()[0] == 0:
        print(f"WARNING: {path} not found -- training on synthetic {what}-shaped data instead "
              f"(write the real files with `python -m hyperion_b200 data prepare --base_dir ...`)", flush=True)
    _warned.add(str(path))


_warned: set = set()


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
    if not synthetic:
        _warn_missing(path, "WikiText-2")
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
    if not synthetic:
        _warn_missing(path, "CIFAR-10")
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


def graph_default(device) -> bool:
    """CUDA-graph the step unless ``HYPERION_GRAPH=0`` (needs a CUDA device and the native library)."""
    from ..ops import _lib

    return torch.device(device).type == "cuda" and _lib.available() and os.environ.get("HYPERION_GRAPH", "1") != "0"


class StepRunner:
    """What every trainer calls per batch: ``loss = runner(batch_tuple)``.

    With ``graph=True`` the whole step (forward, backward, communication kernels, optimizer) is captured in a CUDA graph
    (``GraphedStep``) after ``eager_steps`` ordinary steps and replayed from then on -- the headline configuration of
    ``bench.py``, reachable from ``run_distributed.py --graph`` (default on CUDA).  Batches whose shapes differ from
    the captured ones (the ragged last batch of an epoch) run eagerly.  ``loss_fn(batch)`` returns the loss or
    ``(loss, *extras)``; the same structure comes back (detached).  Without a graph it is just
    ``eng.backward_and_step(loss_fn(batch))``.
    """

    def __init__(self, eng, loss_fn, graph: bool | None = None, eager_steps: int = 2):
        self.eng, self.loss_fn = eng, loss_fn
        dev = getattr(eng.flat, "device", "cpu")
        self.graph = graph_default(dev) if graph is None else (graph and torch.device(dev).type == "cuda")
        self.eager_steps, self.n, self.graphed, self.shapes = eager_steps, 0, None, None
        self.replays = 0

    def _eager(self, batch):
        from ..utils import nvtx

        with nvtx.range("hyperion/forward"):
            out = self.loss_fn(batch)
        loss = out[0] if isinstance(out, (tuple, list)) else out
        self.eng.backward_and_step(loss)
        return tuple(t.detach() for t in out) if isinstance(out, (tuple, list)) else loss.detach()

    def __call__(self, batch):
        self.n += 1
        if not self.graph:
            return self._eager(batch)
        shapes = tuple((tuple(t.shape), t.dtype) for t in batch)
        if self.graphed is None:
            if self.n <= self.eager_steps:
                return self._eager(batch)
            from .engine import GraphedStep

            try:
                # the eager steps above were the warm-up; the capture itself performs no optimizer step
                self.graphed = GraphedStep(self.eng, self.loss_fn, batch, warmup=0)
                self.shapes = shapes
            except Exception as e:  # noqa: BLE001
                import warnings

                warnings.warn(f"CUDA-graph capture failed ({e!r}); continuing eagerly")
                torch.cuda.synchronize()
                self.graph = False
                return self._eager(batch)
        if shapes != self.shapes:
            return self._eager(batch)
        self.replays += 1
        from ..utils import nvtx

        with nvtx.range("hyperion/graph_replay"):
            return self.graphed(batch)
