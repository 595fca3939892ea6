"""Single-GPU trainers of the notebooks: ``train_language_model`` / ``train_cifar_model`` in fp32
(02_development/core_framework.ipynb c6: AdamW 2e-4 / 1e-3, 25 epochs, bare ``state_dict`` saved to
``data/processed/{simple_transformer_lm,resnet18_cifar10}.pt``), their AMP variants (mixed_precision.ipynb c2; saved as
``{"model_state_dict", "scaler_state_dict"}``) and ``profile_amp_training`` (c6: 10 batches, time + peak memory),
plus ``train_one_batch`` with before/after/peak memory (memory_optimization.ipynb c6)."""
from __future__ import annotations

import time
from pathlib import Path

import torch

from .. import ops
from ..data import DeviceResidentLoader, synth
from ..memory import MemoryStats
from ..models import NotebookTransformerLM, resnet18
from .common import load_cifar_tensors, load_wikitext_tensors
from .engine import build_engine


def _device(device=None):
    return torch.device(device) if device is not None else torch.device("cuda" if torch.cuda.is_available() else "cpu")


def train_language_model(model=None, epochs: int = 25, base_dir: str = ".", *, precision: str = "fp32", lr: float = 2e-4,
                         batch_size: int = 32, synthetic: bool = True, rows: int | None = None, device=None,
                         steps_per_epoch: int | None = None, pad_id: int = synth.GPT2_PAD,
                         vocab_size: int = synth.GPT2_VOCAB, use_checkpoint: bool = False, save_path=None, seed: int = 0):
    dev = _device(device)
    torch.manual_seed(seed)
    model = model if model is not None else NotebookTransformerLM(vocab_size, use_checkpoint=use_checkpoint)
    eng = build_engine(model, strategy="single", precision=precision, lr=lr, device=dev)
    ids, _ = load_wikitext_tensors(base_dir, dev, synthetic, rows, vocab=vocab_size, pad_id=pad_id, seed=seed)
    loader = DeviceResidentLoader((ids,), batch_size, shuffle=True, seed=seed)
    losses = []
    for ep in range(epochs):
        loader.set_epoch(ep)
        model.train()
        total, n = torch.zeros((), device=dev), 0
        for (batch,) in loader:
            loss = eng.model.loss(batch[:, :-1].contiguous(), batch[:, 1:].contiguous(), ignore_index=pad_id)
            eng.backward_and_step(loss)
            total += loss.detach().float()
            n += 1
            if steps_per_epoch and n >= steps_per_epoch:
                break
        losses.append(float(total.item()) / max(n, 1))
        print(f"[LM] Epoch {ep + 1}/{epochs} loss: {losses[-1]:.4f}")
    if save_path:
        _save(model, save_path, eng.scaler)
    return losses


def train_cifar_model(model=None, epochs: int = 25, base_dir: str = ".", *, precision: str = "fp32", lr: float = 1e-3,
                      batch_size: int = 64, synthetic: bool = True, rows: int | None = None, device=None,
                      steps_per_epoch: int | None = None, save_path=None, seed: int = 0):
    dev = _device(device)
    torch.manual_seed(seed)
    model = model if model is not None else resnet18(num_classes=10)
    eng = build_engine(model, strategy="single", precision=precision, lr=lr, device=dev)
    x_all, y_all = load_cifar_tensors(base_dir, dev, synthetic, rows, seed)
    loader = DeviceResidentLoader((x_all.to(eng.precision.dtype), y_all), batch_size, shuffle=True, seed=seed)
    hist = []
    for ep in range(epochs):
        loader.set_epoch(ep)
        model.train()
        stats, n = torch.zeros(3, device=dev), 0
        for x, y in loader:
            out = eng.model(x)
            loss = ops.cross_entropy(out, y)
            eng.backward_and_step(loss)
            with torch.no_grad():
                stats[0] += loss.detach().float(); stats[1] += (out.argmax(1) == y).sum(); stats[2] += y.numel()
            n += 1
            if steps_per_epoch and n >= steps_per_epoch:
                break
        s = stats.tolist()
        hist.append((s[0] / max(n, 1), 100.0 * s[1] / max(s[2], 1)))
        print(f"[CIFAR] Epoch {ep + 1}/{epochs} loss: {hist[-1][0]:.4f} acc: {hist[-1][1]:.2f}%")
    if save_path:
        _save(model, save_path, eng.scaler)
    return hist


def _save(model, path, scaler=None) -> None:
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    sd = {k: v.detach().float().cpu() if v.is_floating_point() else v.cpu() for k, v in model.state_dict().items()}
    torch.save({"model_state_dict": sd, "scaler_state_dict": scaler.state_dict()} if scaler is not None else sd, path)


def profile_amp_training(model, batches, *, precision: str = "fp16", lr: float = 2e-4, pad_id: int = synth.GPT2_PAD,
                         device=None) -> dict:
    """Time ``len(batches)`` LM training steps under AMP and report the peak memory (mixed_precision.ipynb c6)."""
    dev = _device(device)
    eng = build_engine(model, strategy="single", precision=precision, lr=lr, device=dev)
    mem = MemoryStats(dev)
    mem.reset_peak()
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    t0 = time.time()
    for b in batches:
        b = b.to(dev)
        loss = eng.model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous(), ignore_index=pad_id)
        eng.backward_and_step(loss)
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    dt = time.time() - t0
    s = mem.sample()
    print(f"AMP Training Time for {len(batches)} batches: {dt:.2f} sec")
    print(f"Peak GPU Memory Usage: {s.peak_mb:.2f} MB")
    return {"time_s": dt, "peak_mb": s.peak_mb, "batches": len(batches)}


def train_one_batch(model, batch, *, precision: str = "fp32", lr: float = 2e-4, pad_id: int = synth.GPT2_PAD, device=None) -> dict:
    """One LM step with allocated-before / allocated-after / PEAK memory (the reference omits the peak, A-16)."""
    dev = _device(device)
    eng = build_engine(model, strategy="single", precision=precision, lr=lr, device=dev)
    mem = MemoryStats(dev)
    mem.reset_peak()
    before = mem.sample()
    b = batch.to(dev)
    loss = eng.model.loss(b[:, :-1].contiguous(), b[:, 1:].contiguous(), ignore_index=pad_id)
    eng.backward_and_step(loss)
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)
    after = mem.sample()
    return {"loss": float(loss.item()), "allocated_before_mb": before.allocated_mb, "allocated_after_mb": after.allocated_mb,
            "peak_mb": after.peak_mb}
