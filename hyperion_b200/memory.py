"""Activation checkpointing and memory statistics.

* ``checkpoint_sequential(functions, segments, input)`` -- same call shape as
  ``torch.utils.checkpoint.checkpoint_sequential`` used by the reference on the LM's
  encoder layers (02_development/memory_optimization.ipynb c3 L18) and on ResNet
  ``[layer1..layer4]`` (c4 L8).  Recompute replays the forward kernels.
* ``MemoryStats`` / ``print_memory`` -- allocated / reserved MB like the reference's
  ``print_memory`` (memory_optimization.ipynb c2 L1-7) plus the *peak*, reset per phase,
  which the reference never records.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
from torch.utils.checkpoint import checkpoint


class _SaltKeeper:
    """Runs ``fn`` for the forward, and again for the recompute in backward with the dropout salt counter rewound to where
    the forward started, so the native Philox dropout kernels (ops/rng.py) regenerate the forward's masks."""

    def __init__(self, fn):
        self.fn, self.salt = fn, None

    def __call__(self, *a, **k):
        from .ops import rng

        if self.salt is None:
            self.salt = rng.get_salt()
            return self.fn(*a, **k)
        cur = rng.get_salt()
        rng.set_salt(self.salt)
        try:
            return self.fn(*a, **k)
        finally:
            rng.set_salt(cur)


def checkpoint_layers(layers, x, **kwargs):
    """Checkpoint every layer of an iterable (one segment per layer)."""
    for layer in layers:
        x = checkpoint(_SaltKeeper(layer), x, use_reentrant=False, **kwargs)
    return x


def checkpoint_sequential(functions, segments: int, input, **kwargs):
    functions = list(functions)
    n = len(functions)
    segments = max(1, min(segments, n))
    per = (n + segments - 1) // segments

    def run(start, end):
        def fn(t):
            for f in functions[start:end]:
                t = f(t, **kwargs) if kwargs else f(t)
            return t
        return fn

    x = input
    for s in range(0, n, per):
        x = checkpoint(_SaltKeeper(run(s, min(n, s + per))), x, use_reentrant=False)
    return x


def enable_checkpointing(model: torch.nn.Module, on: bool = True) -> int:
    """Set ``use_checkpoint`` on every sub-module that supports it; returns how many were switched."""
    k = 0
    for m in model.modules():
        if hasattr(m, "use_checkpoint") and not isinstance(getattr(type(m), "use_checkpoint", None), property):
            m.use_checkpoint = on
            k += 1
    return k


def checkpoint_resnet_blocks(model: torch.nn.Module) -> torch.nn.Module:
    """The reference's helper of that name (memory_optimization.ipynb c4 L1-14: it replaces ``forward`` by one that runs
    ``[layer1..layer4]`` through ``checkpoint_sequential``).  The ResNets here carry the switch already, so this flips
    ``use_checkpoint`` and returns the same module; a torchvision ResNet gets the wrapped forward."""
    if hasattr(model, "use_checkpoint"):
        model.use_checkpoint = True
        return model
    stages = [model.layer1, model.layer2, model.layer3, model.layer4]

    def forward_with_ckpt(x):
        x = model.maxpool(model.relu(model.bn1(model.conv1(x))))
        x = checkpoint_sequential(stages, len(stages), x) if torch.is_grad_enabled() else torch.nn.Sequential(*stages)(x)
        return model.fc(torch.flatten(model.avgpool(x), 1))

    model.forward = forward_with_ckpt
    return model


@dataclass
class MemorySample:
    allocated_mb: float
    reserved_mb: float
    peak_mb: float


class MemoryStats:
    def __init__(self, device=None):
        self.device = device
        self.enabled = torch.cuda.is_available()

    def reset_peak(self) -> None:
        if self.enabled:
            torch.cuda.reset_peak_memory_stats(self.device)

    def sample(self) -> MemorySample:
        if not self.enabled:
            return MemorySample(0.0, 0.0, 0.0)
        mb = 1024 ** 2
        return MemorySample(torch.cuda.memory_allocated(self.device) / mb, torch.cuda.memory_reserved(self.device) / mb,
                            torch.cuda.max_memory_allocated(self.device) / mb)


def get_gpu_memory():
    """(allocated MB, reserved MB) -- the probe's helper (01_hardware_exploration.ipynb c1 L25-28)."""
    s = MemoryStats().sample()
    return s.allocated_mb, s.reserved_mb


def print_memory(prefix: str = "") -> MemorySample:
    s = MemoryStats().sample()
    print(f"{prefix} Allocated: {s.allocated_mb:.2f} MB | Reserved: {s.reserved_mb:.2f} MB | Peak: {s.peak_mb:.2f} MB")
    return s
