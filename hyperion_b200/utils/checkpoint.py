"""Checkpoint formats of the reference, plus the resume path it lacks.

Export formats kept byte-compatible in layout:
* ``{run_id}_model.pt``            ``{"model_state_dict": <unwrapped full state_dict>}`` (DDP: no ``module.``
                                    prefix, distributed_utils.py:195-199; FSDP: gathered to rank 0, :371-385)
* bare ``state_dict`` files         ``simple_transformer_lm.pt`` / ``resnet18_cifar10.pt`` (core_framework.ipynb c9)
* AMP dicts                         ``{"model_state_dict", "scaler_state_dict"}`` (mixed_precision.ipynb c5 L7-15)
* ``{run_id}_model_sharded.pt``    per-rank shard files (the reference's fallback only saved rank 0's shard)

New: ``save_training_state`` / ``load_training_state`` (model + optimizer + scaler + epoch + RNG) behind
``--resume``, and a manifest with sizes and SHA-256 hashes (the README promises hashes and size checks
that no reference code implements, A-2).
"""
from __future__ import annotations

import hashlib
import json
import os
from pathlib import Path

import torch
import torch.distributed as dist


def unwrap(model):
    return getattr(model, "module", model)


def full_state_dict(model, to_cpu: bool = True, dtype: torch.dtype | None = torch.float32) -> dict:
    """Unwrapped state_dict (all ranks may call; for HyperionFSDP every rank must call -- it gathers)."""
    if hasattr(model, "sync_buffers"):
        model.sync_buffers()
    if hasattr(model, "gather_full_state_dict"):
        sd = model.gather_full_state_dict()
    else:
        sd = unwrap(model).state_dict()
    out = {}
    for k, v in sd.items():
        t = v.detach()
        if dtype is not None and t.is_floating_point():
            t = t.to(dtype)
        out[k] = t.cpu() if to_cpu else t.clone()
    return out


def _manifest(path: Path) -> None:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    man = {"file": path.name, "bytes": path.stat().st_size, "sha256": h.hexdigest()}
    Path(str(path) + ".manifest.json").write_text(json.dumps(man, indent=1))


def verify(path) -> bool:
    path = Path(path)
    man = json.loads(Path(str(path) + ".manifest.json").read_text())
    h = hashlib.sha256(path.read_bytes()).hexdigest()
    return man["bytes"] == path.stat().st_size and man["sha256"] == h


def save_model(model, path, rank: int = 0, wrap: bool = True, scaler=None, dtype=torch.float32) -> None:
    """All ranks call (collective for FSDP); rank 0 writes ``{"model_state_dict": ...}``."""
    sd = full_state_dict(model, dtype=dtype)
    if rank == 0:
        path = Path(path)
        path.parent.mkdir(parents=True, exist_ok=True)
        obj = {"model_state_dict": sd} if wrap else sd
        if scaler is not None and wrap:
            obj["scaler_state_dict"] = scaler.state_dict()
        torch.save(obj, path)
        _manifest(path)
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def load_model(model, path, strict: bool = True):
    obj = torch.load(path, map_location="cpu", weights_only=False)
    sd = obj["model_state_dict"] if isinstance(obj, dict) and "model_state_dict" in obj else obj
    return unwrap(model).load_state_dict(sd, strict=strict)


def save_sharded(model, optimizer, path_prefix, rank: int) -> Path:
    """Per-rank shard file: local optimizer shard + (for FSDP) the local parameter shard."""
    p = Path(f"{path_prefix}_rank{rank}.pt")
    p.parent.mkdir(parents=True, exist_ok=True)
    obj = {"rank": rank, "optimizer": optimizer.state_dict() if optimizer is not None else None}
    if hasattr(model, "local_shard_state"):
        obj["param_shard"] = model.local_shard_state()
    torch.save(obj, p)
    return p


def save_training_state(path, model, optimizer, epoch: int, scaler=None, rank: int = 0, extra: dict | None = None) -> None:
    sd = full_state_dict(model, dtype=None)
    if rank == 0:
        path = Path(path)
        path.parent.mkdir(parents=True, exist_ok=True)
        torch.save({
            "model_state_dict": sd, "optimizer_state_dict": optimizer.state_dict() if optimizer is not None else None,
            "scaler_state_dict": scaler.state_dict() if scaler is not None else None, "epoch": epoch,
            "rng_cpu": torch.get_rng_state(), "rng_cuda": torch.cuda.get_rng_state_all() if torch.cuda.is_available() else None,
            "extra": extra or {},
        }, path)
        _manifest(path)
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def load_training_state(path, model, optimizer=None, scaler=None) -> int:
    """Returns the epoch to resume from (saved epoch + 1)."""
    obj = torch.load(path, map_location="cpu", weights_only=False)
    unwrap(model).load_state_dict(obj["model_state_dict"])
    if optimizer is not None and obj.get("optimizer_state_dict") is not None:
        try:
            optimizer.load_state_dict(obj["optimizer_state_dict"])
        except Exception:
            pass  # a different world size changes shard shapes; weights are still restored
    if scaler is not None and obj.get("scaler_state_dict") is not None:
        scaler.load_state_dict(obj["scaler_state_dict"])
    if obj.get("rng_cpu") is not None:
        torch.set_rng_state(obj["rng_cpu"])
    return int(obj["epoch"]) + 1
