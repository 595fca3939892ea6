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
from pathlib import Path

import torch
import torch.distributed as dist


def unwrap(model):
    return getattr(model, "module", model)


def full_state_dict(model, to_cpu: bool = True, dtype: torch.dtype | None = torch.float32, optimizer=None) -> dict:
    """Unwrapped state_dict (all ranks may call; for HyperionFSDP every rank must call -- it gathers).  With
    ``optimizer`` (a ``FusedAdamW``) the fp32 master weights replace the upcast 16-bit compute copy, so the file holds
    true fp32 parameters like the reference's AMP checkpoints."""
    from ..ops.conv import flush_bn_counters

    flush_bn_counters()
    if hasattr(model, "sync_buffers"):
        model.sync_buffers()
    if hasattr(model, "gather_full_state_dict"):
        sd = model.gather_full_state_dict()
    else:
        sd = dict(unwrap(model).state_dict())
        overlay = optimizer.master_state_overlay() if hasattr(optimizer, "master_state_overlay") else {}
        if overlay:
            by_ptr = {p.data_ptr(): n for n, p in unwrap(model).named_parameters()}
            for k, v in list(sd.items()):
                n = by_ptr.get(v.data_ptr(), k)   # tied parameters share storage
                if n in overlay and overlay[n].shape == v.shape:
                    sd[k] = overlay[n]
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


def save_model(model, path, rank: int = 0, wrap: bool = True, scaler=None, dtype=torch.float32, optimizer=None) -> None:
    """All ranks call (collective for FSDP); rank 0 writes ``{"model_state_dict": ...}``."""
    sd = full_state_dict(model, dtype=dtype, optimizer=optimizer)
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


def _load_weights(model, sd, optimizer=None, strict: bool = True):
    """Weights into the module AND into every derived copy: FSDP's local 16-bit / fp32 master shards, or the
    optimizer's fp32 master (otherwise the first step would write the stale master back over the loaded weights)."""
    if hasattr(model, "refresh_shards"):          # HyperionFSDP
        return model.load_state_dict(sd, strict=strict)
    res = unwrap(model).load_state_dict(sd, strict=strict)
    if optimizer is not None and hasattr(optimizer, "refresh_master"):
        optimizer.refresh_master(sd)
    return res


def load_model(model, path, strict: bool = True, optimizer=None):
    """Weights-only load.  Pass the engine's ``optimizer`` (``FusedAdamW``) so its fp32 master follows the file."""
    obj = torch.load(path, map_location="cpu", weights_only=False)
    sd = obj["model_state_dict"] if isinstance(obj, dict) and "model_state_dict" in obj else obj
    return _load_weights(model, sd, optimizer, strict)


def save_sharded(model, optimizer, path_prefix, rank: int) -> Path:
    """Per-rank shard file: local optimizer shard + (for FSDP) the local parameter shard."""
    p = Path(f"{path_prefix}_rank{rank}.pt")
    p.parent.mkdir(parents=True, exist_ok=True)
    obj = {"rank": rank, "optimizer": optimizer.state_dict() if optimizer is not None else None}
    if hasattr(model, "local_shard_state"):
        obj["param_shard"] = model.local_shard_state()
    torch.save(obj, p)
    return p


def _opt_shard_path(path, rank: int) -> Path:
    return Path(f"{path}.optshard_rank{rank}.pt")


def save_training_state(path, model, optimizer, epoch: int, scaler=None, rank: int = 0, extra: dict | None = None) -> None:
    """Collective.  Rank 0 writes model + (replicated) optimizer + scaler + epoch + RNG; under HyperionFSDP every rank
    also writes its optimizer shard next to it (``<path>.optshard_rank<r>.pt``)."""
    sd = full_state_dict(model, dtype=None, optimizer=optimizer)
    if hasattr(model, "optimizer_state_dict"):    # HyperionFSDP: sharded state, one file per rank
        Path(path).parent.mkdir(parents=True, exist_ok=True)
        torch.save({k: (v.detach().cpu() if torch.is_tensor(v) else v) for k, v in model.optimizer_state_dict().items()},
                   _opt_shard_path(path, rank))
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
    import warnings

    obj = torch.load(path, map_location="cpu", weights_only=False)
    _load_weights(model, obj["model_state_dict"], optimizer)
    if optimizer is not None and obj.get("optimizer_state_dict") is not None:
        try:
            optimizer.load_state_dict(obj["optimizer_state_dict"])
        except Exception as e:  # noqa: BLE001 -- e.g. a different flat layout; weights + re-seeded master are still right
            warnings.warn(f"optimizer state in {path} could not be loaded ({e!r}); moments start from zero")
            if hasattr(optimizer, "refresh_master"):
                optimizer.refresh_master(obj["model_state_dict"])
    if hasattr(model, "load_optimizer_state_dict"):   # HyperionFSDP: this rank's shard
        rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        sp = _opt_shard_path(path, rank)
        if sp.exists():
            try:
                model.load_optimizer_state_dict(torch.load(sp, map_location="cpu", weights_only=False))
            except ValueError as e:
                warnings.warn(f"{e}; moments start from zero, weights are restored")
                model.refresh_shards(obj["model_state_dict"])
        else:
            warnings.warn(f"no optimizer shard {sp.name}; moments start from zero")
    if scaler is not None and obj.get("scaler_state_dict") is not None:
        scaler.load_state_dict(obj["scaler_state_dict"])
    if obj.get("rng_cpu") is not None:
        torch.set_rng_state(obj["rng_cpu"])
    if obj.get("rng_cuda") is not None and torch.cuda.is_available():
        try:
            torch.cuda.set_rng_state_all(obj["rng_cuda"][:torch.cuda.device_count()])
        except Exception as e:  # noqa: BLE001
            warnings.warn(f"CUDA RNG state not restored: {e!r}")
    return int(obj["epoch"]) + 1
