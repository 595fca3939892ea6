"""Random state of the native dropout kernels (``csrc/common/philox.cuh``).

Every mask is a pure function of ``(seed, step, salt, element index)``:

* ``seed`` and ``step`` live in a device tensor (``int64[2]``) that the kernels read, so nothing random is baked into a
  captured CUDA graph.  The training engine calls :func:`advance` once per optimizer step (inside the captured step), so
  every replay draws fresh masks; a backward -- or an activation-checkpoint recompute -- that runs before the advance
  regenerates the forward's mask bit for bit, which is why no mask is ever stored.
* ``salt`` is a host counter handed out per dropout call (:func:`next_salt`); it tells the calls of one step apart.
  Activation checkpointing saves and restores it around a recompute (:func:`get_salt` / :func:`set_salt`).

The reference gets its masks from torch's global Philox generator through ``nn.Dropout`` / SDPA
(``nn.TransformerEncoderLayer`` defaults, 02_development/distributed_utils.py:79; ``lora_dropout=0.05``, :469-473).
"""
from __future__ import annotations

import torch

_state: dict = {}     # device index -> int64[2] tensor (seed, step)
_salt = 0
_seed = 0x5EED_B200


def _rank() -> int:
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized():
            return dist.get_rank()
    except Exception:  # noqa: BLE001
        pass
    return 0


def _seed_for(key: int) -> int:
    # every rank (and every device of a process) draws its own masks
    return _seed + 7919 * key + 104729 * _rank()


def manual_seed(seed: int) -> None:
    """Re-seed every device's dropout state (and restart the per-call salt).  Written in place, so a captured CUDA graph
    that reads the state keeps working."""
    global _seed, _salt
    _seed, _salt = int(seed), 0
    for key, t in _state.items():
        t[0] = _seed_for(key)
        t[1] = 0


def state(device) -> torch.Tensor:
    device = torch.device(device)
    key = device.index if device.index is not None else torch.cuda.current_device()
    t = _state.get(key)
    if t is None:
        t = torch.tensor([_seed_for(key), 0], dtype=torch.int64, device=torch.device("cuda", key))
        _state[key] = t
    return t


def advance(device) -> None:
    """One optimizer step has passed (device-side increment: capturable, no host sync)."""
    device = torch.device(device)
    if device.type == "cuda" and (device.index if device.index is not None else torch.cuda.current_device()) in _state:
        state(device)[1:2].add_(1)


def next_salt() -> int:
    global _salt
    _salt = (_salt + 1) & 0x7FFFFFFF
    return _salt


def get_salt() -> int:
    return _salt


def set_salt(v: int) -> None:
    global _salt
    _salt = int(v)


def reference_keep_mask(seed: int, step: int, salt: int, plane, rows: int, cols: int, p: float):
    """numpy restatement of the kernels' mask (Philox4x32-10, one call per 16 columns, byte >= round(256 p) keeps):
    returns ``(keep[rows, cols] bool, scale)``.  ``plane``: ``0x80000000`` for the 2-D masks, ``b * H + h`` for attention.
    Test oracle only."""
    import numpy as np

    thresh = min(255, max(1, int(p * 256.0 + 0.5)))
    groups = (cols + 15) // 16
    M0, M1, W0, W1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0x9E3779B9), np.uint64(0xBB67AE85)
    mask32 = np.uint64(0xFFFFFFFF)
    g, r = np.meshgrid(np.arange(groups, dtype=np.uint64), np.arange(rows, dtype=np.uint64))
    c0, c1 = g.copy(), r.copy()
    c2 = np.full_like(c0, np.uint64(int(plane) & 0xFFFFFFFF))
    c3 = np.full_like(c0, np.uint64(salt & 0xFFFFFFFF))
    k0, k1 = np.uint64(seed & 0xFFFFFFFF), np.uint64(step & 0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & mask32, p1 >> np.uint64(32), p1 & mask32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & mask32, lo1, (hi0 ^ c3 ^ k1) & mask32, lo0
        k0, k1 = (k0 + W0) & mask32, (k1 + W1) & mask32
    words = np.stack([c0, c1, c2, c3], axis=-1)                       # [rows, groups, 4]
    by = np.stack([(words >> np.uint64(8 * i)) & np.uint64(0xFF) for i in range(4)], axis=-1)  # [rows, groups, 4 words, 4 bytes]
    keep = (by.reshape(rows, groups * 16) >= np.uint64(thresh))[:, :cols]
    return keep, 256.0 / (256 - thresh)
