"""Process-group runtime: the reference's ``_local_gpu`` / ``setup`` / ``cleanup``
(02_development/distributed_utils.py:96-125) with a working CPU/gloo path.

Differences from the reference (deliberate):
* ``backend="gloo"`` really works: ``torch.cuda.set_device`` is only called when a GPU exists.
* ``backend=None`` picks nccl with CUDA, gloo otherwise.
* seeds are set identically on every rank (the reference never seeds, A-2 / A-7).
"""
from __future__ import annotations

import datetime
import os

import torch
import torch.distributed as dist


def _local_gpu(rank: int) -> int:
    """Map global rank -> local device index."""
    n = torch.cuda.device_count()
    if n == 0:
        return 0
    return int(os.environ.get("LOCAL_RANK", rank % n)) % n


def default_backend() -> str:
    return "nccl" if torch.cuda.is_available() else "gloo"


def device_for(rank: int) -> torch.device:
    return torch.device("cuda", _local_gpu(rank)) if torch.cuda.is_available() else torch.device("cpu")


def setup(rank: int, world_size: int, backend: str | None = None, timeout_min: float = 10.0, seed: int | None = 0) -> torch.device:
    """Initialise the default process group from torchrun's env (``env://``) and bind the device."""
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    backend = backend or default_backend()
    dev = device_for(rank) if backend != "gloo" or torch.cuda.is_available() else torch.device("cpu")
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    if not dist.is_initialized():
        kwargs = {}
        if backend == "nccl" and dev.type == "cuda":
            kwargs["device_id"] = dev
        dist.init_process_group(backend, rank=rank, world_size=world_size, init_method="env://",
                                timeout=datetime.timedelta(minutes=timeout_min), **kwargs)
    if seed is not None:
        torch.manual_seed(seed)
        if dev.type == "cuda":
            torch.cuda.manual_seed_all(seed)
            from ..ops import rng

            rng.manual_seed(seed)   # the native dropout kernels' (seed, step) state; ranks differ through the device index
    if rank == 0:
        print(f"[Rank 0] PG ready on {os.environ['MASTER_ADDR']}:{os.environ['MASTER_PORT']} (backend={backend})")
    return dev


_holds = 0


def hold_process_group() -> None:
    """Keep the process group alive across several trainers run in ONE process (``--model all``): until the matching
    ``release_process_group()``, ``cleanup()`` only synchronises.  The reference tears the group down and re-creates it between
    its four jobs (run_distributed.py:134-139 -> distributed_utils.py:101-125); re-initialising ``env://`` in the same
    process races the previous store's teardown (gloo refuses the connection), and the peer-memory arenas are tied to the group."""
    global _holds
    _holds += 1


def release_process_group() -> None:
    global _holds
    _holds = max(0, _holds - 1)
    if _holds == 0:
        cleanup()


def cleanup() -> None:
    if dist.is_initialized():
        dist.barrier()
        if _holds == 0:
            dist.destroy_process_group()


def world() -> tuple[int, int]:
    """(rank, world_size); (0, 1) when no process group exists."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def max_over_ranks(value: float, device=None) -> float:
    """Max-reduce a host scalar over all ranks (device-timed step times are reported this way)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    dev = device if device is not None else (torch.device("cuda", torch.cuda.current_device())
                                             if dist.get_backend() == "nccl" else torch.device("cpu"))
    t = torch.tensor([float(value)], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
