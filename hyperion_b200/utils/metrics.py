"""Per-run CSV metrics with the reference's exact headers and file names
(02_development/distributed_utils.py:140-147, 215-222, 301-306, 438-444):

    run_id = f"{kind}_{world}gpus_{%Y%m%d_%H%M%S}"
    {run_id}_metrics.csv : LM DDP/FSDP  epoch,loss,duration,gpus
                           CIFAR        epoch,loss,accuracy,duration,gpus
                           Llama        epoch,loss,duration_s,gpus,mode

Rank 0 writes; rows are appended per epoch.  Extra device-timed columns (step ms max over ranks,
tokens/s, peak MB) go to a *second* file ``{run_id}_device_metrics.csv`` so the reference's
``create_scaling_report`` keeps parsing the first one unchanged.
"""
from __future__ import annotations

import csv
import datetime
from pathlib import Path

HEADERS = {
    "language_ddp": ["epoch", "loss", "duration", "gpus"],
    "language_fsdp": ["epoch", "loss", "duration", "gpus"],
    "cifar_ddp": ["epoch", "loss", "accuracy", "duration", "gpus"],
    "llama": ["epoch", "loss", "duration_s", "gpus", "mode"],
    # jobs the reference does not have get their own file-name prefix, so a scaling report never mixes two models
    "gpt2_fsdp": ["epoch", "loss", "duration", "gpus"],
    "resnet50_ddp": ["epoch", "loss", "accuracy", "duration", "gpus"],
    "vit_ddp": ["epoch", "loss", "accuracy", "duration", "gpus"],
}
DEVICE_HEADER = ["epoch", "steps", "step_ms_max", "throughput", "unit", "peak_mem_mb", "gpus"]


def make_run_id(kind: str, world_size: int, now: datetime.datetime | None = None) -> str:
    now = now or datetime.datetime.now()
    return f"{kind}_{world_size}gpus_{now.strftime('%Y%m%d_%H%M%S')}"


def shared_run_id(kind: str, world_size: int) -> str:
    """Rank 0's run id on every rank (each rank reading its own clock can straddle a second boundary, and the per-rank
    shard files of one run would then carry different names)."""
    rid = make_run_id(kind, world_size)
    try:
        import torch.distributed as dist

        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            box = [rid]
            dist.broadcast_object_list(box, src=0)
            rid = box[0]
    except Exception:  # noqa: BLE001 -- logging must never take a run down
        pass
    return rid


class CsvLogger:
    def __init__(self, out_dir, kind: str, world_size: int, rank: int = 0, header: list[str] | None = None,
                 run_id: str | None = None):
        self.kind, self.rank, self.world_size = kind, rank, world_size
        self.out_dir = Path(out_dir)
        self.run_id = run_id or shared_run_id(kind, world_size)
        self.header = header or HEADERS[kind]
        self.path = self.out_dir / f"{self.run_id}_metrics.csv"
        self.device_path = self.out_dir / f"{self.run_id}_device_metrics.csv"
        if rank == 0:
            self.out_dir.mkdir(parents=True, exist_ok=True)
            with open(self.path, "w", newline="") as f:
                csv.writer(f).writerow(self.header)
            with open(self.device_path, "w", newline="") as f:
                csv.writer(f).writerow(DEVICE_HEADER)

    def log(self, *row) -> None:
        if self.rank != 0:
            return
        assert len(row) == len(self.header), f"expected {self.header}, got {row}"
        with open(self.path, "a", newline="") as f:
            csv.writer(f).writerow(row)

    def log_device(self, *row) -> None:
        if self.rank != 0:
            return
        with open(self.device_path, "a", newline="") as f:
            csv.writer(f).writerow(row)
