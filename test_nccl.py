#!/usr/bin/env python3
"""Communication sanity check -- the reference's ``02_development/test_nccl.py`` (:8-44), but asserting.

    torchrun --standalone --nproc_per_node=N test_nccl.py [--full]

Each rank prints its device, joins the process group (30 s timeout like the reference), all-reduces
``ones(1) * rank`` and CHECKS the result equals W(W-1)/2 (the reference only prints it).  ``--full`` additionally runs
the peer-memory self-test (P2P / NVLS mapping, fused all-reduce / reduce-scatter / all-gather vs NCCL).  On a CPU-only
machine the same check runs over gloo.  Exit code 1 on any failure.
"""
import datetime
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent))

import torch
import torch.distributed as dist


def main() -> int:
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29500")
    cuda = torch.cuda.is_available()
    try:
        if cuda:
            torch.cuda.set_device(local % torch.cuda.device_count())
            dev = torch.device("cuda", local % torch.cuda.device_count())
            print(f"[rank {rank}] device: {torch.cuda.get_device_name(dev)}", flush=True)
        else:
            dev = torch.device("cpu")
            print(f"[rank {rank}] device: cpu (gloo)", flush=True)
        dist.init_process_group("nccl" if cuda else "gloo", rank=rank, world_size=world,
                                timeout=datetime.timedelta(seconds=30))
        t = torch.ones(1, device=dev) * rank
        print(f"[rank {rank}] before all_reduce: {t.item()}", flush=True)
        dist.all_reduce(t)
        print(f"[rank {rank}] after all_reduce: {t.item()}", flush=True)
        expected = world * (world - 1) / 2
        if t.item() != expected:
            print(f"[rank {rank}] FAILED: expected {expected}", flush=True)
            return 1
        if "--full" in sys.argv and cuda and world > 1:
            dist.destroy_process_group()
            from hyperion_b200.parallel.selftest import run

            return 1 if run(False, None) else 0
        dist.destroy_process_group()
        return 0
    except Exception as e:  # noqa: BLE001
        print(f"[rank {rank}] exception: {e!r}", flush=True)
        return 1


if __name__ == "__main__":
    sys.exit(main())
