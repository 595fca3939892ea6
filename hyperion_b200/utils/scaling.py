"""Scaling-experiment orchestrator: the reference's ``run_scaling_experiment``
(02_development/distributed_utils.py:780-831) -- for each GPU count launch
``torchrun --standalone --nproc_per_node=N run_distributed.py --model=...``, then build the report.
Here it is a plain launcher (not run from inside a torchrun worker, A-15) and uses a fresh rendezvous
port per run."""
from __future__ import annotations

import os
import subprocess
import sys
import time
from pathlib import Path

from .report import create_scaling_report

ROOT = Path(__file__).resolve().parent.parent.parent


def run_scaling_experiment(model: str, gpu_counts, epochs: int = 3, base_dir: str = ".", extra_args=(),
                           port0: int = 29600, sleep_s: float = 1.0, dry_run: bool = False) -> list[list[str]]:
    cmds = []
    for i, n in enumerate(gpu_counts):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
               "--master-addr", "127.0.0.1", "--master-port", str(port0 + i), str(ROOT / "run_distributed.py"),
               f"--model={model}", f"--epochs={epochs}", f"--base_dir={base_dir}", *extra_args]
        cmds.append(cmd)
        if dry_run:
            continue
        print(f"\n### Scaling run: {model} on {n} GPU(s)\n", flush=True)
        try:
            subprocess.run(cmd, check=True)
        except subprocess.CalledProcessError as e:
            print(f"run with {n} GPUs failed: {e}")
        time.sleep(sleep_s)
    if not dry_run:
        create_scaling_report(os.path.join(base_dir, "data/distributed"))
    return cmds
