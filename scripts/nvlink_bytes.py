"""NVLink bytes per training step, read from the NVML link counters, against the algorithmic bytes of the collectives.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 \
        scripts/nvlink_bytes.py --config gpt2_fsdp [--reshard] [--ag-blocks] [--steps 20]

Every rank reads the data TX / RX counters of its own GPU (all links summed, NVML field values) before and after K steps of
the bench case (CUDA graph, same code path as bench.py) and prints KiB per step.  Algorithmic bytes per step and GPU for an
FSDP engine with S bytes of 16-bit parameters on W ranks: reduce-scatter of the gradients receives S (W-1)/W, the parameter
all-gather receives S (W-1)/W (twice that in FULL_SHARD mode); a DDP all-reduce moves 2 S (W-1)/W each way (two-shot) or
~S with the in-switch NVLS reduction.  ncu cannot wrap a multi-rank command, so this is the byte evidence for the
communication kernels (profiling recipe: counters instead of a profiler).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist


def nvlink_kib(handle, nv):
    """(tx KiB, rx KiB) summed over all links of one GPU."""
    ids = [nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, nv.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX]
    try:
        vals = nv.nvmlDeviceGetFieldValues(handle, [(i, 0xFFFFFFFF) for i in ids])   # scope = all links
    except TypeError:
        vals = nv.nvmlDeviceGetFieldValues(handle, ids)
    out = []
    for v in vals:
        out.append(int(v.value.ullVal) if v.nvmlReturn == 0 else -1)
    return tuple(out)


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="gpt2_fsdp")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--reshard", action="store_true")
    ap.add_argument("--ag-blocks", action="store_true")
    a = ap.parse_args()
    if a.reshard:
        os.environ["HYPERION_FSDP_RESHARD"] = "1"
    if a.ag_blocks:
        os.environ["HYPERION_FSDP_AG_BLOCKS"] = "1"
    import pynvml as nv

    from hyperion_b200.bench.cases import make_case
    from hyperion_b200.parallel.runtime import setup

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = setup(rank, world, "nccl")
    nv.nvmlInit()
    try:
        h = nv.nvmlDeviceGetHandleByUUID(("GPU-" + str(torch.cuda.get_device_properties(dev).uuid)).encode())
    except Exception:  # noqa: BLE001
        h = nv.nvmlDeviceGetHandleByIndex(dev.index)
    case = make_case(a.config, dev, rank, world)
    case.enable_graph()
    for _ in range(5):
        case.step(case.device_batch())
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    t0, r0 = nvlink_kib(h, nv)
    for _ in range(a.steps):
        case.step(case.device_batch())
    torch.cuda.synchronize(dev)
    dist.barrier()
    torch.cuda.synchronize(dev)
    t1, r1 = nvlink_kib(h, nv)
    eng = case.eng
    nbytes = eng.flat.numel * 2 if eng.flat is not None else 0
    if hasattr(eng.model, "shard_numel"):
        nbytes = eng.model.shard_numel * world * 2
    frac = (world - 1) / world
    rec = {"rank": rank, "world": world, "config": a.config, "reshard": a.reshard, "ag_blocks": a.ag_blocks, "steps": a.steps,
           "tx_mib_per_step": (t1 - t0) / 1024 / a.steps, "rx_mib_per_step": (r1 - r0) / 1024 / a.steps,
           "param_bytes_16bit_mib": nbytes / 2 ** 20,
           "algorithmic_rx_mib_per_step": {"reduce_scatter": nbytes * frac / 2 ** 20,
                                           "all_gather": nbytes * frac / 2 ** 20 * (2 if a.reshard else 1)},
           "counters_valid": min(t0, r0, t1, r1) >= 0}
    print("NVLINK " + json.dumps(rec), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
