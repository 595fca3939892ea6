#!/usr/bin/env python3
"""Headline benchmark (driver contract; see the task statement and BASELINE.json).

    python bench.py --gpus 1 --steps 10 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus 8 --steps 10 --warmup 3

Flagship = BASELINE config 4: GPT-2 medium (354.8 M parameters), FSDP full-shard state, bf16, WikiText-2-shaped
synthetic tokens, sequence length 1024, weak scaling (fixed per-GPU batch); metric = whole-job tokens/s, device-timed
with CUDA events, max over ranks.  ``--config resnet50_ddp | vit_ddp | lm_ddp | llama_lora`` select the other
BASELINE configs.  Prints ONE JSON line on rank 0.

``--impl reference`` drives the UNMODIFIED reference trainers installed in ``baseline/_ref``
(``baseline/install_reference.sh``, ``baseline/reference_arm.py``) for the jobs the reference contains and this image
can import: ``--config lm_ddp | lm_fsdp | cifar_ddp``.  Its GPT-2 / Llama path needs ``accelerate`` and ``peft`` (absent),
so for ``gpt2_fsdp`` / ``llama_*`` / ``resnet50_ddp`` / ``vit_ddp`` that arm prints ``{"impl": "reference", "unavailable": ...}``
and ``--impl torch`` (stock torch FSDP/DDP + NCCL + cuBLAS, ``baseline/torch_baseline.py``) is the operative baseline.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["hyperion", "reference", "torch"], default="hyperion")
    ap.add_argument("--config", default="gpt2_fsdp", choices=["gpt2_fsdp", "resnet50_ddp", "vit_ddp", "lm_ddp", "lm_fsdp", "cifar_ddp", "llama_lora", "llama_fsdp"])
    ap.add_argument("--batch", type=int, default=None, help="per-GPU batch")
    ap.add_argument("--seq", type=int, default=None)
    ap.add_argument("--size", default=None, help="model size override (gpt2: medium/small/tiny; llama: 7b/tiny)")
    ap.add_argument("--no-fused-comm", dest="fused_comm", action="store_false", default=None)
    ap.add_argument("--no-e2e", dest="e2e", action="store_false", default=True)
    ap.add_argument("--ag-blocks", action="store_true", help="FSDP: block-sized units gather inside their first GEMM (K-C3 A/B)")
    ap.add_argument("--reshard", action="store_true", help="FSDP configs: FULL_SHARD semantics (free after forward, re-gather in backward)")
    ap.add_argument("--no-graph", dest="graph", action="store_false", default=True, help="do not capture the step in a CUDA graph")
    return ap.parse_args()


def main() -> int:
    args = parse()
    if args.impl == "reference":
        from baseline.reference_arm import run as run_reference

        return run_reference(args)
    if args.impl == "torch":
        from baseline.torch_baseline import run as run_torch

        return run_torch(args)

    import torch
    import torch.distributed as dist

    from hyperion_b200.ops import _lib
    from hyperion_b200.parallel.runtime import max_over_ranks, setup
    from hyperion_b200.utils.timing import ClockSampler

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if world != args.gpus and rank == 0:
        print(f"[bench] warning: --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)
    if not torch.cuda.is_available():
        print(json.dumps({"error": "no CUDA device", "metric": "tokens/s"}))
        return 1
    if world > 1:
        dev = setup(rank, world, "nccl")
    else:
        dev = torch.device("cuda", 0)
        torch.cuda.set_device(dev)
        torch.manual_seed(0)
    assert _lib.available(), "libhyperion_b200.so must be built (python -m hyperion_b200.build)"

    from hyperion_b200.bench.cases import make_case

    if args.reshard:
        os.environ["HYPERION_FSDP_RESHARD"] = "1"
    if args.ag_blocks:
        os.environ["HYPERION_FSDP_AG_BLOCKS"] = "1"
    case = make_case(args.config, dev, rank, world, batch=args.batch, seq=args.seq, size=args.size,
                     fused_comm=args.fused_comm)
    if getattr(case.eng.model, "fused", False) and hasattr(case.eng.model, "_ag_handles"):
        case.config["comm"] = (f"fused NVLink kernels: reduce-scatter(+AdamW), all-gather, "
                               f"ag_gemm x{len(case.eng.model._ag_handles)} units")
    if hasattr(case.eng.model, "reshard"):
        case.config["fsdp_mode"] = ("full_shard(reshard_after_forward)" if case.eng.model.reshard else
                                    "shard optimizer state + gradients, 16-bit parameters resident (one all-gather per step)")

    graphed = False
    if args.graph:
        graphed = case.enable_graph()
    case.config["cuda_graph"] = graphed

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    # ---------------- device-timed region: inputs resident on the device
    for _ in range(args.warmup):
        case.step(case.device_batch())
    sync_all()
    sampler = ClockSampler(dev.index or 0).start() if rank == 0 else None
    launches0 = _lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    import time as _time

    e0.record()
    h0 = _time.perf_counter()
    for _ in range(args.steps):
        case.step(case.device_batch())
    host_ms = (_time.perf_counter() - h0) * 1e3 / args.steps  # CPU time to ENQUEUE one step (launch-bound check)
    e1.record()
    sync_all()
    launches = _lib.launch_count() - launches0
    ms = max_over_ranks(e0.elapsed_time(e1), dev)
    clocks = sampler.stop() if sampler is not None else None
    ms_per_step = ms / args.steps
    value = case.items_per_step * world / (ms_per_step / 1e3)

    # ---------------- end-to-end region: pinned host batch -> H2D every step, loss D2H every step
    e2e = None
    if args.e2e:
        # the public input pipeline: pinned host batches, DevicePrefetcher issues the copy of batch i+1 on a side stream
        # before the loop blocks on loss i; every step still copies its own inputs and reads its loss back
        from hyperion_b200.data import DevicePrefetcher

        host_batches = [case.host_batch(i) for i in range(max(2, min(args.steps, 4)))]
        for db in DevicePrefetcher((host_batches[i % len(host_batches)] for i in range(2)), dev):
            float(case.step(db).item())
        sync_all()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        for db in DevicePrefetcher((host_batches[i % len(host_batches)] for i in range(args.steps)), dev):
            float(case.step(db).item())
        f1.record()
        sync_all()
        ms2 = max_over_ranks(f0.elapsed_time(f1), dev) / args.steps
        e2e = {"value": case.items_per_step * world / (ms2 / 1e3), "unit": case.unit, "ms_per_step": ms2,
               "h2d_bytes_per_step": case.h2d_bytes, "d2h_bytes_per_step": case.d2h_bytes}

    if rank == 0:
        out = {
            "metric": case.metric, "value": value, "unit": case.unit, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": case.dtype, "data": "synthetic", "impl": "hyperion",
            "config": case.config, "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "host_enqueue_ms_per_step": host_ms,
            "peak_mem_gb": torch.cuda.max_memory_allocated(dev) / 2 ** 30,
        }
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
