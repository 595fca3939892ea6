#!/usr/bin/env python3
"""Launcher for the distributed trainers -- same flags and ``--model`` choices as the reference's
``02_development/run_distributed.py`` (:38-67), started the same way:

    torchrun --standalone --nproc_per_node=N run_distributed.py --model language_ddp --epochs 5 --base_dir .

Reference flags kept: --model {language_ddp,cifar,language_fsdp,llama,all,scaling} --epochs --base_dir --hf_token
--model_id --lora --batch_size --progress_every --scaling_gpus.  Additions: --model {resnet50,vit,gpt2}
(BASELINE configs 2-4), --precision, --synthetic / --real-data, --steps, --backend (gloo runs on CPU),
--no-fused-comm (NCCL baseline path), --lora-strategy {fsdp,ddp}, --llama-size, --checkpointing, --seed, --config,
--graph / --no-graph (whole-step CUDA graph, default on CUDA), --resume <train_state.pt>, --reshard (FSDP FULL_SHARD).

Unlike the reference (A-15), ``--model scaling`` is an ordinary launcher: run it WITHOUT torchrun; it spawns one
torchrun per GPU count and then writes the scaling report.
"""
from __future__ import annotations

import argparse
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent))


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Hyperion-B200 launcher")
    p.add_argument("--model", default="language_ddp",
                   choices=["language_ddp", "cifar", "language_fsdp", "llama", "all", "scaling", "resnet50", "vit", "gpt2"])
    p.add_argument("--epochs", type=int, default=5)
    p.add_argument("--base_dir", type=str, default=os.environ.get("HYPERION_BASE_DIR", "."))
    p.add_argument("--hf_token", type=str, default=None, help="accepted for compatibility; nothing is downloaded")
    p.add_argument("--model_id", type=str, default="NousResearch/Llama-2-7b-hf",
                   help="a local Hugging Face checkpoint directory (config.json + safetensors) to start from; a hub name is accepted "
                        "for compatibility and means random weights of that architecture (nothing is downloaded)")
    p.add_argument("--lora", action="store_true", help="LoRA adapters (r=16, alpha=32, q/k/v/o) instead of full fine-tuning")
    p.add_argument("--batch_size", type=int, default=None, help="per-GPU batch size (Llama default 1)")
    p.add_argument("--progress_every", type=int, default=50)
    p.add_argument("--scaling_gpus", type=str, default="1,2,4,8")
    # additions
    p.add_argument("--precision", choices=["fp32", "fp16", "bf16"], default=None)
    p.add_argument("--synthetic", dest="synthetic", action="store_true", default=True)
    p.add_argument("--real-data", dest="synthetic", action="store_false", help="read data/processed/* under --base_dir")
    p.add_argument("--steps", type=int, default=None, help="cap steps per epoch")
    p.add_argument("--backend", choices=["nccl", "gloo"], default=None)
    p.add_argument("--no-fused-comm", dest="fused_comm", action="store_false", default=None)
    p.add_argument("--lora-strategy", choices=["fsdp", "ddp"], default="fsdp")
    p.add_argument("--llama-size", choices=["7b", "tiny"], default="7b")
    p.add_argument("--gpt2-size", choices=["medium", "small", "tiny"], default="medium")
    p.add_argument("--seq-len", type=int, default=None)
    p.add_argument("--checkpointing", action="store_true")
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--config", type=str, default=None, help="JSON file with the default_config.json schema")
    p.add_argument("--no-save", dest="save", action="store_false", default=True)
    p.add_argument("--save", dest="save_gpt2", action="store_true", default=False,
                   help="--model gpt2 only: also write the 1.4 GB model file + resumable train state (the other jobs save by default)")
    p.add_argument("--graph", dest="graph", action="store_true", default=None,
                   help="capture the whole training step in a CUDA graph (default on CUDA; HYPERION_GRAPH=0 also disables it)")
    p.add_argument("--no-graph", dest="graph", action="store_false")
    p.add_argument("--resume", type=str, default=None,
                   help="a <run_id>_train_state.pt written by an earlier run: weights, optimizer state (sharded under FSDP), "
                        "loss scaler, RNG and epoch are restored and training continues at the next epoch")
    p.add_argument("--reshard", dest="reshard", action="store_true", default=None,
                   help="FSDP: free each unit's gathered parameters after its forward and re-gather them in backward "
                        "(FULL_SHARD semantics; default keeps the 16-bit copy resident)")
    return p


def forwarded_args(argv: list[str]) -> list[str]:
    """What a scaling sweep hands down to every run it launches: the command line as given, minus the options the sweep sets itself
    (``--model``, ``--epochs``, ``--base_dir``, ``--scaling_gpus``) -- so ``--backend``, ``--steps``, ``--batch_size``, ``--lora``,
    ``--llama-size`` ... apply to the child runs too."""
    own = ("--model", "--epochs", "--base_dir", "--scaling_gpus")
    out, skip = [], False
    for tok in argv:
        if skip:
            skip = False
            continue
        if tok in own:
            skip = True
            continue
        if tok.split("=", 1)[0] in own:
            continue
        out.append(tok)
    return out


def run_selected(which: str, args, rank: int, world: int) -> None:
    from hyperion_b200 import train

    common = dict(backend=args.backend, synthetic=args.synthetic, steps_per_epoch=args.steps, fused_comm=args.fused_comm,
                  seed=args.seed, save=args.save, resume=args.resume, graph=args.graph)
    bs = {} if args.batch_size is None else {"batch_size": args.batch_size}
    if which == "language_ddp":
        train.train_language_model_ddp(rank, world, epochs=args.epochs, base_dir=args.base_dir, precision=args.precision,
                                       **common, **bs)
    elif which == "cifar":
        train.train_cifar_model_ddp(rank, world, epochs=args.epochs, base_dir=args.base_dir, precision=args.precision,
                                    **common, **bs)
    elif which in ("resnet50", "vit"):
        train.train_cifar_model_ddp(rank, world, epochs=args.epochs, base_dir=args.base_dir,
                                    precision=args.precision or "bf16", arch=which, checkpointing=args.checkpointing or which == "vit",
                                    **common, **bs)
    elif which == "language_fsdp":
        train.train_language_model_fsdp(rank, world, epochs=args.epochs, base_dir=args.base_dir, precision=args.precision,
                                        reshard_after_forward=args.reshard, **common, **bs)
    elif which == "gpt2":
        train.train_gpt2_fsdp(rank, world, epochs=args.epochs, base_dir=args.base_dir, backend=args.backend,
                              precision=args.precision or "bf16", size=args.gpt2_size, seq_len=args.seq_len or 1024,
                              steps_per_epoch=args.steps or 20, fused_comm=args.fused_comm, seed=args.seed,
                              checkpointing=args.checkpointing, save=args.save_gpt2,
                              resume=args.resume, graph=args.graph, reshard_after_forward=args.reshard, model_id=args.model_id, **bs)
    elif which == "llama":
        train.train_llama_fsdp(rank, world, epochs=args.epochs, base_dir=args.base_dir, hf_token=args.hf_token,
                               model_id=args.model_id, lora=args.lora, batch_size=args.batch_size or 1,
                               progress_every=args.progress_every, backend=args.backend, size=args.llama_size,
                               seq_len=args.seq_len or 128, steps_per_epoch=args.steps or 50,
                               lora_strategy=args.lora_strategy, fused_comm=args.fused_comm, seed=args.seed,
                               save=args.save, checkpointing=args.checkpointing, resume=args.resume, graph=args.graph,
                               reshard_after_forward=args.reshard)
    else:
        raise ValueError(f"unknown model: {which}")


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    if args.config:
        from hyperion_b200.config import HyperionConfig

        cfg = HyperionConfig.load(args.config)
        if args.precision is None and cfg.optimization.enable_amp:
            args.precision = cfg.optimization.amp_dtype
        if args.backend is None:
            args.backend = cfg.distributed.backend if cfg.distributed.backend in ("nccl", "gloo") else None
        args.checkpointing = args.checkpointing or cfg.optimization.enable_gradient_checkpointing

    from hyperion_b200.utils.report import create_scaling_report

    if args.model == "scaling":
        from hyperion_b200.utils.scaling import run_scaling_experiment

        gpus = [int(x) for x in args.scaling_gpus.split(",")]
        extra = forwarded_args(sys.argv[1:] if argv is None else list(argv))
        for mdl in ["language_ddp", "cifar", "language_fsdp", "llama"]:
            run_scaling_experiment(mdl, gpus, epochs=args.epochs, base_dir=args.base_dir, extra_args=extra)
        create_scaling_report(os.path.join(args.base_dir, "data/distributed"))
        return 0

    try:
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    except KeyError:
        sys.exit("Please launch me with torchrun -- RANK/WORLD_SIZE not found "
                 "(e.g. torchrun --standalone --nproc_per_node=1 run_distributed.py ...).")
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if local_rank == 0:
        import torch

        print(f"torch {torch.__version__} | rank {rank}/{world} | GPU {local_rank} / node has {torch.cuda.device_count()} devices")
    if args.model == "all":
        from hyperion_b200.parallel.runtime import hold_process_group, release_process_group

        hold_process_group()                # one process group for the four jobs (each trainer's cleanup() only synchronises)
        try:
            for mdl in ["language_ddp", "cifar", "language_fsdp", "llama"]:
                if local_rank == 0:
                    print(f"\n=== [{mdl}] starting ===\n")
                run_selected(mdl, args, rank, world)
        finally:
            release_process_group()
    else:
        run_selected(args.model, args, rank, world)
    if rank == 0:
        create_scaling_report(os.path.join(args.base_dir, "data/distributed"))
    return 0


if __name__ == "__main__":
    sys.exit(main())
