"""FSDP trainers: ``train_language_model_fsdp`` (reference: 02_development/distributed_utils.py:290-406) and the
GPT-2 job BASELINE config 4 adds.

Reference recipe kept: SimpleTransformerLM, FULL_SHARD-style sharding with a size-based wrap policy
(``min_num_params=100_000`` -> 8 units), bf16 parameters / reduction, AdamW 1e-4, **no GradScaler**,
``clip_grad_norm_(1.0)``, CSV ``epoch,loss,duration,gpus``, final ``{"model_state_dict": full fp32 dict}``.
Corrected: the clip uses the exact GLOBAL norm (A-4), every rank takes part in the state-dict gather (A-3), all
ranks start from the same weights (A-7).
"""
from __future__ import annotations

import os

import torch

from ..data import DeviceResidentLoader, synth
from ..models import GPT2Config, GPT2LMHeadModel, SimpleTransformerLM
from ..models.gpt2 import Block as GPT2Block
from ..parallel.runtime import cleanup, setup
from ..utils.checkpoint import load_training_state, save_model, save_sharded, save_training_state
from ..utils.metrics import CsvLogger
from .common import EpochTimer, StepRunner, load_wikitext_tensors, mean_over_ranks
from .engine import build_engine


def train_language_model_fsdp(rank: int, world_size: int, epochs: int = 5, base_dir: str = ".", *,
                              backend: str | None = None, precision: str | None = None, synthetic: bool = True,
                              steps_per_epoch: int | None = None, batch_size: int = 32, rows: int | None = None,
                              fused_comm: bool | None = None, seed: int = 0, save: bool = True, clip: float = 1.0,
                              model_kwargs: dict | None = None, vocab_size: int = synth.GPT2_VOCAB,
                              pad_id: int = synth.GPT2_PAD, min_num_params: int = 100_000, resume: str | None = None,
                              graph: bool | None = None, reshard_after_forward: bool | None = None):
    dev = setup(rank, world_size, backend, seed=seed)
    precision = precision or ("bf16" if dev.type == "cuda" else "fp32")
    out_dir = os.path.join(base_dir, "data/distributed")
    logger = CsvLogger(out_dir, "language_fsdp", world_size, rank)
    ids, _ = load_wikitext_tensors(base_dir, dev, synthetic, rows, vocab=vocab_size, pad_id=pad_id, seed=seed)
    loader = DeviceResidentLoader((ids,), batch_size, rank, world_size, shuffle=True, seed=seed)
    model = SimpleTransformerLM(vocab_size, **(model_kwargs or {}))
    eng = build_engine(model, strategy="fsdp", precision=precision, lr=1e-4, clip=clip, device=dev,
                       fused_comm=fused_comm, fsdp_min_num_params=min_num_params, reshard_after_forward=reshard_after_forward)
    start_epoch = load_training_state(resume, eng.model, None, eng.scaler) if resume else 0
    eng.model.train()
    runner = StepRunner(eng, lambda b: eng.model.loss(b[0][:, :-1].contiguous(), b[0][:, 1:].contiguous(), ignore_index=pad_id),
                        graph=graph)
    timer = EpochTimer(dev)
    history = []
    for ep in range(start_epoch, epochs):
        loader.set_epoch(ep)
        total = torch.zeros((), device=dev)
        nsteps = 0
        timer.start()
        for batch in loader:
            loss = runner(batch)
            total += loss.float()
            nsteps += 1
            if steps_per_epoch and nsteps >= steps_per_epoch:
                break
        wall_s, dev_ms = timer.stop()
        avg = mean_over_ranks(total / max(nsteps, 1))
        history.append(avg)
        if rank == 0:
            print(f"[LM-FSDP] Epoch {ep + 1}/{epochs} loss={avg:.4f} time={wall_s:.2f}s")
        logger.log(ep + 1, avg, wall_s, world_size)
        toks = nsteps * batch_size * world_size * (ids.shape[1] - 1)
        logger.log_device(ep + 1, nsteps, dev_ms / max(nsteps, 1), toks / (dev_ms / 1e3), "tokens/s",
                          torch.cuda.max_memory_allocated(dev) / 2 ** 20 if dev.type == "cuda" else 0.0, world_size)
    if save:
        save_model(eng.model, os.path.join(out_dir, f"{logger.run_id}_model.pt"), rank)          # all ranks gather
        save_sharded(eng.model, None, os.path.join(out_dir, f"{logger.run_id}_model_sharded"), rank)
        save_training_state(os.path.join(out_dir, f"{logger.run_id}_train_state.pt"), eng.model, None, epochs - 1,
                            eng.scaler, rank)     # + one optimizer shard file per rank
    cleanup()
    return history


def train_gpt2_fsdp(rank: int, world_size: int, epochs: int = 1, base_dir: str = ".", *, backend: str | None = None,
                    precision: str = "bf16", size: str = "medium", seq_len: int = 1024, batch_size: int = 8,
                    steps_per_epoch: int = 20, fused_comm: bool | None = None, seed: int = 0, clip: float = 1.0,
                    save: bool = False, checkpointing: bool = False, resume: str | None = None, graph: bool | None = None,
                    reshard_after_forward: bool | None = None, model_id: str | None = None):
    """GPT-2 (medium by default) under HyperionFSDP on WikiText-2-shaped synthetic tokens, seq 1024 (config 4).  ``model_id``: a local
    Hugging Face GPT-2 checkpoint directory to start from (random weights of ``size`` otherwise)."""
    dev = setup(rank, world_size, backend, seed=seed)
    from ..models.hf import gpt2_from_pretrained, is_checkpoint_dir, read_config

    pretrained = is_checkpoint_dir(model_id) and read_config(model_id).get("model_type") == "gpt2"
    cfg = {"medium": GPT2Config.medium, "small": GPT2Config.small, "tiny": GPT2Config.tiny}[size]()
    seq_len = min(seq_len, cfg.n_positions)
    out_dir = os.path.join(base_dir, "data/distributed")
    logger = CsvLogger(out_dir, "gpt2_fsdp", world_size, rank, run_id=None)
    if pretrained:
        model = gpt2_from_pretrained(model_id)
        cfg = model.config
        seq_len = min(seq_len, cfg.n_positions)
        if rank == 0:
            print(f"[GPT2-FSDP] weights read from {model_id}")
    else:
        model = GPT2LMHeadModel(cfg)
    model.use_checkpoint = checkpointing
    units = [m for m in model.modules() if isinstance(m, GPT2Block)]
    eng = build_engine(model, strategy="fsdp", precision=precision if dev.type == "cuda" else "fp32", lr=1e-4, clip=clip,
                       device=dev, fused_comm=fused_comm, fsdp_units=units, reshard_after_forward=reshard_after_forward)
    start_epoch = load_training_state(resume, eng.model, None, eng.scaler) if resume else 0
    runner = StepRunner(eng, lambda b: eng.model.loss(b[0]), graph=graph)
    n = batch_size * world_size * steps_per_epoch
    ids, _ = synth.token_batch(n, seq_len, cfg.vocab_size, dev, seed)
    loader = DeviceResidentLoader((ids,), batch_size, rank, world_size, shuffle=True, seed=seed)
    timer = EpochTimer(dev)
    history = []
    for ep in range(start_epoch, epochs):
        loader.set_epoch(ep)
        total = torch.zeros((), device=dev)
        nsteps = 0
        timer.start()
        for batch in loader:
            loss = runner(batch)
            total += loss.float()
            nsteps += 1
        wall_s, dev_ms = timer.stop()
        avg = mean_over_ranks(total / max(nsteps, 1))
        history.append(avg)
        toks = nsteps * batch_size * world_size * seq_len
        if rank == 0:
            print(f"[GPT2-FSDP] Epoch {ep + 1}/{epochs} loss={avg:.4f} time={wall_s:.2f}s tokens/s={toks / (dev_ms / 1e3):.0f}")
        logger.log(ep + 1, avg, wall_s, world_size)
        logger.log_device(ep + 1, nsteps, dev_ms / max(nsteps, 1), toks / (dev_ms / 1e3), "tokens/s",
                          torch.cuda.max_memory_allocated(dev) / 2 ** 20 if dev.type == "cuda" else 0.0, world_size)
    if save:
        save_model(eng.model, os.path.join(out_dir, f"{logger.run_id}_model.pt"), rank)
        save_training_state(os.path.join(out_dir, f"{logger.run_id}_train_state.pt"), eng.model, None, epochs - 1,
                            eng.scaler, rank)
    cleanup()
    return history
