"""DDP trainers: ``train_language_model_ddp`` and ``train_cifar_model_ddp`` with the reference's signatures,
hyper-parameters, CSV schema and checkpoint layout (02_development/distributed_utils.py:132-200, 208-278):

* LM:    bs 32/GPU, SimpleTransformerLM(50257), AdamW 2e-4, CE(ignore_index=pad 50256), next-token shift
         ``ids[:, :-1] -> ids[:, 1:]``, fp16 AMP; CSV ``epoch,loss,duration,gpus``; ``{run_id}_model.pt``.
* CIFAR: bs 64/GPU, resnet18(num_classes=10), AdamW 1e-3, CE, fp16 AMP, accuracy; CSV
         ``epoch,loss,accuracy,duration,gpus``.

What differs (by design): gradients are reduced by the fused NVLink all-reduce kernel from autograd hooks
(HyperionDDP), the loss/accuracy are accumulated on the device (one sync per epoch instead of ``loss.item()``
every step, distributed_utils.py:181, 249-250), data can be synthetic and device-resident, and the backend may
be gloo/CPU (BASELINE config 1).
"""
from __future__ import annotations

import os

import torch

from .. import ops
from ..data import DeviceResidentLoader, synth
from ..models import SimpleTransformerLM, resnet18, resnet50, vit_b_16
from ..parallel.runtime import cleanup, setup
from ..utils.checkpoint import load_training_state, save_model, save_training_state
from ..utils.metrics import CsvLogger
from .common import EpochTimer, StepRunner, load_cifar_tensors, load_wikitext_tensors, mean_over_ranks
from .engine import build_engine


def train_language_model_ddp(rank: int, world_size: int, epochs: int = 5, base_dir: str = ".", *, backend: str | None = None,
                             precision: str | None = None, synthetic: bool = True, steps_per_epoch: int | None = None,
                             batch_size: int = 32, rows: int | None = None, fused_comm: bool | None = None,
                             seed: int = 0, resume: str | None = None, save: bool = True, model_kwargs: dict | None = None,
                             vocab_size: int = synth.GPT2_VOCAB, pad_id: int = synth.GPT2_PAD, strategy: str = "ddp",
                             graph: bool | None = None):
    dev = setup(rank, world_size, backend, seed=seed)
    precision = precision or ("fp16" if dev.type == "cuda" else "fp32")
    out_dir = os.path.join(base_dir, "data/distributed")
    logger = CsvLogger(out_dir, "language_ddp", world_size, rank)
    ids, _mask = load_wikitext_tensors(base_dir, dev, synthetic, rows, vocab=vocab_size, pad_id=pad_id, seed=seed)
    loader = DeviceResidentLoader((ids,), batch_size, rank, world_size, shuffle=True, seed=seed)
    model = SimpleTransformerLM(vocab_size, **(model_kwargs or {}))
    eng = build_engine(model, strategy=strategy, precision=precision, lr=2e-4, device=dev, fused_comm=fused_comm)
    start_epoch = load_training_state(resume, eng.model, eng.optimizer, eng.scaler) if resume else 0
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
            print(f"[LM-DDP] Epoch {ep + 1}/{epochs} loss={avg:.4f} time={wall_s:.2f}s")
        logger.log(ep + 1, avg, wall_s, world_size)
        toks = nsteps * batch_size * world_size * (ids.shape[1] - 1)
        logger.log_device(ep + 1, nsteps, dev_ms / max(nsteps, 1), toks / (dev_ms / 1e3), "tokens/s",
                          torch.cuda.max_memory_allocated(dev) / 2 ** 20 if dev.type == "cuda" else 0.0, world_size)
    if save:
        save_model(eng.model, os.path.join(out_dir, f"{logger.run_id}_model.pt"), rank, optimizer=eng.optimizer)
        save_training_state(os.path.join(out_dir, f"{logger.run_id}_train_state.pt"), eng.model, eng.optimizer,
                            epochs - 1, eng.scaler, rank)
    cleanup()
    return history


VISION_MODELS = {
    "resnet18": (lambda: resnet18(num_classes=10), 32, 10),
    "resnet50": (lambda: resnet50(num_classes=1000), 224, 1000),
    "vit": (lambda: vit_b_16(num_classes=1000), 224, 1000),
}


def train_cifar_model_ddp(rank: int, world_size: int, epochs: int = 5, base_dir: str = ".", *, backend: str | None = None,
                          precision: str | None = None, synthetic: bool = True, steps_per_epoch: int | None = None,
                          batch_size: int = 64, rows: int | None = None, fused_comm: bool | None = None, seed: int = 0,
                          arch: str = "resnet18", checkpointing: bool = False, save: bool = True, resume: str | None = None,
                          graph: bool | None = None):
    """ResNet-18 on CIFAR-10-shaped data (the reference job); ``arch`` generalises it to ResNet-50 / ViT-B/16
    at 224x224 (BASELINE configs 2 and 3)."""
    dev = setup(rank, world_size, backend, seed=seed)
    precision = precision or ("fp16" if dev.type == "cuda" else "fp32")
    out_dir = os.path.join(base_dir, "data/distributed")
    logger = CsvLogger(out_dir, "cifar_ddp" if arch == "resnet18" else f"{arch}_ddp", world_size, rank)
    make, size, ncls = VISION_MODELS[arch]
    if size == 32:
        x_all, y_all = load_cifar_tensors(base_dir, dev, synthetic, rows, seed)
    else:
        n = rows or batch_size * world_size * (steps_per_epoch or 8)
        x_all, y_all = synth.image_batch(n, size, ncls, dev, seed)
    model = make()
    if checkpointing:
        from ..memory import enable_checkpointing

        enable_checkpointing(model, True)
    eng = build_engine(model, strategy="ddp", precision=precision, lr=1e-3, device=dev, fused_comm=fused_comm)
    x_all = x_all.to(eng.precision.dtype)
    loader = DeviceResidentLoader((x_all, y_all), batch_size, rank, world_size, shuffle=True, seed=seed)
    start_epoch = load_training_state(resume, eng.model, eng.optimizer, eng.scaler) if resume else 0
    eng.model.train()

    def loss_fn(b):
        out = eng.model(b[0])
        with torch.no_grad():
            correct = (out.argmax(1) == b[1]).sum()
        return ops.cross_entropy(out, b[1]), correct

    runner = StepRunner(eng, loss_fn, graph=graph)
    timer = EpochTimer(dev)
    history = []
    for ep in range(start_epoch, epochs):
        loader.set_epoch(ep)
        stats = torch.zeros(3, device=dev)  # loss sum, correct, total
        nsteps = 0
        timer.start()
        for batch in loader:
            loss, correct = runner(batch)
            stats[0] += loss.float()
            stats[1] += correct
            stats[2] += batch[1].numel()
            nsteps += 1
            if steps_per_epoch and nsteps >= steps_per_epoch:
                break
        wall_s, dev_ms = timer.stop()
        if world_size > 1:
            torch.distributed.all_reduce(stats)
        s = stats.tolist()
        avg_loss = s[0] / max(nsteps * world_size, 1)
        acc = 100.0 * s[1] / max(s[2], 1)
        history.append((avg_loss, acc))
        if rank == 0:
            print(f"[CIFAR-DDP] Epoch {ep + 1}/{epochs} loss={avg_loss:.4f} acc={acc:.2f}% time={wall_s:.2f}s")
        logger.log(ep + 1, avg_loss, acc, wall_s, world_size)
        logger.log_device(ep + 1, nsteps, dev_ms / max(nsteps, 1), nsteps * batch_size * world_size / (dev_ms / 1e3),
                          "samples/s", torch.cuda.max_memory_allocated(dev) / 2 ** 20 if dev.type == "cuda" else 0.0, world_size)
    if save:
        save_model(eng.model, os.path.join(out_dir, f"{logger.run_id}_model.pt"), rank, optimizer=eng.optimizer)
        save_training_state(os.path.join(out_dir, f"{logger.run_id}_train_state.pt"), eng.model, eng.optimizer,
                            epochs - 1, eng.scaler, rank)
    cleanup()
    return history
