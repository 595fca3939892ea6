"""``train_llama_fsdp``: Llama-2 fine-tuning, full FSDP or LoRA (reference: 02_development/distributed_utils.py:415-554).

Reference recipe kept: bf16 weights, AdamW(lr 1e-5, wd 0.01), batch 1/GPU, seq 128, ``labels = input_ids.clone()``
(pad positions are NOT masked to -100, distributed_utils.py:517), HF-style shifted CE inside the model, clip 1.0, CSV
``epoch,loss,duration_s,gpus,mode`` with integer seconds and mode ``lora_bf16`` / ``fsdp_bf16``; save: FSDP ->
``{run_id}_fsdp_bf16.pt`` full state dict, LoRA -> peft ``save_pretrained`` directory.

Differences: the architecture is built locally from ``LlamaConfig`` with random-init weights (no network);
synthetic ids are drawn from ``[0, vocab)`` with pad = eos = 2 (the reference feeds GPT-2 ids up to 50256 into a
32000-row embedding, A-5); ``--lora`` can run under DDP (as the reference) **or under FSDP** (BASELINE config 5);
the adapter export unwraps the parallel wrapper (A-6).
"""
from __future__ import annotations

import os

import torch

from ..data import DeviceResidentLoader, synth
from ..lora import inject_lora, save_pretrained, trainable_parameters
from ..models import LlamaConfig, LlamaForCausalLM
from ..models.llama import LlamaDecoderLayer
from ..parallel.runtime import cleanup, setup
from ..utils.checkpoint import load_training_state, save_model, save_training_state
from ..utils.metrics import CsvLogger
from .common import EpochTimer, StepRunner, mean_over_ranks
from .engine import build_engine

CONFIGS = {"7b": LlamaConfig.llama2_7b, "tiny": LlamaConfig.tiny}


def train_llama_fsdp(rank: int, world_size: int, epochs: int = 1, base_dir: str = ".", hf_token: str | None = None,
                     model_id: str = "NousResearch/Llama-2-7b-hf", lora: bool = False, batch_size: int = 1,
                     progress_every: int = 50, *, backend: str | None = None, size: str = "7b", seq_len: int = 128,
                     steps_per_epoch: int | None = 50, lora_strategy: str = "fsdp", fused_comm: bool | None = None,
                     seed: int = 0, save: bool = True, checkpointing: bool = False, config: LlamaConfig | None = None,
                     resume: str | None = None, graph: bool | None = None, reshard_after_forward: bool | None = None):
    dev = setup(rank, world_size, backend, seed=seed)
    from ..models.hf import is_checkpoint_dir, llama_config_from_hf, read_config

    pretrained = config is None and is_checkpoint_dir(model_id)      # a local HF checkpoint directory: its weights, its config
    cfg = config or (llama_config_from_hf(read_config(model_id)) if pretrained else CONFIGS[size]())
    if dev.type != "cuda" and config is None and not pretrained and size == "7b" and os.environ.get("HYPERION_ALLOW_CPU_7B") != "1":
        # fp32 weights + gradients + AdamW state of 6.7 B parameters per rank: > 100 GB of host memory, and the OOM killer's choice of
        # victim is not ours to make.  The CPU / gloo path exists for the host-side logic; it has a model of its own size.
        raise RuntimeError("Llama-2-7B on a machine without a GPU needs > 100 GB of host memory per rank; pass size='tiny' "
                           "(--llama-size tiny) for CPU runs, or set HYPERION_ALLOW_CPU_7B=1 to go ahead")
    mode = "lora_bf16" if lora else "fsdp_bf16"
    out_dir = os.path.join(base_dir, "data/distributed")
    logger = CsvLogger(out_dir, "llama", world_size, rank)
    precision = "bf16" if dev.type == "cuda" else "fp32"
    dtype = torch.bfloat16 if dev.type == "cuda" else torch.float32
    # build directly in the compute dtype on the device (7B fp32 on the host would be 27 GB per rank)
    prev = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        with torch.device(dev):
            model = LlamaForCausalLM(cfg)
    finally:
        torch.set_default_dtype(prev)
    if pretrained:
        from ..models.hf import _copy_in, iter_tensors

        _copy_in(model, iter_tensors(model_id), lambda k, v: None if k.endswith("rotary_emb.inv_freq") else (k, v))
        if rank == 0:
            print(f"[Llama] weights read from {model_id} ({model.num_parameters():,} parameters)")
    elif rank == 0 and model_id and os.path.sep in str(model_id) and not str(model_id).startswith("NousResearch/"):
        print(f"[Llama] {model_id} is not a local checkpoint directory (no config.json): training from random weights")
    model.model.use_checkpoint = checkpointing
    if lora:
        inject_lora(model, r=16, alpha=32, dropout=0.05)
        if rank == 0:
            print(f"[Llama] LoRA adapters: {trainable_parameters(model):,} trainable parameters")
    units = [m for m in model.modules() if isinstance(m, LlamaDecoderLayer)]
    strategy = ("fsdp" if lora_strategy == "fsdp" else "ddp") if lora else "fsdp"
    eng = build_engine(model, strategy=strategy, precision=precision, lr=1e-5, weight_decay=0.01, clip=1.0, device=dev,
                       fused_comm=fused_comm, fsdp_units=units, reshard_after_forward=reshard_after_forward)
    start_epoch = load_training_state(resume, eng.model, eng.optimizer, eng.scaler) if resume else 0
    eng.model.train()
    runner = StepRunner(eng, lambda b: eng.model(input_ids=b[0], attention_mask=b[1], labels=b[0].clone()).loss, graph=graph)
    steps = steps_per_epoch or (synth.WIKITEXT2_ROWS["train"] // (batch_size * world_size))
    n = batch_size * world_size * steps
    ids, mask = synth.token_batch(n, seq_len, cfg.vocab_size, dev, seed, pad_id=cfg.pad_token_id, pad_fraction=0.4)
    loader = DeviceResidentLoader((ids, mask), batch_size, rank, world_size, shuffle=True, seed=seed)
    timer = EpochTimer(dev)
    history = []
    for ep in range(start_epoch, epochs):
        loader.set_epoch(ep)
        total = torch.zeros((), device=dev)
        nsteps = 0
        timer.start()
        for step, batch in enumerate(loader):
            loss = runner(batch)
            total += loss.float()
            nsteps += 1
            if rank == 0 and progress_every and (step + 1) % progress_every == 0:
                print(f"[Llama] epoch {ep + 1} step {step + 1}/{len(loader)}", flush=True)
        wall_s, dev_ms = timer.stop()
        avg = mean_over_ranks(total / max(nsteps, 1))
        history.append(avg)
        toks = nsteps * batch_size * world_size * seq_len
        if rank == 0:
            print(f"[Llama-{mode}] Epoch {ep + 1}/{epochs} loss={avg:.4f} time={wall_s:.1f}s tokens/s={toks / (dev_ms / 1e3):.0f}")
        logger.log(ep + 1, avg, int(wall_s), world_size, mode)
        logger.log_device(ep + 1, nsteps, dev_ms / max(nsteps, 1), toks / (dev_ms / 1e3), "tokens/s",
                          torch.cuda.max_memory_allocated(dev) / 2 ** 20 if dev.type == "cuda" else 0.0, world_size)
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)  # pending all-gathers on the comm stream must land before tensors are read
    if save:
        if lora:
            if rank == 0:
                save_pretrained(eng.raw_model, os.path.join(out_dir, f"{logger.run_id}_{mode}"), model_id)
        else:
            save_model(eng.model, os.path.join(out_dir, f"{logger.run_id}_{mode}.pt"), rank, dtype=None)
        save_training_state(os.path.join(out_dir, f"{logger.run_id}_train_state.pt"), eng.model, eng.optimizer, epochs - 1,
                            eng.scaler, rank)
    cleanup()
    return history
