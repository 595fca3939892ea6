"""Benchmark cases for ``bench.py``: one per BASELINE.json config.  A case owns the model + engine and exposes

    step(device_batch)       one full training step (forward, backward, communication, optimizer), no host sync
    step_e2e(host_batch)     the same step through the public API with the batch copied from pinned host memory
                             and the loss read back to the host (what a user's loop does)
"""
from __future__ import annotations

import torch

from ..data import synth
from ..train.engine import build_engine


class Case:
    metric = "tokens/s"
    unit = "tokens/s"
    dtype = "bf16"

    def __init__(self, dev, rank, world):
        self.dev, self.rank, self.world = dev, rank, world
        self._k = 0
        self.batches: list = []
        self.h2d_bytes = 0
        self.d2h_bytes = 4

    def device_batch(self):
        self._k = (self._k + 1) % len(self.batches)
        return self.batches[self._k]

    def host_batch(self, i: int):
        b = self.batches[i % len(self.batches)]
        hb = tuple(t.cpu().pin_memory() for t in b)
        self.h2d_bytes = sum(t.numel() * t.element_size() for t in hb)
        return hb

    def to_device(self, hb):
        return tuple(t.to(self.dev, non_blocking=True) for t in hb)

    def loss_fn(self, batch):
        raise NotImplementedError

    graphed = None

    def enable_graph(self) -> bool:
        """Capture the whole step in a CUDA graph (falls back to eager and says so if capture fails)."""
        from ..train.engine import GraphedStep

        try:
            self.graphed = GraphedStep(self.eng, self.loss_fn, self.batches[0])
            return True
        except Exception as e:  # noqa: BLE001
            import sys

            print(f"[bench] CUDA-graph capture failed, running eager: {e!r}", file=sys.stderr)
            torch.cuda.synchronize()
            self.graphed = None
            return False

    def step(self, batch):
        if self.graphed is not None:
            return self.graphed(batch)
        loss = self.loss_fn(batch)
        self.eng.backward_and_step(loss)
        return loss

    def step_e2e(self, host_batch) -> float:
        loss = self.step(self.to_device(host_batch))
        return float(loss.item())


class GPT2FSDP(Case):
    """BASELINE config 4: GPT-2 medium, FSDP full-shard state, bf16, seq 1024, WikiText-2-shaped synthetic tokens."""

    def __init__(self, dev, rank, world, batch=None, seq=None, size=None, fused_comm=None):
        super().__init__(dev, rank, world)
        from ..models import GPT2Config, GPT2LMHeadModel
        from ..models.gpt2 import Block

        size = size or "medium"
        cfg = {"medium": GPT2Config.medium, "small": GPT2Config.small, "tiny": GPT2Config.tiny}[size]()
        self.seq = min(seq or 1024, cfg.n_positions)
        self.bs = batch or 8
        model = GPT2LMHeadModel(cfg)
        units = [m for m in model.modules() if isinstance(m, Block)]
        self.eng = build_engine(model, strategy="fsdp", precision="bf16", lr=1e-4, clip=1.0, device=dev,
                                fused_comm=fused_comm, fsdp_units=units)
        self.items_per_step = self.bs * self.seq
        for k in range(4):
            ids, _ = synth.token_batch(self.bs, self.seq, cfg.vocab_size, dev, seed=1000 * rank + k)
            self.batches.append((ids,))
        nparam = sum(p.numel() for p in model.parameters())
        self.metric = "GPT-2 medium FSDP tokens/s" if size == "medium" else f"GPT-2 {size} FSDP tokens/s"
        self.config = {"model": f"gpt2-{size}", "params": nparam, "global_batch": self.bs * world, "seq_len": self.seq,
                       "parallelism": f"fsdp{world}" if world > 1 else "fsdp1(no_shard)", "optimizer": "AdamW lr=1e-4 clip=1.0",
                       "comm": "fused NVLink kernels" if getattr(self.eng.model, "fused", False) else ("none" if world == 1 else "nccl"),
                       "l2": "working set (710 MB bf16 weights + activations per step) exceeds the 126 MB L2; no explicit flush"}

    def loss_fn(self, batch):
        return self.eng.model.loss(batch[0])


class LMDDP(Case):
    """The reference's own DDP job: SimpleTransformerLM, bs 32/GPU, seq 127, fp16 AMP (distributed_utils.py:132-200)."""
    dtype = "fp16"

    def __init__(self, dev, rank, world, batch=None, seq=None, size=None, fused_comm=None):
        super().__init__(dev, rank, world)
        from ..models import SimpleTransformerLM

        self.bs, self.seq = batch or 32, seq or 128
        model = SimpleTransformerLM(synth.GPT2_VOCAB)
        self.eng = build_engine(model, strategy="ddp", precision="fp16", lr=2e-4, device=dev, fused_comm=fused_comm)
        self.items_per_step = self.bs * (self.seq - 1)
        for k in range(4):
            ids, _ = synth.wikitext2_like_arrays(self.bs, self.seq, seed=1000 * rank + k)
            self.batches.append((torch.from_numpy(ids).to(dev),))
        self.metric = "SimpleTransformerLM DDP tokens/s"
        self.config = {"model": "SimpleTransformerLM(50257,256,4,2)", "global_batch": self.bs * world, "seq_len": self.seq - 1,
                       "parallelism": f"dp{world}", "l2": "fp16 weights 57 MB + 817 MB-equivalent logits chunks exceed L2"}

    def loss_fn(self, batch):
        ids = batch[0]
        return self.eng.model.loss(ids[:, :-1].contiguous(), ids[:, 1:].contiguous(), ignore_index=synth.GPT2_PAD)


class LMFSDP(LMDDP):
    """The reference's own FSDP job: SimpleTransformerLM, FULL_SHARD, size-based wrap (100 k), bf16, AdamW 1e-4, clip 1.0,
    bs 32/GPU (distributed_utils.py:290-352)."""
    dtype = "bf16"

    def __init__(self, dev, rank, world, batch=None, seq=None, size=None, fused_comm=None):
        Case.__init__(self, dev, rank, world)
        from ..models import SimpleTransformerLM

        self.bs, self.seq = batch or 32, seq or 128
        model = SimpleTransformerLM(synth.GPT2_VOCAB)
        self.eng = build_engine(model, strategy="fsdp", precision="bf16", lr=1e-4, clip=1.0, device=dev, fused_comm=fused_comm,
                                fsdp_min_num_params=100_000)
        self.items_per_step = self.bs * (self.seq - 1)
        for k in range(4):
            ids, _ = synth.wikitext2_like_arrays(self.bs, self.seq, seed=1000 * rank + k)
            self.batches.append((torch.from_numpy(ids).to(dev),))
        self.metric = "SimpleTransformerLM FSDP tokens/s"
        self.config = {"model": "SimpleTransformerLM(50257,256,4,2)", "global_batch": self.bs * world, "seq_len": self.seq - 1,
                       "parallelism": f"fsdp{world}", "l2": "bf16 weights 57 MB + logits chunks exceed L2"}


class CifarDDP(Case):
    """The reference's own vision job: ResNet-18 (10 classes) on CIFAR-shaped 32x32 images, DDP, fp16 AMP, AdamW 1e-3,
    bs 64/GPU (distributed_utils.py:207-279)."""
    unit = "samples/s"
    dtype = "fp16"

    def __init__(self, dev, rank, world, batch=None, fused_comm=None):
        super().__init__(dev, rank, world)
        from .. import ops
        from ..models import resnet18

        self.ops = ops
        self.bs = batch or 64
        model = resnet18(num_classes=10)
        self.eng = build_engine(model, strategy="ddp", precision="fp16", lr=1e-3, device=dev, fused_comm=fused_comm)
        self.items_per_step = self.bs
        for k in range(4):
            x, y = synth.image_batch(self.bs, 32, 10, dev, seed=1000 * rank + k, dtype=torch.float32)
            self.batches.append((x, y))
        self.metric = "ResNet-18 CIFAR DDP samples/s"
        self.config = {"model": "resnet18(num_classes=10)", "global_batch": self.bs * world, "image": 32,
                       "parallelism": f"dp{world}", "input": "fp32 images, cast to fp16 on the device inside the step",
                       "l2": "small working set; 256 MB L2 flush not applied (launch-bound job)"}

    def loss_fn(self, batch):
        x, y = batch
        return self.ops.cross_entropy(self.eng.model(x.to(torch.float16)), y)


class VisionDDP(Case):
    """BASELINE configs 2 / 3: ResNet-50 DDP bf16, ViT-B/16 DDP bf16 with activation checkpointing (224x224)."""
    unit = "samples/s"

    def __init__(self, dev, rank, world, arch="resnet50", batch=None, fused_comm=None):
        super().__init__(dev, rank, world)
        from .. import ops
        from ..memory import enable_checkpointing
        from ..models import resnet50, vit_b_16

        self.ops = ops
        self.bs = batch or 64
        model = resnet50() if arch == "resnet50" else vit_b_16()
        if arch == "vit":
            enable_checkpointing(model, True)
        self.eng = build_engine(model, strategy="ddp", precision="bf16", lr=1e-3, device=dev, fused_comm=fused_comm)
        self.items_per_step = self.bs
        for k in range(2):
            x, y = synth.image_batch(self.bs, 224, 1000, dev, seed=1000 * rank + k, dtype=torch.float32)
            self.batches.append((x, y))
        self.metric = f"{'ResNet-50' if arch == 'resnet50' else 'ViT-B/16'} DDP samples/s"
        self.config = {"model": arch, "global_batch": self.bs * world, "image": 224, "parallelism": f"dp{world}",
                       "checkpointing": arch == "vit", "input": "fp32 images (as the torch arm), cast to bf16 on the device inside the step",
                       "l2": "activations per step exceed the 126 MB L2"}

    def loss_fn(self, batch):
        x, y = batch
        return self.ops.cross_entropy(self.eng.model(x.to(torch.bfloat16)), y)


class LlamaLoRA(Case):
    """BASELINE config 5: Llama-2-7B LoRA (r=16, q/k/v/o) under FSDP, bf16, seq 128."""

    def __init__(self, dev, rank, world, batch=None, seq=None, size=None, fused_comm=None, lora=True):
        super().__init__(dev, rank, world)
        from ..lora import inject_lora
        from ..models import LlamaConfig, LlamaForCausalLM
        from ..models.llama import LlamaDecoderLayer

        size = size or "7b"
        cfg = LlamaConfig.llama2_7b() if size == "7b" else LlamaConfig.tiny()
        self.bs, self.seq = batch or 4, seq or 128
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            with torch.device(dev):
                model = LlamaForCausalLM(cfg)
        finally:
            torch.set_default_dtype(prev)
        if lora:
            inject_lora(model, r=16, alpha=32, dropout=0.05)
        units = [m for m in model.modules() if isinstance(m, LlamaDecoderLayer)]
        self.eng = build_engine(model, strategy="fsdp", precision="bf16", lr=1e-5, weight_decay=0.01, clip=1.0, device=dev,
                                fused_comm=fused_comm, fsdp_units=units)
        self.items_per_step = self.bs * self.seq
        for k in range(4):
            ids, mask = synth.token_batch(self.bs, self.seq, cfg.vocab_size, dev, seed=1000 * rank + k, pad_id=cfg.pad_token_id,
                                          pad_fraction=0.3)
            self.batches.append((ids, mask))
        self.metric = f"Llama-2 {size} LoRA FSDP tokens/s" if lora else f"Llama-2 {size} FSDP tokens/s"
        self.config = {"model": f"llama2-{size}", "lora": "r16 a32 qkvo" if lora else None, "global_batch": self.bs * world, "seq_len": self.seq,
                       "parallelism": f"fsdp{world}", "l2": "13.5 GB bf16 weights exceed L2"}

    def loss_fn(self, batch):
        ids, mask = batch
        return self.eng.model(input_ids=ids, attention_mask=mask, labels=ids.clone()).loss


def make_case(name: str, dev, rank, world, batch=None, seq=None, size=None, fused_comm=None) -> Case:
    if name == "gpt2_fsdp":
        return GPT2FSDP(dev, rank, world, batch, seq, size, fused_comm)
    if name == "lm_ddp":
        return LMDDP(dev, rank, world, batch, seq, size, fused_comm)
    if name == "lm_fsdp":
        return LMFSDP(dev, rank, world, batch, seq, size, fused_comm)
    if name == "cifar_ddp":
        return CifarDDP(dev, rank, world, batch, fused_comm)
    if name == "resnet50_ddp":
        return VisionDDP(dev, rank, world, "resnet50", batch, fused_comm)
    if name == "vit_ddp":
        return VisionDDP(dev, rank, world, "vit", batch, fused_comm)
    if name == "llama_lora":
        return LlamaLoRA(dev, rank, world, batch, seq, size, fused_comm)
    if name == "llama_fsdp":
        return LlamaLoRA(dev, rank, world, batch, seq, size, fused_comm, lora=False)
    raise ValueError(name)
