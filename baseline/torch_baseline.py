"""Stock-PyTorch baseline harness: the reference-equivalent loop -- same model, hyper-parameters and
wrapper configuration as the reference's trainers -- on stock torch DDP / FSDP1 + NCCL + cuBLAS/cuDNN/SDPA, timed exactly
like the product (CUDA events, K steps after W warm-ups, max over ranks).  None of the framework's kernels, models or
engines are used here.  ``python bench.py --impl torch --config ...`` runs it; it prints the same JSON line with
``"impl": "torch"``.

* gpt2_fsdp   : transformers ``GPT2LMHeadModel(GPT2Config(n_embd=1024, n_layer=24, n_head=16))`` (random init), torch
                ``FSDP(FULL_SHARD, transformer_auto_wrap_policy{GPT2Block}, MixedPrecision(bf16 x3))``, AdamW 1e-4,
                ``clip_grad_norm_(1.0)`` -- the recipe of distributed_utils.py:318-352 applied to BASELINE config 4.
* lm_ddp      : ``nn.TransformerEncoder`` LM exactly as distributed_utils.py:75-88 + DDP + fp16 autocast + GradScaler.
* lm_fsdp     : the same LM under torch ``FSDP(FULL_SHARD, size_based_auto_wrap_policy(100 k), MixedPrecision(bf16 x3))``,
                AdamW 1e-4, ``clip_grad_norm_(1.0)`` (distributed_utils.py:290-352).
* resnet50_ddp: torchvision resnet50 + DDP + bf16 autocast, channels_last.
* vit_ddp     : torchvision vit_b_16 + DDP + bf16 autocast, ``torch.utils.checkpoint`` around every encoder block (the
                product arm checkpoints per block too).
* cifar_ddp   : torchvision resnet18(num_classes=10) + DDP + fp16 autocast + GradScaler, 32x32 (distributed_utils.py:207-279).
* llama_lora  : transformers ``LlamaForCausalLM(LlamaConfig())`` (Llama-2-7B shape, random init, bf16) with a hand-rolled
                LoRA (r 16, alpha 32, dropout 0.05 on q/k/v/o -- the reference's LoraConfig, distributed_utils.py:469-473;
                peft itself is not installed) under torch FSDP FULL_SHARD with one unit per decoder layer
                (``use_orig_params=True`` so frozen and trainable parameters can share a unit), AdamW 1e-5 wd 0.01, clip 1.0.
* llama_fsdp  : the same model, all parameters trainable, FSDP FULL_SHARD per decoder layer (distributed_utils.py:477-500).

Every JSON line carries a ``clocks`` record sampled during the device-timed region, like the product arm.
"""
from __future__ import annotations

import functools
import json
import os
import sys

import torch
import torch.distributed as dist
import torch.nn as nn


def _setup():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1 or True:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.manual_seed(0)
    return rank, world, dev


class _RefLM(nn.Module):
    def __init__(self, vocab_size, emb_dim=256, n_heads=4, n_layers=2):
        super().__init__()
        self.embed = nn.Embedding(vocab_size, emb_dim)
        layer = nn.TransformerEncoderLayer(d_model=emb_dim, nhead=n_heads)
        self.tr = nn.TransformerEncoder(layer, num_layers=n_layers)
        self.fc = nn.Linear(emb_dim, vocab_size)

    def forward(self, x, **_):
        x = self.embed(x).permute(1, 0, 2)
        return self.fc(self.tr(x).permute(1, 0, 2))


def run(args) -> int:
    from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
    from torch.distributed.fsdp import MixedPrecision, ShardingStrategy
    from torch.distributed.fsdp.wrap import transformer_auto_wrap_policy
    from torch.nn.parallel import DistributedDataParallel as DDP

    rank, world, dev = _setup()
    cfgname = args.config
    scaler = None
    if cfgname == "gpt2_fsdp":
        from transformers import GPT2Config, GPT2LMHeadModel
        from transformers.models.gpt2.modeling_gpt2 import GPT2Block

        size = args.size or "medium"
        dims = {"medium": (1024, 24, 16), "small": (768, 12, 12)}[size]
        seq, bs = args.seq or 1024, args.batch or 8
        hf = GPT2Config(n_embd=dims[0], n_layer=dims[1], n_head=dims[2], n_positions=1024, resid_pdrop=0.0, embd_pdrop=0.0,
                        attn_pdrop=0.0, attn_implementation="sdpa")
        model = GPT2LMHeadModel(hf).to(dev)
        mp = MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16)
        model = FSDP(model, sharding_strategy=ShardingStrategy.FULL_SHARD, mixed_precision=mp, device_id=dev,
                     auto_wrap_policy=functools.partial(transformer_auto_wrap_policy, transformer_layer_cls={GPT2Block}),
                     use_orig_params=False)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4)
        batches = [torch.randint(0, 50257, (bs, seq), device=dev) for _ in range(4)]
        items, unit, metric, dtype = bs * seq, "tokens/s", f"GPT-2 {size} FSDP tokens/s", "bf16"

        def step(b):
            opt.zero_grad(set_to_none=True)
            loss = model(input_ids=b, labels=b).loss
            loss.backward()
            model.clip_grad_norm_(1.0)
            opt.step()
            return loss
        config = {"model": f"gpt2-{size}", "global_batch": bs * world, "seq_len": seq, "parallelism": f"fsdp{world}"}
    elif cfgname == "lm_fsdp":
        from torch.distributed.fsdp.wrap import size_based_auto_wrap_policy

        bs, seq = args.batch or 32, args.seq or 128
        mp = MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16)
        model = FSDP(_RefLM(50257).to(dev), sharding_strategy=ShardingStrategy.FULL_SHARD, mixed_precision=mp, device_id=dev,
                     auto_wrap_policy=functools.partial(size_based_auto_wrap_policy, min_num_params=100_000))
        opt = torch.optim.AdamW(model.parameters(), lr=1e-4)
        crit = nn.CrossEntropyLoss(ignore_index=50256)
        batches = [torch.randint(0, 50256, (bs, seq), device=dev) for _ in range(4)]
        items, unit, metric, dtype = bs * (seq - 1), "tokens/s", "SimpleTransformerLM FSDP tokens/s", "bf16"

        def step(b):
            x, y = b[:, :-1], b[:, 1:]
            opt.zero_grad(set_to_none=True)
            logits = model(x)
            loss = crit(logits.reshape(-1, logits.size(-1)), y.reshape(-1))
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.parameters(), 1.0)   # as the reference does (local norm under FSDP)
            opt.step()
            return loss
        config = {"model": "SimpleTransformerLM(50257,256,4,2)", "global_batch": bs * world, "seq_len": seq - 1, "parallelism": f"fsdp{world}"}
    elif cfgname in ("llama_lora", "llama_fsdp"):
        import math

        from transformers import LlamaConfig, LlamaForCausalLM
        from transformers.models.llama.modeling_llama import LlamaDecoderLayer

        size = args.size or "7b"
        bs, seq = args.batch or 4, args.seq or 128
        hf = (LlamaConfig(attn_implementation="sdpa") if size == "7b" else
              LlamaConfig(hidden_size=256, intermediate_size=688, num_hidden_layers=2, num_attention_heads=4,
                          num_key_value_heads=4, vocab_size=32000, attn_implementation="sdpa"))
        prev = torch.get_default_dtype()
        torch.set_default_dtype(torch.bfloat16)
        try:
            with torch.device(dev):
                model = LlamaForCausalLM(hf)
        finally:
            torch.set_default_dtype(prev)
        lora = cfgname == "llama_lora"
        if lora:
            class LoRALinear(nn.Module):
                def __init__(self, base: nn.Linear, r=16, alpha=32, p=0.05):
                    super().__init__()
                    self.base, self.scale = base, alpha / r
                    self.lora_A = nn.Linear(base.in_features, r, bias=False, device=base.weight.device, dtype=base.weight.dtype)
                    self.lora_B = nn.Linear(r, base.out_features, bias=False, device=base.weight.device, dtype=base.weight.dtype)
                    nn.init.kaiming_uniform_(self.lora_A.weight, a=math.sqrt(5))
                    nn.init.zeros_(self.lora_B.weight)
                    self.drop = nn.Dropout(p)

                def forward(self, x):
                    return self.base(x) + self.lora_B(self.lora_A(self.drop(x))) * self.scale

            for p_ in model.parameters():
                p_.requires_grad_(False)
            for layer in model.model.layers:
                for nm in ("q_proj", "k_proj", "v_proj", "o_proj"):
                    setattr(layer.self_attn, nm, LoRALinear(getattr(layer.self_attn, nm)))
        mp = MixedPrecision(param_dtype=torch.bfloat16, reduce_dtype=torch.bfloat16, buffer_dtype=torch.bfloat16)
        model = FSDP(model, sharding_strategy=ShardingStrategy.FULL_SHARD, mixed_precision=mp, device_id=dev,
                     auto_wrap_policy=functools.partial(transformer_auto_wrap_policy, transformer_layer_cls={LlamaDecoderLayer}),
                     use_orig_params=True)
        opt = torch.optim.AdamW([p_ for p_ in model.parameters() if p_.requires_grad], lr=1e-5, weight_decay=0.01)
        g = torch.Generator().manual_seed(rank)
        batches = []
        for _ in range(4):
            ids = torch.randint(0, hf.vocab_size - 1, (bs, seq), generator=g)
            lens = torch.randint(int(seq * 0.7), seq + 1, (bs,), generator=g)
            mask = (torch.arange(seq)[None] < lens[:, None]).long()
            batches.append((torch.where(mask.bool(), ids, torch.full_like(ids, 2)).to(dev), mask.to(dev)))
        items, unit, dtype = bs * seq, "tokens/s", "bf16"
        metric = f"Llama-2 {size} LoRA FSDP tokens/s" if lora else f"Llama-2 {size} FSDP tokens/s"

        def step(b):
            ids, mask = b
            opt.zero_grad(set_to_none=True)
            loss = model(input_ids=ids, attention_mask=mask, labels=ids.clone()).loss
            loss.backward()
            model.clip_grad_norm_(1.0)
            opt.step()
            return loss
        config = {"model": f"llama2-{size}", "global_batch": bs * world, "seq_len": seq, "parallelism": f"fsdp{world}",
                  "lora": "r16 a32 qkvo, hand-rolled (peft absent)" if lora else None}
    elif cfgname == "lm_ddp":
        bs, seq = args.batch or 32, args.seq or 128
        model = DDP(_RefLM(50257).to(dev), device_ids=[dev.index])
        opt = torch.optim.AdamW(model.parameters(), lr=2e-4)
        crit = nn.CrossEntropyLoss(ignore_index=50256)
        scaler = torch.amp.GradScaler("cuda")
        batches = [torch.randint(0, 50256, (bs, seq), device=dev) for _ in range(4)]
        items, unit, metric, dtype = bs * (seq - 1), "tokens/s", "SimpleTransformerLM DDP tokens/s", "fp16"

        def step(b):
            x, y = b[:, :-1], b[:, 1:]
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.float16):
                logits = model(x)
                loss = crit(logits.reshape(-1, logits.size(-1)), y.reshape(-1))
            scaler.scale(loss).backward()
            scaler.step(opt)
            scaler.update()
            return loss
        config = {"model": "SimpleTransformerLM(50257,256,4,2)", "global_batch": bs * world, "seq_len": seq - 1, "parallelism": f"dp{world}"}
    elif cfgname == "cifar_ddp":
        import torchvision

        bs = args.batch or 64
        model = DDP(torchvision.models.resnet18(num_classes=10).to(dev), device_ids=[dev.index])
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
        crit = nn.CrossEntropyLoss()
        scaler = torch.amp.GradScaler("cuda")
        batches = [(torch.randn(bs, 3, 32, 32, device=dev), torch.randint(0, 10, (bs,), device=dev)) for _ in range(4)]
        items, unit, metric, dtype = bs, "samples/s", "ResNet-18 CIFAR DDP samples/s", "fp16"

        def step(b):
            x, y = b
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.float16):
                loss = crit(model(x), y)
            scaler.scale(loss).backward()
            scaler.step(opt)
            scaler.update()
            return loss
        config = {"model": "resnet18(num_classes=10)", "global_batch": bs * world, "image": 32, "parallelism": f"dp{world}"}
    elif cfgname in ("resnet50_ddp", "vit_ddp"):
        import torchvision

        bs = args.batch or 64
        net = torchvision.models.resnet50() if cfgname == "resnet50_ddp" else torchvision.models.vit_b_16()
        if cfgname == "vit_ddp":
            from torch.utils.checkpoint import checkpoint

            class _Ckpt(nn.Module):   # activation checkpointing per encoder block, as the product arm runs it
                def __init__(self, blk):
                    super().__init__()
                    self.blk = blk

                def forward(self, x):
                    return checkpoint(self.blk, x, use_reentrant=False) if self.training and torch.is_grad_enabled() else self.blk(x)
            net.encoder.layers = nn.Sequential(*[_Ckpt(b) for b in net.encoder.layers])
        model = DDP(net.to(dev).to(memory_format=torch.channels_last), device_ids=[dev.index])
        opt = torch.optim.AdamW(model.parameters(), lr=1e-3)
        crit = nn.CrossEntropyLoss()
        batches = [(torch.randn(bs, 3, 224, 224, device=dev).contiguous(memory_format=torch.channels_last),
                    torch.randint(0, 1000, (bs,), device=dev)) for _ in range(2)]
        items, unit, dtype = bs, "samples/s", "bf16"
        metric = "ResNet-50 DDP samples/s" if cfgname == "resnet50_ddp" else "ViT-B/16 DDP samples/s"

        def step(b):
            x, y = b
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = crit(model(x), y)
            loss.backward()
            opt.step()
            return loss
        config = {"model": cfgname.split("_")[0], "global_batch": bs * world, "image": 224, "parallelism": f"dp{world}",
                  "checkpointing": cfgname == "vit_ddp", "input": "fp32 images"}
    else:
        if rank == 0:
            print(json.dumps({"impl": "torch", "unavailable": f"no torch baseline for {cfgname}"}))
        dist.destroy_process_group()
        return 0

    def sync_all():
        torch.cuda.synchronize(dev)
        dist.barrier()
        torch.cuda.synchronize(dev)

    def maxr(v):
        t = torch.tensor([v], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from hyperion_b200.utils.timing import ClockSampler   # measurement only: nvidia-smi sampling thread

    for i in range(args.warmup):
        step(batches[i % len(batches)])
    sync_all()
    sampler = ClockSampler(dev.index or 0).start() if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(batches[i % len(batches)])
    e1.record()
    sync_all()
    ms = maxr(e0.elapsed_time(e1)) / args.steps
    clocks = sampler.stop() if sampler is not None else None
    # end to end: pinned host batch -> device, loss.item() each step (what the reference's loops do)
    to_host = lambda b: tuple(t.cpu().pin_memory() for t in (b if isinstance(b, tuple) else (b,)))
    hbs = [to_host(b) for b in batches]
    h2d = sum(t.numel() * t.element_size() for t in hbs[0])

    def e2e_step(hb):
        db = tuple(t.to(dev, non_blocking=True) for t in hb)
        return float(step(db if len(db) > 1 else db[0]).item())
    for i in range(2):
        e2e_step(hbs[i % len(hbs)])
    sync_all()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for i in range(args.steps):
        e2e_step(hbs[i % len(hbs)])
    f1.record()
    sync_all()
    ms2 = maxr(f0.elapsed_time(f1)) / args.steps
    if rank == 0:
        print(json.dumps({"metric": metric, "value": items * world / (ms / 1e3), "unit": unit, "n_gpus": world, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                          "vs_baseline": None, "dtype": dtype, "data": "synthetic", "impl": "torch", "config": config, "clocks": clocks,
                          "e2e": {"value": items * world / (ms2 / 1e3), "unit": unit, "ms_per_step": ms2,
                                  "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
                          "gpu_launches": 0, "peak_mem_gb": torch.cuda.max_memory_allocated(dev) / 2 ** 30}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0
