"""GPT-2 (BASELINE config 4: GPT-2 medium -- 24 layers, d 1024, 16 heads, ctx 1024,
vocab 50257, ~354.8 M parameters).  Not a model of the reference itself (it only uses
the GPT-2 *tokenizer*, 02_development/distributed_utils.py:154); added because
BASELINE.json names it.  Parameter names follow HF GPT-2 (``wte wpe h.N.ln_1
h.N.attn.c_attn ...``) but weights use the ``nn.Linear`` ``[out, in]`` layout.

Pre-norm blocks; every residual add is fused into the epilogue of the GEMM that
produces the branch output; GELU(tanh) and its derivative live in GEMM epilogues;
the tied LM head + cross-entropy run as one op that produces the loss and all gradients in the forward
(16-bit logits are held while it runs, see ops.functional._LinearCEFn).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

from .. import ops


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 1024
    n_layer: int = 24
    n_head: int = 16
    layer_norm_eps: float = 1e-5
    dropout: float = 0.0

    @staticmethod
    def medium() -> "GPT2Config":
        return GPT2Config()

    @staticmethod
    def small() -> "GPT2Config":
        return GPT2Config(n_embd=768, n_layer=12, n_head=12)

    @staticmethod
    def tiny() -> "GPT2Config":
        return GPT2Config(vocab_size=512, n_positions=64, n_embd=64, n_layer=2, n_head=2)


class _Attn(nn.Module):
    def __init__(self, c: GPT2Config):
        super().__init__()
        self.n_head = c.n_head
        self.c_attn = nn.Linear(c.n_embd, 3 * c.n_embd)
        self.c_proj = nn.Linear(c.n_embd, c.n_embd)

    def forward(self, x, residual):
        B, S, E = x.shape
        qkv = ops.linear(x, self.c_attn.weight, self.c_attn.bias)
        o = ops.attention_qkvpacked(qkv.reshape(B, S, 3, self.n_head, E // self.n_head), causal=True)
        o = o.transpose(1, 2).reshape(B, S, E)
        return ops.linear(o, self.c_proj.weight, self.c_proj.bias, residual=residual)


class _MLP(nn.Module):
    def __init__(self, c: GPT2Config):
        super().__init__()
        self.c_fc = nn.Linear(c.n_embd, 4 * c.n_embd)
        self.c_proj = nn.Linear(4 * c.n_embd, c.n_embd)

    def forward(self, x, residual):
        return ops.mlp(x, self.c_fc.weight, self.c_fc.bias, self.c_proj.weight, self.c_proj.bias, act="gelu_tanh",
                       residual=residual)


class Block(nn.Module):
    def __init__(self, c: GPT2Config):
        super().__init__()
        self.ln_1 = nn.LayerNorm(c.n_embd, eps=c.layer_norm_eps)
        self.attn = _Attn(c)
        self.ln_2 = nn.LayerNorm(c.n_embd, eps=c.layer_norm_eps)
        self.mlp = _MLP(c)

    def forward(self, x):
        # layer_norm_fork hands back x for the skip connection, so the backward sums both branches in the norm kernel
        h, skip = ops.layer_norm_fork(x, self.ln_1.weight, self.ln_1.bias, self.ln_1.eps)
        x = self.attn(h, residual=skip)
        h, skip = ops.layer_norm_fork(x, self.ln_2.weight, self.ln_2.bias, self.ln_2.eps)
        return self.mlp(h, residual=skip)


class GPT2LMHeadModel(nn.Module):
    def __init__(self, config: GPT2Config | None = None):
        super().__init__()
        c = config or GPT2Config()
        self.config = c
        self.wte = nn.Embedding(c.vocab_size, c.n_embd)
        self.wpe = nn.Embedding(c.n_positions, c.n_embd)
        self.h = nn.ModuleList([Block(c) for _ in range(c.n_layer)])
        self.ln_f = nn.LayerNorm(c.n_embd, eps=c.layer_norm_eps)
        self.use_checkpoint = False
        self.apply(self._init)
        for name, p in self.named_parameters():
            if name.endswith("c_proj.weight"):
                nn.init.normal_(p, mean=0.0, std=0.02 / (2 * c.n_layer) ** 0.5)

    @staticmethod
    def _init(m):
        if isinstance(m, (nn.Linear, nn.Embedding)):
            nn.init.normal_(m.weight, mean=0.0, std=0.02)
            if isinstance(m, nn.Linear) and m.bias is not None:
                nn.init.zeros_(m.bias)

    def features(self, input_ids):
        B, S = input_ids.shape
        pos = torch.arange(S, device=input_ids.device)
        x = ops.embedding(input_ids, self.wte.weight) + ops.embedding(pos, self.wpe.weight)[None]
        if self.use_checkpoint and self.training and torch.is_grad_enabled():
            from ..memory import checkpoint_layers

            x = checkpoint_layers(self.h, x)
        else:
            for blk in self.h:
                x = blk(x)
        return ops.layer_norm(x, self.ln_f.weight, self.ln_f.bias, self.ln_f.eps)

    def forward(self, input_ids, **_):
        return ops.linear(self.features(input_ids), self.wte.weight)

    def loss(self, input_ids, labels=None, ignore_index: int = -100):
        """Shifted next-token loss (HF convention: labels = input_ids, shift inside)."""
        labels = input_ids if labels is None else labels
        h = self.features(input_ids)[:, :-1].contiguous()
        return ops.linear_cross_entropy(h, self.wte.weight, None, labels[:, 1:].contiguous(), ignore_index)
