"""Llama-2 decoder (``LlamaForCausalLM`` architecture) with HF-compatible ``state_dict``
keys (``model.embed_tokens``, ``model.layers.N.{self_attn.{q,k,v,o}_proj, mlp.{gate,up,
down}_proj, input_layernorm, post_attention_layernorm}``, ``model.norm``, ``lm_head``).

The reference loads ``AutoModelForCausalLM.from_pretrained("NousResearch/Llama-2-7b-hf",
torch_dtype=bf16)`` (02_development/distributed_utils.py:465-468, 484-487), which needs the
network; here the architecture is built locally with random-init weights (``LlamaConfig.
llama2_7b()`` has the 7B shape: hidden 4096, inter 11008, 32 layers, 32 heads, vocab 32000).

RMSNorm, RoPE, SwiGLU and causal + key-padding attention run on the framework's kernels;
residual adds ride in GEMM epilogues; the LM head is fused with the shifted cross-entropy.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

from .. import ops


@dataclass
class LlamaConfig:
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 32
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    pad_token_id: int = 2  # the reference sets pad = eos (distributed_utils.py:458-459)

    @staticmethod
    def llama2_7b() -> "LlamaConfig":
        return LlamaConfig()

    @staticmethod
    def tiny() -> "LlamaConfig":
        return LlamaConfig(vocab_size=512, hidden_size=128, intermediate_size=352, num_hidden_layers=2,
                           num_attention_heads=2, num_key_value_heads=2)


class LlamaRMSNorm(nn.Module):
    def __init__(self, dim, eps):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.eps = eps

    def forward(self, x):
        return ops.rms_norm(x, self.weight, self.eps)

    def fork(self, x):
        """(norm(x), x) with the skip-connection gradient summed inside the norm backward (ops.rms_norm_fork)."""
        return ops.rms_norm_fork(x, self.weight, self.eps)


def _proj(module, x, residual=None):
    """Apply a (possibly LoRA-wrapped) bias-free projection."""
    if hasattr(module, "lora_forward"):
        return module.lora_forward(x, residual=residual)
    return ops.linear(x, module.weight, module.bias, residual=residual)


class LlamaAttention(nn.Module):
    def __init__(self, c: LlamaConfig):
        super().__init__()
        self.nh, self.nkv = c.num_attention_heads, c.num_key_value_heads
        self.hd = c.hidden_size // c.num_attention_heads
        self.theta = c.rope_theta
        self.q_proj = nn.Linear(c.hidden_size, self.nh * self.hd, bias=False)
        self.k_proj = nn.Linear(c.hidden_size, self.nkv * self.hd, bias=False)
        self.v_proj = nn.Linear(c.hidden_size, self.nkv * self.hd, bias=False)
        self.o_proj = nn.Linear(self.nh * self.hd, c.hidden_size, bias=False)

    def forward(self, x, attention_mask, residual):
        B, S, _ = x.shape
        q = _proj(self.q_proj, x).reshape(B, S, self.nh, self.hd)
        k = _proj(self.k_proj, x).reshape(B, S, self.nkv, self.hd)
        v = _proj(self.v_proj, x).reshape(B, S, self.nkv, self.hd)
        q = ops.rope(q, self.theta, layout="bshd").transpose(1, 2)
        k = ops.rope(k, self.theta, layout="bshd").transpose(1, 2)
        v = v.transpose(1, 2)
        if self.nkv != self.nh:  # grouped-query attention: expand kv heads
            rep = self.nh // self.nkv
            k = k.repeat_interleave(rep, dim=1)
            v = v.repeat_interleave(rep, dim=1)
        kpm = attention_mask.bool() if attention_mask is not None else None
        o = ops.attention(q, k, v, causal=True, key_padding_mask=kpm)
        o = o.transpose(1, 2).reshape(B, S, self.nh * self.hd)
        return _proj(self.o_proj, o, residual=residual)


class LlamaMLP(nn.Module):
    def __init__(self, c: LlamaConfig):
        super().__init__()
        self.gate_proj = nn.Linear(c.hidden_size, c.intermediate_size, bias=False)
        self.up_proj = nn.Linear(c.hidden_size, c.intermediate_size, bias=False)
        self.down_proj = nn.Linear(c.intermediate_size, c.hidden_size, bias=False)

    def forward(self, x, residual):
        h = ops.swiglu(_proj(self.gate_proj, x), _proj(self.up_proj, x))
        return _proj(self.down_proj, h, residual=residual)


class LlamaDecoderLayer(nn.Module):
    def __init__(self, c: LlamaConfig):
        super().__init__()
        self.self_attn = LlamaAttention(c)
        self.mlp = LlamaMLP(c)
        self.input_layernorm = LlamaRMSNorm(c.hidden_size, c.rms_norm_eps)
        self.post_attention_layernorm = LlamaRMSNorm(c.hidden_size, c.rms_norm_eps)

    def forward(self, x, attention_mask=None):
        h, skip = self.input_layernorm.fork(x)
        x = self.self_attn(h, attention_mask, residual=skip)
        h, skip = self.post_attention_layernorm.fork(x)
        return self.mlp(h, residual=skip)


class LlamaModel(nn.Module):
    def __init__(self, c: LlamaConfig):
        super().__init__()
        self.embed_tokens = nn.Embedding(c.vocab_size, c.hidden_size)
        self.layers = nn.ModuleList([LlamaDecoderLayer(c) for _ in range(c.num_hidden_layers)])
        self.norm = LlamaRMSNorm(c.hidden_size, c.rms_norm_eps)
        self.use_checkpoint = False

    def forward(self, input_ids, attention_mask=None):
        x = ops.embedding(input_ids, self.embed_tokens.weight)
        if self.use_checkpoint and self.training and torch.is_grad_enabled():
            from ..memory import checkpoint_layers

            x = checkpoint_layers(self.layers, x, attention_mask=attention_mask)
        else:
            for layer in self.layers:
                x = layer(x, attention_mask)
        return self.norm(x)


class CausalLMOutput:
    def __init__(self, loss=None, logits=None):
        self.loss, self.logits = loss, logits


class LlamaForCausalLM(nn.Module):
    def __init__(self, config: LlamaConfig | None = None):
        super().__init__()
        self.config = config or LlamaConfig()
        self.model = LlamaModel(self.config)
        self.lm_head = nn.Linear(self.config.hidden_size, self.config.vocab_size, bias=False)
        for m in self.modules():
            if isinstance(m, (nn.Linear, nn.Embedding)):
                nn.init.normal_(m.weight, mean=0.0, std=0.02)

    def forward(self, input_ids, attention_mask=None, labels=None):
        """HF-style call: with ``labels`` returns ``.loss`` (shifted CE, -100 ignored) and no logits."""
        h = self.model(input_ids, attention_mask)
        if labels is None:
            return CausalLMOutput(logits=_proj(self.lm_head, h))
        hs = h[:, :-1].contiguous()
        loss = ops.linear_cross_entropy(hs, self.lm_head.weight, None, labels[:, 1:].contiguous(), -100)
        return CausalLMOutput(loss=loss)

    def num_parameters(self) -> int:
        return sum(p.numel() for p in self.parameters())
