"""Transformer encoder building blocks whose ``state_dict`` keys equal those of
``torch.nn.TransformerEncoder`` / ``nn.TransformerEncoderLayer`` /
``nn.MultiheadAttention`` (the classes the reference instantiates at
02_development/distributed_utils.py:79-80, Phase 1/baseline_performance.ipynb c0
L56-67 and 02_development/compilation_optimization.py:57-71) but whose forward
runs on the framework's fused ops (``hyperion_b200.ops``).

Layout: activations are batch-first ``[B, S, E]`` internally; callers with
sequence-first tensors transpose at the boundary (attention over the sequence
dimension is the same function either way).
"""
from __future__ import annotations

import torch
import torch.nn as nn

from .. import ops


class MultiheadSelfAttention(nn.Module):
    """Parameter names match ``nn.MultiheadAttention``: in_proj_weight/bias, out_proj.weight/bias."""

    def __init__(self, embed_dim: int, num_heads: int, dropout: float = 0.0):
        super().__init__()
        assert embed_dim % num_heads == 0
        self.embed_dim, self.num_heads, self.dropout = embed_dim, num_heads, dropout
        self.head_dim = embed_dim // num_heads
        self.in_proj_weight = nn.Parameter(torch.empty(3 * embed_dim, embed_dim))
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * embed_dim))
        self.out_proj = nn.Linear(embed_dim, embed_dim)
        nn.init.xavier_uniform_(self.in_proj_weight)
        nn.init.zeros_(self.out_proj.bias)

    def forward(self, x, causal: bool = False, key_padding_mask=None, residual=None):
        B, S, E = x.shape
        qkv = ops.linear(x, self.in_proj_weight, self.in_proj_bias)
        qkv = qkv.reshape(B, S, 3, self.num_heads, self.head_dim)
        p = self.dropout if self.training else 0.0
        o = ops.attention_qkvpacked(qkv, causal=causal, key_padding_mask=key_padding_mask, dropout_p=p)
        o = o.transpose(1, 2).reshape(B, S, E)
        return ops.linear(o, self.out_proj.weight, self.out_proj.bias, residual=residual)


class TransformerEncoderLayer(nn.Module):
    """Post-norm (default) or pre-norm encoder layer; keys: self_attn, linear1, linear2, norm1, norm2."""

    def __init__(self, d_model: int, nhead: int, dim_feedforward: int = 2048, dropout: float = 0.1,
                 activation: str = "relu", norm_first: bool = False, layer_norm_eps: float = 1e-5):
        super().__init__()
        self.self_attn = MultiheadSelfAttention(d_model, nhead, dropout)
        self.linear1 = nn.Linear(d_model, dim_feedforward)
        self.linear2 = nn.Linear(dim_feedforward, d_model)
        self.norm1 = nn.LayerNorm(d_model, eps=layer_norm_eps)
        self.norm2 = nn.LayerNorm(d_model, eps=layer_norm_eps)
        self.p = dropout
        self.activation = activation
        self.norm_first = norm_first

    def _ffn(self, x, residual=None, out_dropout: bool = True):
        """linear2(dropout(act(linear1(x)))) [+ dropout2 + residual].  ``out_dropout=False``: the caller folds dropout2
        into the norm kernel that consumes the result (post-norm)."""
        if self.training and self.p > 0.0:
            if not out_dropout:
                return ops.mlp(x, self.linear1.weight, self.linear1.bias, self.linear2.weight, self.linear2.bias,
                               act=self.activation, residual=residual, dropout_p=self.p)
            y = ops.dropout(ops.mlp(x, self.linear1.weight, self.linear1.bias, self.linear2.weight, self.linear2.bias,
                                    act=self.activation, dropout_p=self.p), self.p, True)
            return y if residual is None else y + residual
        return ops.mlp(x, self.linear1.weight, self.linear1.bias, self.linear2.weight, self.linear2.bias,
                       act=self.activation, residual=residual)

    def forward(self, x, causal: bool = False, key_padding_mask=None):
        drop = self.training and self.p > 0.0
        if self.norm_first:
            h, skip = ops.layer_norm_fork(x, self.norm1.weight, self.norm1.bias, self.norm1.eps)
            if drop:
                x = skip + ops.dropout(self.self_attn(h, causal, key_padding_mask), self.p, True)
            else:
                x = self.self_attn(h, causal, key_padding_mask, residual=skip)
            h, skip = ops.layer_norm_fork(x, self.norm2.weight, self.norm2.bias, self.norm2.eps)
            return self._ffn(h, residual=skip)
        # post-norm: x = LN(x + sublayer(x)); the residual add is fused into the norm kernel
        # dropout1 / dropout2 of the residual branches ride inside the add + norm kernels (z = dropout(branch) + x)
        pd = self.p if drop else 0.0
        a = self.self_attn(x, causal, key_padding_mask)
        x = ops.layer_norm(a, self.norm1.weight, self.norm1.bias, self.norm1.eps, residual=x, dropout_p=pd)
        f = self._ffn(x, out_dropout=False)
        return ops.layer_norm(f, self.norm2.weight, self.norm2.bias, self.norm2.eps, residual=x, dropout_p=pd)


class TransformerEncoder(nn.Module):
    """Stack with keys ``layers.N.*`` (+ optional final ``norm``), like ``nn.TransformerEncoder``."""

    def __init__(self, make_layer, num_layers: int, norm: nn.Module | None = None):
        super().__init__()
        self.layers = nn.ModuleList([make_layer() for _ in range(num_layers)])
        self.norm = norm
        self.use_checkpoint = False

    def forward(self, x, causal: bool = False, key_padding_mask=None):
        if self.use_checkpoint and self.training and torch.is_grad_enabled():
            from ..memory import checkpoint_layers

            x = checkpoint_layers(self.layers, x, causal=causal, key_padding_mask=key_padding_mask)
        else:
            for layer in self.layers:
                x = layer(x, causal, key_padding_mask)
        if self.norm is not None:
            x = ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        return x
