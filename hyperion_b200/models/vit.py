"""ViT-B/16 with torchvision-compatible ``state_dict`` keys (``conv_proj``,
``class_token``, ``encoder.pos_embedding``, ``encoder.layers.encoder_layer_N.{ln_1,
self_attention, ln_2, mlp.0, mlp.3}``, ``encoder.ln``, ``heads.head``).

The reference asks for ``torchvision.models.vit_b_16(pretrained=False)`` (Phase 1/
baseline_performance.ipynb c0 L25-39) but its published run fell back to a CNN
(SURVEY.md A-1); BASELINE config 3 wants the real ViT-B/16 with activation
checkpointing.  The 16x16/16 patch embedding is a GEMM over unfolded patches and
runs on gemm_tcgen05; blocks are pre-norm with GELU(erf).
"""
from __future__ import annotations

from collections import OrderedDict

import torch
import torch.nn as nn

from .. import ops
from .transformer import MultiheadSelfAttention


class _MLP(nn.Sequential):
    """Keys ``0`` (Linear), ``3`` (Linear) as in torchvision's MLPBlock (GELU and Dropout sit at 1, 2, 4)."""

    def __init__(self, dim, hidden, dropout):
        super().__init__(nn.Linear(dim, hidden), nn.GELU(), nn.Dropout(dropout), nn.Linear(hidden, dim), nn.Dropout(dropout))


class EncoderBlock(nn.Module):
    def __init__(self, dim, heads, hidden, dropout=0.0):
        super().__init__()
        self.ln_1 = nn.LayerNorm(dim, eps=1e-6)
        self.self_attention = MultiheadSelfAttention(dim, heads, 0.0)
        self.ln_2 = nn.LayerNorm(dim, eps=1e-6)
        self.mlp = _MLP(dim, hidden, dropout)

    def forward(self, x):
        h, skip = ops.layer_norm_fork(x, self.ln_1.weight, self.ln_1.bias, 1e-6)
        x = self.self_attention(h, residual=skip)
        h, skip = ops.layer_norm_fork(x, self.ln_2.weight, self.ln_2.bias, 1e-6)
        return ops.mlp(h, self.mlp[0].weight, self.mlp[0].bias, self.mlp[3].weight, self.mlp[3].bias, act="gelu", residual=skip)


class Encoder(nn.Module):
    def __init__(self, seq_len, layers, dim, heads, hidden, dropout=0.0):
        super().__init__()
        self.pos_embedding = nn.Parameter(torch.empty(1, seq_len, dim).normal_(std=0.02))
        self.layers = nn.Sequential(OrderedDict((f"encoder_layer_{i}", EncoderBlock(dim, heads, hidden, dropout)) for i in range(layers)))
        self.ln = nn.LayerNorm(dim, eps=1e-6)
        self.use_checkpoint = False

    def forward(self, x):
        x = x + self.pos_embedding
        if self.use_checkpoint and self.training and torch.is_grad_enabled():
            from ..memory import checkpoint_layers

            x = checkpoint_layers(list(self.layers), x)
        else:
            for blk in self.layers:
                x = blk(x)
        return ops.layer_norm(x, self.ln.weight, self.ln.bias, 1e-6)


class VisionTransformer(nn.Module):
    def __init__(self, image_size=224, patch_size=16, num_layers=12, num_heads=12, hidden_dim=768, mlp_dim=3072,
                 num_classes=1000, dropout=0.0):
        super().__init__()
        self.patch_size, self.hidden_dim = patch_size, hidden_dim
        self.conv_proj = nn.Conv2d(3, hidden_dim, patch_size, patch_size)
        seq = (image_size // patch_size) ** 2 + 1
        self.class_token = nn.Parameter(torch.zeros(1, 1, hidden_dim))
        self.encoder = Encoder(seq, num_layers, hidden_dim, num_heads, mlp_dim, dropout)
        self.heads = nn.Sequential(OrderedDict(head=nn.Linear(hidden_dim, num_classes)))

    def forward(self, x):
        B, C, H, W = x.shape
        p = self.patch_size
        # patch embedding == GEMM: [B*nh*nw, C*p*p] @ W[hidden, C*p*p]^T
        patches = x.reshape(B, C, H // p, p, W // p, p).permute(0, 2, 4, 1, 3, 5).reshape(B, (H // p) * (W // p), C * p * p)
        tok = ops.linear(patches, self.conv_proj.weight.reshape(self.hidden_dim, -1), self.conv_proj.bias)
        tok = torch.cat([self.class_token.expand(B, -1, -1).to(tok.dtype), tok], dim=1)
        tok = self.encoder(tok)
        return ops.linear(tok[:, 0].contiguous(), self.heads.head.weight, self.heads.head.bias)


def vit_b_16(num_classes: int = 1000) -> VisionTransformer:
    return VisionTransformer(224, 16, 12, 12, 768, 3072, num_classes)


def create_vit_model(allow_fallback: bool = False):
    """Factory name from baseline_performance.ipynb c0 L25; returns the real ViT-B/16 (no silent CNN fallback)."""
    return vit_b_16()
