"""The reference's toy language models, state_dict-compatible, on fused ops.

Three divergent definitions exist in the reference:

* ``SimpleTransformerLM`` -- 02_development/distributed_utils.py:75-88; attributes
  ``embed`` / ``tr`` / ``fc``; 256-d, 4 heads, 2 post-norm ReLU layers, ff 2048,
  dropout 0.1, **no positional encoding and no causal mask**; ignores extra kwargs.
* ``NotebookTransformerLM`` -- core_framework.ipynb c4 L3-18 (and the
  mixed_precision / memory_optimization copies); attributes ``embedding`` /
  ``transformer`` / ``fc``; ``max_len`` stored but unused; ``use_checkpoint`` from
  memory_optimization.ipynb c3 L3-23.
* ``CompileBenchLM`` -- compilation_optimization.py:57-71; 768-d, 12 heads, 4 GELU
  layers, ff 3072, sequence-first input ``idx[T, B]``.

The output head is exposed separately (``features`` + ``fc``) so trainers can use
the fused linear+cross-entropy op (loss and gradients in one pass over L2-sized slices of the logits).
"""
from __future__ import annotations

import torch.nn as nn

from .. import ops
from .transformer import TransformerEncoder, TransformerEncoderLayer


class SimpleTransformerLM(nn.Module):
    def __init__(self, vocab_size: int, emb_dim: int = 256, n_heads: int = 4, n_layers: int = 2,
                 dim_feedforward: int = 2048, dropout: float = 0.1):
        super().__init__()
        self.embed = nn.Embedding(vocab_size, emb_dim)
        self.tr = TransformerEncoder(lambda: TransformerEncoderLayer(emb_dim, n_heads, dim_feedforward, dropout), n_layers)
        self.fc = nn.Linear(emb_dim, vocab_size)

    def features(self, x):
        return self.tr(ops.embedding(x, self.embed.weight))

    def forward(self, x, **_):
        return ops.linear(self.features(x), self.fc.weight, self.fc.bias)

    def loss(self, x, y, ignore_index: int = -100):
        """Mean next-token CE via the fused vocabulary-projection + cross-entropy op."""
        return ops.linear_cross_entropy(self.features(x), self.fc.weight, self.fc.bias, y, ignore_index)


class NotebookTransformerLM(nn.Module):
    def __init__(self, vocab_size: int, emb_dim: int = 256, n_heads: int = 4, n_layers: int = 2, max_len: int = 128,
                 use_checkpoint: bool = False, dropout: float = 0.1):
        super().__init__()
        self.max_len = max_len
        self.embedding = nn.Embedding(vocab_size, emb_dim)
        self.transformer = TransformerEncoder(lambda: TransformerEncoderLayer(emb_dim, n_heads, 2048, dropout), n_layers)
        self.transformer.use_checkpoint = use_checkpoint
        self.fc = nn.Linear(emb_dim, vocab_size)

    @property
    def use_checkpoint(self) -> bool:
        return self.transformer.use_checkpoint

    def features(self, x):
        return self.transformer(ops.embedding(x, self.embedding.weight))

    def forward(self, x, **_):
        return ops.linear(self.features(x), self.fc.weight, self.fc.bias)

    def loss(self, x, y, ignore_index: int = -100):
        return ops.linear_cross_entropy(self.features(x), self.fc.weight, self.fc.bias, y, ignore_index)


class CompileBenchLM(nn.Module):
    """compilation_optimization.py:57-71 -- attributes ``embed`` / ``tr`` / ``fc``; input ``idx[T, B]``."""

    def __init__(self, vocab: int = 50257, d_model: int = 768, nhead: int = 12, dim_ff: int = 3072, n_layers: int = 4):
        super().__init__()
        self.embed = nn.Embedding(vocab, d_model)
        self.tr = TransformerEncoder(lambda: TransformerEncoderLayer(d_model, nhead, dim_ff, 0.1, activation="gelu"), n_layers)
        self.fc = nn.Linear(d_model, vocab)

    def forward(self, idx):
        x = ops.embedding(idx.t(), self.embed.weight)  # [B, T, E]; attention runs over T
        x = self.tr(x)
        return ops.linear(x, self.fc.weight, self.fc.bias).transpose(0, 1)
