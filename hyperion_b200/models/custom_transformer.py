"""``create_custom_transformer`` of the baseline benchmark: a bare 6-layer encoder
(d 512, 8 heads, ff 2048, batch-first), float input ``[B, 16, 512]`` -> same shape, no
embedding and no head (Phase 1/baseline_performance.ipynb c0 L56-67).  The module
*is* the encoder, so its keys are ``layers.N.*`` exactly like the reference's
``nn.TransformerEncoder`` instance."""
from __future__ import annotations

from .transformer import TransformerEncoder, TransformerEncoderLayer


class CustomTransformer(TransformerEncoder):
    def __init__(self, d_model: int = 512, nhead: int = 8, num_layers: int = 6, dim_feedforward: int = 2048,
                 dropout: float = 0.1):
        super().__init__(lambda: TransformerEncoderLayer(d_model, nhead, dim_feedforward, dropout), num_layers)


def create_custom_transformer(vocab_size: int = 10000, d_model: int = 512, nhead: int = 8, num_layers: int = 6):
    """Same signature as the reference factory; ``vocab_size`` is accepted and unused there too."""
    return CustomTransformer(d_model, nhead, num_layers)
