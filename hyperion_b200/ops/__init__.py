"""Hot ops of the framework: autograd functions over the sm_100a kernels (see functional.py)."""
from . import _lib
from .functional import (attention, attention_qkvpacked, cross_entropy, dropout, embedding, layer_norm, layer_norm_fork, linear, linear_cross_entropy, lora_linear,
                         mlp, rms_norm, rms_norm_fork, rope, swiglu)

__all__ = ["attention", "attention_qkvpacked", "cross_entropy", "dropout", "embedding", "layer_norm", "layer_norm_fork", "linear", "linear_cross_entropy", "lora_linear", "mlp",
           "rms_norm", "rms_norm_fork", "rope", "swiglu", "native_available"]


def native_available() -> bool:
    return _lib.available()
