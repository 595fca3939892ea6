"""Scaled-dot-product attention front end.

``attention(q, k, v)`` takes ``[B, H, S, D]`` tensors.  On CUDA with head dim 64
or 128 and a 16-bit dtype it runs the hand-written tcgen05 flash kernels
(``csrc/attention.cu``: S = QK^T and O += PV on the tensor cores with TMEM
accumulators, online softmax in registers, variants none / causal /
causal+key-padding, optional Philox dropout on the probabilities that the backward
regenerates instead of storing).  Elsewhere (CPU, fp32, odd head dims) it evaluates the same
maths with plain PyTorch ops -- the oracle for the kernel tests.

Reference call sites: nn.MultiheadAttention inside nn.TransformerEncoderLayer
(02_development/distributed_utils.py:79), HF Llama SDPA (distributed_utils.py:465).
"""
from __future__ import annotations

import math

import torch
import torch.nn.functional as F

from . import _lib


def attention_reference(q, k, v, causal=False, key_padding_mask=None, dropout_p=0.0, scale=None):
    """Plain-PyTorch attention (fp32 softmax), used on CPU and as the test oracle."""
    D = q.shape[-1]
    scale = scale if scale is not None else 1.0 / math.sqrt(D)
    s = torch.matmul(q.float(), k.float().transpose(-1, -2)) * scale
    Sq, Sk = s.shape[-2], s.shape[-1]
    if causal:
        m = torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).tril(diagonal=Sk - Sq)
        s = s.masked_fill(~m, float("-inf"))
    if key_padding_mask is not None:
        s = s.masked_fill(~key_padding_mask[:, None, None, :].bool(), float("-inf"))
    p = torch.softmax(s, dim=-1)
    p = torch.nan_to_num(p)  # fully-masked rows -> zeros
    if dropout_p > 0.0:
        p = F.dropout(p, dropout_p)
    return torch.matmul(p, v.float()).to(q.dtype)


def _flash_supported(q, k, v, dropout_p) -> bool:
    if not (q.is_cuda and _lib.available()):
        return False
    lib = _lib.raw()
    if lib is None or not hasattr(lib, "hyp_flash_attn_fwd"):
        return False
    if dropout_p > 0.0:
        from .functional import native_dropout_enabled

        if not (native_dropout_enabled() and hasattr(lib, "hyp_flash_attn_fwd_drop")):
            return False
    return q.dtype in (torch.float16, torch.bfloat16) and q.shape[-1] in (64, 128)


def attention_qkvpacked(qkv, causal=False, key_padding_mask=None, dropout_p=0.0, scale=None):
    """Attention over the packed projection output ``qkv[B, S, 3, H, D]``; returns ``[B, H, S, D]``."""
    import os

    if (qkv.is_contiguous() and _flash_supported(qkv, qkv, qkv, dropout_p) and qkv.data_ptr() % 16 == 0
            and os.environ.get("HYPERION_ATTN", "flash") == "flash"):
        from .flash import flash_attention_qkvpacked

        return flash_attention_qkvpacked(qkv, causal=causal, key_padding_mask=key_padding_mask, scale=scale,
                                         dropout_p=dropout_p)
    q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    return attention(q, k, v, causal, key_padding_mask, dropout_p, scale)


def attention(q, k, v, causal=False, key_padding_mask=None, dropout_p=0.0, scale=None):
    import os

    if _flash_supported(q, k, v, dropout_p) and os.environ.get("HYPERION_ATTN", "flash") == "flash":
        from .flash import flash_attention

        return flash_attention(q, k, v, causal=causal, key_padding_mask=key_padding_mask, scale=scale, dropout_p=dropout_p)
    if q.is_cuda:
        # Library fallback (cuDNN / flash SDPA) for shapes the native kernel does not cover yet.
        mask = None
        if key_padding_mask is not None:
            mask = key_padding_mask[:, None, None, :].bool()
            if causal:
                Sq, Sk = q.shape[-2], k.shape[-2]
                mask = mask & torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).tril(diagonal=Sk - Sq)
            return F.scaled_dot_product_attention(q, k, v, attn_mask=mask, dropout_p=dropout_p, scale=scale)
        return F.scaled_dot_product_attention(q, k, v, dropout_p=dropout_p, is_causal=causal, scale=scale)
    return attention_reference(q, k, v, causal, key_padding_mask, dropout_p, scale)
