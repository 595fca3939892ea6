"""Backward of the tcgen05 flash attention (``hyp_flash_attn_bwd``): dq, dk, dv for ``[B, H, S, D]`` views."""
from __future__ import annotations

import torch

from ._lib import DT, ptr
from .flash import _declare, _strides_bhs, _tma_ok, bwd_call


def flash_bwd(q, k, v, o, lse, do, causal: bool, kpm, scale: float, drop=None):
    _declare()
    B, H, Sq, D = q.shape
    Sk = k.shape[2]
    if not (do.stride(3) == 1 and _tma_ok(do)):
        do = do.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)  # [B,S,H,D] memory, like o
    dq_acc = torch.empty((B, H, Sq, D), device=q.device, dtype=torch.float32)  # cleared by the kernel
    dkv = torch.empty((2, B, Sk, H, D), device=q.device, dtype=q.dtype)
    dk, dv = dkv[0].permute(0, 2, 1, 3), dkv[1].permute(0, 2, 1, 3)
    delta = torch.empty((B, H, Sq), device=q.device, dtype=torch.float32)
    bwd_call((ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq_acc), ptr(dk), ptr(dv),
              ptr(delta), ptr(kpm), B, H, Sq, Sk, D, *_strides_bhs(q), *_strides_bhs(k), *_strides_bhs(v), *_strides_bhs(o),
              *_strides_bhs(do), *_strides_bhs(dk), int(causal), float(scale), DT[q.dtype]), drop)
    return dq_acc.to(q.dtype), dk, dv
