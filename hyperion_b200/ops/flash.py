"""Autograd wrapper of the tcgen05 flash-attention kernels (``csrc/attention.cu``).

Inputs are ``[B, H, S, D]`` *views* with unit stride in D (e.g. slices of the packed ``[B, S, 3, H, D]`` projection
output); the kernels read them in place through 4-D TMA descriptors.  The output is produced in ``[B, S, H, D]`` memory
order and returned as a ``[B, H, S, D]`` view, so the usual ``transpose(1, 2).reshape(B, S, E)`` that follows is free.
"""
from __future__ import annotations

import math

import torch

from . import _lib, rng
from ._lib import DT, call, ptr, stream_ptr

_declared = False


def _declare():
    global _declared
    if _declared:
        return
    _lib.declare("hyp_flash_attn_fwd", "ppppp" + "iiiii" + "l" * 12 + "p" + "i" + "f" + "i" + "p")
    lib = _lib.require()
    if hasattr(lib, "hyp_flash_attn_bwd"):
        _lib.declare("hyp_flash_attn_bwd", "pppppp" + "ppp" + "pp" + "iiiii" + "l" * 18 + "ifip")
        _lib.declare("hyp_flash_dq_convert", "pp" + "iiii" + "lll" + "ip")
    if hasattr(lib, "hyp_flash_attn_fwd_drop"):
        _lib.declare("hyp_flash_attn_fwd_drop", "ppppp" + "iiiii" + "l" * 12 + "p" + "i" + "f" + "i" + "pif" + "p")
        _lib.declare("hyp_flash_attn_bwd_drop", "pppppp" + "ppp" + "pp" + "iiiii" + "l" * 18 + "ifi" + "pif" + "p")
    _declared = True


def _strides_bhs(t: torch.Tensor):
    assert t.dim() == 4 and t.stride(3) == 1, "expected [B, H, S, D] with unit stride in D"
    return t.stride(0), t.stride(1), t.stride(2)


def _tma_ok(t: torch.Tensor) -> bool:
    return t.data_ptr() % 16 == 0 and all(s % 8 == 0 for s in t.stride()[:3])


def make_drop(device, dropout_p: float):
    """``(rng_state, salt, p)`` of one dropout call, or None (see ops/rng.py)."""
    return (rng.state(device), rng.next_salt(), float(dropout_p)) if dropout_p > 0.0 else None


def bwd_call(args_before_stream, drop):
    """hyp_flash_attn_bwd / _drop with the forward's dropout triple."""
    if drop is None:
        call("hyp_flash_attn_bwd", *args_before_stream, stream_ptr())
    else:
        call("hyp_flash_attn_bwd_drop", *args_before_stream, ptr(drop[0]), int(drop[1]), float(drop[2]), stream_ptr())


def flash_fwd(q, k, v, causal: bool, key_padding_mask, scale: float, drop=None):
    """Returns (o as [B,H,S,D] view over [B,S,H,D] memory, lse [B,H,Sq] fp32).  ``drop``: see :func:`make_drop`."""
    _declare()
    B, H, Sq, D = q.shape
    Sk = k.shape[2]
    q, k, v = (t if _tma_ok(t) else t.contiguous() for t in (q, k, v))
    o_mem = torch.empty((B, Sq, H, D), device=q.device, dtype=q.dtype)
    o = o_mem.permute(0, 2, 1, 3)
    lse = torch.empty((B, H, Sq), device=q.device, dtype=torch.float32)
    kpm = None
    if key_padding_mask is not None:
        kpm = key_padding_mask.to(torch.uint8).contiguous()
    args = (ptr(q), ptr(k), ptr(v), ptr(o_mem), ptr(lse), B, H, Sq, Sk, D, *_strides_bhs(q),
            *_strides_bhs(k), *_strides_bhs(v), *_strides_bhs(o), ptr(kpm), int(causal), float(scale), DT[q.dtype])
    if drop is None:
        call("hyp_flash_attn_fwd", *args, stream_ptr())
    else:
        call("hyp_flash_attn_fwd_drop", *args, ptr(drop[0]), int(drop[1]), float(drop[2]), stream_ptr())
    return o, lse, (q, k, v, kpm)


class _FlashFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, causal, key_padding_mask, scale, dropout_p=0.0):
        ctx.drop = make_drop(q.device, dropout_p)
        o, lse, (qc, kc, vc, kpm) = flash_fwd(q, k, v, causal, key_padding_mask, scale, ctx.drop)
        ctx.save_for_backward(qc, kc, vc, o, lse, kpm if kpm is not None else torch.empty(0, device=q.device))
        ctx.causal, ctx.scale, ctx.has_kpm = causal, scale, kpm is not None
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse, kpm = ctx.saved_tensors
        kpm = kpm if ctx.has_kpm else None
        lib = _lib.require()
        if hasattr(lib, "hyp_flash_attn_bwd"):
            from .flash_bwd import flash_bwd

            dq, dk, dv = flash_bwd(q, k, v, o, lse, do, ctx.causal, kpm, ctx.scale, ctx.drop)
            return dq, dk, dv, None, None, None, None
        # recompute-based backward in plain PyTorch (only used when the backward kernel is not built)
        from .attn import attention_reference

        with torch.enable_grad():
            qq, kk, vv = (t.detach().requires_grad_(True) for t in (q, k, v))
            out = attention_reference(qq, kk, vv, ctx.causal, kpm.bool() if kpm is not None else None, 0.0, ctx.scale)
            dq, dk, dv = torch.autograd.grad(out, (qq, kk, vv), do)
        return dq, dk, dv, None, None, None, None


def flash_attention(q, k, v, causal: bool = False, key_padding_mask=None, scale: float | None = None,
                    dropout_p: float = 0.0):
    scale = scale if scale is not None else 1.0 / math.sqrt(q.shape[-1])
    return _FlashFn.apply(q, k, v, causal, key_padding_mask, scale, dropout_p)


class _FlashPackedFn(torch.autograd.Function):
    """Attention over the packed projection output ``qkv[B, S, 3, H, D]``; the backward writes dq / dk / dv straight
    into one ``[B, S, 3, H, D]`` gradient buffer (no per-slice zero-fill + copy + add in autograd)."""

    @staticmethod
    def forward(ctx, qkv, causal, key_padding_mask, scale, dropout_p=0.0):
        q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        ctx.drop = make_drop(qkv.device, dropout_p)
        o, lse, (_, _, _, kpm) = flash_fwd(q, k, v, causal, key_padding_mask, scale, ctx.drop)
        ctx.save_for_backward(qkv, o, lse, kpm if kpm is not None else torch.empty(0, device=qkv.device))
        ctx.causal, ctx.scale, ctx.has_kpm = causal, scale, kpm is not None
        return o

    @staticmethod
    def backward(ctx, do):
        qkv, o, lse, kpm = ctx.saved_tensors
        kpm = kpm if ctx.has_kpm else None
        _declare()
        B, S, _, H, D = qkv.shape
        q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        if not (do.stride(3) == 1 and _tma_ok(do)):
            do = do.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
        dqkv = torch.empty_like(qkv)
        dk, dv = dqkv[:, :, 1].permute(0, 2, 1, 3), dqkv[:, :, 2].permute(0, 2, 1, 3)
        dq_acc = torch.empty((B, H, S, D), device=qkv.device, dtype=torch.float32)  # cleared by the kernel
        delta = torch.empty((B, H, S), device=qkv.device, dtype=torch.float32)
        bwd_call((ptr(q), ptr(k), ptr(v), ptr(o), ptr(do), ptr(lse), ptr(dq_acc), ptr(dk), ptr(dv),
                  ptr(delta), ptr(kpm), B, H, S, S, D, *_strides_bhs(q), *_strides_bhs(k), *_strides_bhs(v), *_strides_bhs(o),
                  *_strides_bhs(do), *_strides_bhs(dk), int(ctx.causal), float(ctx.scale), DT[qkv.dtype]), ctx.drop)
        dq = dqkv[:, :, 0].permute(0, 2, 1, 3)
        call("hyp_flash_dq_convert", ptr(dq_acc), ptr(dq), B, H, S, D, *_strides_bhs(dq), DT[qkv.dtype], stream_ptr())
        return dqkv, None, None, None, None


def flash_attention_qkvpacked(qkv, causal: bool = False, key_padding_mask=None, scale: float | None = None,
                              dropout_p: float = 0.0):
    """``qkv``: ``[B, S, 3, H, D]`` (contiguous).  Returns ``[B, H, S, D]`` (a view over ``[B, S, H, D]`` memory)."""
    scale = scale if scale is not None else 1.0 / math.sqrt(qkv.shape[-1])
    return _FlashPackedFn.apply(qkv, causal, key_padding_mask, scale, dropout_p)
