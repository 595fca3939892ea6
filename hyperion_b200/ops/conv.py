"""Conv -> BatchNorm -> ReLU (+ residual) blocks for the ResNets.

Data runs channels-last (NHWC).  Convolution forward is the tcgen05 implicit GEMM of
``csrc/conv_igemm.cu`` (one 4-D TMA box per filter tap, no im2col buffer) whenever the
input channels are a multiple of 64; 1x1 stride-1 convolutions are routed to the plain
tcgen05 GEMM (an NHWC 1x1 conv *is* ``[N*H*W, Cin] @ W[Cout, Cin]^T``); the weight gradient is the
pixel-reduction implicit GEMM ``conv_wgrad_kernel``; the 3-channel stem and -- by default, see ``_igemm_dgrad_on`` --
the data gradient are on cuDNN (a native stride-1 dgrad exists: ``HYPERION_CONV_DGRAD=igemm``).  In training the
BatchNorm-apply + ReLU + residual-add tail is one fused pass (``csrc/bn_act.cu``);
in eval mode BatchNorm is folded into the convolution's epilogue together with the
residual add and the ReLU, so a whole Conv-BN-(+res)-ReLU block is one kernel.

Reference call sites: torchvision resnet18/50 Conv2d->BatchNorm2d->ReLU
(02_development/distributed_utils.py:229; Phase 1/baseline_performance.ipynb c0 L23).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib


def to_channels_last(x: torch.Tensor) -> torch.Tensor:
    return x.contiguous(memory_format=torch.channels_last) if x.is_cuda and x.dim() == 4 else x


def _igemm_ok(x, conv: nn.Conv2d) -> bool:
    import os

    return (x.is_cuda and _lib.available() and x.dim() == 4 and x.dtype in (torch.float16, torch.bfloat16)
            and conv.weight.dtype == x.dtype and conv.groups == 1 and conv.dilation == (1, 1)
            and conv.stride[0] == conv.stride[1] and conv.stride[0] in (1, 2)
            and conv.padding[0] == conv.padding[1] and isinstance(conv.padding[0], int)
            and conv.padding_mode == "zeros" and conv.in_channels % 64 == 0 and conv.out_channels % 32 == 0
            and x.is_contiguous(memory_format=torch.channels_last)
            and os.environ.get("HYPERION_CONV", "igemm") == "igemm")


def _igemm_dgrad_on() -> bool:
    import os

    # default cuDNN: measured per part on the step benches (profiles/bench/r2/r2n_*): ResNet-50 14.93 ms all-native,
    # 14.64 with cuDNN dgrad, 14.72 with cuDNN wgrad, 14.44 with both; ResNet-18/CIFAR 1.41 / 1.19 / 1.48 / 1.26 -- the native
    # dgrad (forward kernel over dy with a filter that has to be flipped and re-laid-out every step, stride 1 only) loses
    # on both, the native wgrad wins on the small model and is within 1.4 % on the large one, so it stays the default
    return os.environ.get("HYPERION_CONV_DGRAD", "cudnn") == "igemm"


def _igemm_wgrad_on() -> bool:
    import os

    return os.environ.get("HYPERION_CONV_WGRAD", "igemm") == "igemm"


def conv2d_igemm(x, weight, stride: int, pad: int, scale=None, shift=None, residual=None, relu: bool = False):
    """Raw call of ``hyp_conv2d_fwd`` on channels-last tensors (no autograd): returns y [N, K, Ho, Wo] channels-last."""
    N, C, H, W = x.shape
    K, _, R, S = weight.shape
    w = weight.contiguous(memory_format=torch.channels_last)   # memory order [K, R, S, C]
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    y = torch.empty((N, K, Ho, Wo), device=x.device, dtype=x.dtype, memory_format=torch.channels_last)
    if residual is not None:
        assert residual.shape == y.shape and residual.is_contiguous(memory_format=torch.channels_last)
    _lib.call("hyp_conv2d_fwd", _lib.ptr(x), _lib.ptr(w), _lib.ptr(y), N, H, W, C, K, R, S, stride, pad,
              _lib.ptr(scale), _lib.ptr(shift), _lib.ptr(residual), int(relu), _lib.DT[x.dtype], _lib.stream_ptr())
    return y


def conv2d_wgrad_igemm(x, dy, weight_shape, stride: int, pad: int, out_flat=None):
    """dW of ``conv2d(x, w)`` on the tcgen05 implicit GEMM: fp32 accumulation in [K, R, S, C] order, rounded to x's
    dtype; returns the [K, C, R, S]-shaped channels_last view (backed by ``out_flat`` when given)."""
    from . import native

    N, C, H, W = x.shape
    K, _, R, S = weight_shape
    acc = torch.zeros(K * R * S * C, device=x.device, dtype=torch.float32)
    _lib.call("hyp_conv2d_wgrad", _lib.ptr(x), _lib.ptr(dy), _lib.ptr(acc), N, H, W, C, K, R, S, stride, pad,
              _lib.DT[x.dtype], _lib.stream_ptr())
    flat = native.cast(acc, x.dtype, out=out_flat)
    return flat.view(K, R, S, C).permute(0, 3, 1, 2)


class _ConvFn(torch.autograd.Function):
    """Convolution on conv_igemm_tcgen05: forward, stride-1 dgrad (same kernel, flipped filter) and wgrad (pixel-reduction
    GEMM with split reduction); stride-2 dgrad and weights that are not channels_last go through ATen (cuDNN)."""

    @staticmethod
    def forward(ctx, x, weight, stride, pad):
        from . import gradslot

        ctx.save_for_backward(x, weight)
        ctx.stride, ctx.pad = stride, pad
        ctx.slot = gradslot.slot_of(weight)
        return conv2d_igemm(x, weight, stride, pad)

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        dy = dy.contiguous(memory_format=torch.channels_last)
        K, C, R, S = weight.shape
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        dx = None
        if need_x and ctx.stride == 1 and K % 64 == 0 and C % 32 == 0 and R - 1 - ctx.pad >= 0 and _igemm_dgrad_on():
            # stride-1 dgrad is the same implicit GEMM over dy with the filter flipped and its channel roles swapped:
            # dx = conv(dy, w'[c, r, s, k] = w[k, R-1-r, S-1-s, c], pad' = R-1-pad)
            wt = weight.flip(2, 3).transpose(0, 1).contiguous(memory_format=torch.channels_last)
            dx = conv2d_igemm(dy, wt, 1, R - 1 - ctx.pad)
            need_x = False
        dw = None
        if (need_w and C % 64 == 0 and K % 64 == 0 and weight.is_contiguous(memory_format=torch.channels_last)
                and _igemm_wgrad_on()):
            from . import gradslot

            dst = gradslot.claim(ctx.slot, weight.dtype)
            if dst is not None and not dst.is_contiguous(memory_format=torch.channels_last):
                dst = None
            out_flat = dst.permute(0, 2, 3, 1).reshape(-1) if dst is not None else None   # the slot's [K,R,S,C] memory
            dw = conv2d_wgrad_igemm(x, dy, weight.shape, ctx.stride, ctx.pad, out_flat)
            if dst is not None:
                dw = dst
            need_w = False
        if need_x or need_w:
            gx, gw, _ = torch.ops.aten.convolution_backward(
                dy, x, weight, None, [ctx.stride] * 2, [ctx.pad] * 2, [1, 1], False, [0, 0], 1, [need_x, need_w, False])
            dx = dx if dx is not None else gx
            dw = dw if dw is not None else gw
        return dx, dw, None, None


def _conv(x, conv: nn.Conv2d):
    from . import functional as Fn

    if (x.is_cuda and _lib.available() and conv.kernel_size == (1, 1) and conv.stride == (1, 1) and conv.padding == (0, 0)
            and conv.groups == 1 and x.dtype in (torch.float16, torch.bfloat16) and conv.weight.dtype == x.dtype
            and x.is_contiguous(memory_format=torch.channels_last) and conv.in_channels % 8 == 0
            and conv.out_channels % 8 == 0):
        # NHWC 1x1 convolution == GEMM over the pixel rows
        N, C, H, W = x.shape
        x2 = x.permute(0, 2, 3, 1).reshape(N * H * W, C)
        y2 = Fn.linear(x2, conv.weight.reshape(conv.out_channels, C), conv.bias)
        return y2.reshape(N, H, W, conv.out_channels).permute(0, 3, 1, 2)
    if _igemm_ok(x, conv):
        y = _ConvFn.apply(x, conv.weight, conv.stride[0], conv.padding[0])
        return y if conv.bias is None else y + conv.bias.view(1, -1, 1, 1)
    return F.conv2d(x, conv.weight, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups)


_bn_declared = False


def _declare_bn():
    global _bn_declared
    if not _bn_declared:
        _lib.declare("hyp_bn_act_fwd", "pppppppppp" + "li" + "ff" + "ii" + "p")
        _lib.declare("hyp_bn_act_bwd", "ppppppppp" + "li" + "ii" + "p")
        if hasattr(_lib.raw(), "hyp_bn_act_fwd_p16"):
            _lib.declare("hyp_bn_act_fwd_p16", "pppppppppp" + "li" + "ff" + "ii" + "p")
            _lib.declare("hyp_bn_act_bwd_p16", "ppppppppp" + "li" + "ii" + "pp" + "p")
        _bn_declared = True


# BatchNorm ``num_batches_tracked`` counters: one fused add for all layers instead of one tiny kernel per layer and step.
_pending_nbt: list = []


def _count_batch(bn) -> None:
    if bn.num_batches_tracked is not None:
        _pending_nbt.append(bn.num_batches_tracked)
        if len(_pending_nbt) >= 512:
            flush_bn_counters()


def flush_bn_counters() -> None:
    """Apply the pending ``num_batches_tracked += 1`` of every BatchNorm that ran since the last flush (one multi-tensor
    kernel).  The training engine calls it once per step; checkpoint export calls it before reading the buffers."""
    if _pending_nbt:
        torch._foreach_add_(list(_pending_nbt), 1)
        _pending_nbt.clear()


class _BnActFn(torch.autograd.Function):
    """Training-mode BatchNorm2d + (residual add) + ReLU on a channels-last tensor: csrc/bn_act.cu."""

    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, eps, momentum, relu):
        _declare_bn()
        N, C, H, W = x.shape
        rows = N * H * W
        y = torch.empty_like(x, memory_format=torch.channels_last)
        # one zero-fill serves the forward statistics and, later, the backward reductions of this layer
        zbuf = torch.zeros(4 * C, device=x.device, dtype=torch.float32)
        sums, ctx.dsums = zbuf[:2 * C], zbuf[2 * C:]
        mean = torch.empty(C, device=x.device, dtype=torch.float32)
        rstd = torch.empty(C, device=x.device, dtype=torch.float32)
        # 16-bit gamma / beta go to the kernels as they are (hyp_bn_act_*_p16); fp32 ones take the original entry points
        p16 = gamma.dtype == x.dtype and beta.dtype == x.dtype and hasattr(_lib.raw(), "hyp_bn_act_fwd_p16")
        gf, bf = (gamma.detach(), beta.detach()) if p16 else (gamma.float(), beta.float())
        ctx.p16 = p16
        _lib.call("hyp_bn_act_fwd_p16" if p16 else "hyp_bn_act_fwd", _lib.ptr(x), _lib.ptr(residual), _lib.ptr(gf), _lib.ptr(bf), _lib.ptr(y), _lib.ptr(sums),
                  _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(running_mean), _lib.ptr(running_var), rows, C, float(eps),
                  float(momentum), int(relu), _lib.DT[x.dtype], _lib.stream_ptr())
        ctx.save_for_backward(x, y, mean, rstd, gf)
        ctx.relu, ctx.has_res, ctx.pdtype = relu, residual is not None, gamma.dtype
        from . import gradslot

        ctx.slots = (gradslot.slot_of(gamma), gradslot.slot_of(beta))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, mean, rstd, gf = ctx.saved_tensors
        N, C, H, W = x.shape
        dy = dy.contiguous(memory_format=torch.channels_last)
        dx = torch.empty_like(x, memory_format=torch.channels_last)
        dres = torch.empty_like(x, memory_format=torch.channels_last) if ctx.has_res else None
        dsums = ctx.dsums
        if ctx.p16:
            # dgamma / dbeta are rounded by the dx kernel itself, straight into the parameters' slots of the flat gradient
            # buffer when they can be claimed (no cast launches, no accumulate copies)
            from . import gradslot

            dg = gradslot.claim(ctx.slots[0], ctx.pdtype)
            db = gradslot.claim(ctx.slots[1], ctx.pdtype)
            if dg is None:
                dg = torch.empty(C, device=x.device, dtype=ctx.pdtype)
            if db is None:
                db = torch.empty(C, device=x.device, dtype=ctx.pdtype)
            _lib.call("hyp_bn_act_bwd_p16", _lib.ptr(dy), _lib.ptr(x), _lib.ptr(y), _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(gf),
                      _lib.ptr(dsums), _lib.ptr(dx), _lib.ptr(dres), N * H * W, C, int(ctx.relu), _lib.DT[x.dtype],
                      _lib.ptr(dg), _lib.ptr(db), _lib.stream_ptr())
            return dx, dg, db, dres, None, None, None, None, None
        _lib.call("hyp_bn_act_bwd", _lib.ptr(dy), _lib.ptr(x), _lib.ptr(y), _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(gf),
                  _lib.ptr(dsums), _lib.ptr(dx), _lib.ptr(dres), N * H * W, C, int(ctx.relu), _lib.DT[x.dtype], _lib.stream_ptr())
        return dx, dsums[C:].to(ctx.pdtype), dsums[:C].to(ctx.pdtype), dres, None, None, None, None, None


def _bn_native_ok(y, bn, residual) -> bool:
    return (bn.training and y.is_cuda and _lib.available() and y.dtype in (torch.float16, torch.bfloat16)
            and y.dim() == 4 and y.shape[1] % 8 == 0 and y.is_contiguous(memory_format=torch.channels_last)
            and bn.track_running_stats and bn.momentum is not None
            and (residual is None or (residual.shape == y.shape and residual.is_contiguous(memory_format=torch.channels_last))))


def _folded_bn(conv: nn.Conv2d, bn: nn.BatchNorm2d):
    """fp32 per-channel (scale, shift) of an eval-mode BatchNorm (and the conv bias) folded into the convolution;
    cached on the BatchNorm module until one of its tensors is modified or replaced."""
    ts = [t for t in (bn.weight, bn.bias, bn.running_mean, bn.running_var, conv.bias) if t is not None]
    key = tuple((t.data_ptr(), t._version) for t in ts) + (bn.eps,)
    cached = getattr(bn, "_hyp_folded", None)
    if cached is not None and cached[0] == key:
        return cached[1], cached[2]
    with torch.no_grad():
        scale = torch.rsqrt(bn.running_var.float() + bn.eps)
        if bn.affine:
            scale = scale * bn.weight.float()
        shift = -bn.running_mean.float() * scale
        if bn.affine:
            shift = shift + bn.bias.float()
        if conv.bias is not None:
            shift = shift + conv.bias.float() * scale
        scale, shift = scale.contiguous(), shift.contiguous()
    bn._hyp_folded = (key, scale, shift)
    return scale, shift


def conv_bn_act(x, conv: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool = True, residual=None):
    if (not bn.training and not torch.is_grad_enabled() and bn.track_running_stats and _igemm_ok(x, conv)
            and (residual is None or residual.is_contiguous(memory_format=torch.channels_last))):
        # inference: BatchNorm folded into the convolution epilogue (+ conv bias, residual, ReLU) -- one kernel per block
        scale, shift = _folded_bn(conv, bn)
        return conv2d_igemm(x, conv.weight, conv.stride[0], conv.padding[0], scale, shift, residual, relu)
    y = _conv(x, conv)
    if _bn_native_ok(y, bn, residual):
        _count_batch(bn)
        return _BnActFn.apply(y, bn.weight, bn.bias, residual, bn.running_mean, bn.running_var, bn.eps, bn.momentum, relu)
    if bn.training:
        if bn.num_batches_tracked is not None:
            bn.num_batches_tracked.add_(1)
    if y.device.type == "cpu" and y.dtype != torch.float32:
        # ATen's CPU batch norm wants one dtype for weight and running statistics; 16-bit modules keep fp32 statistics
        w = bn.weight.float() if bn.weight is not None else None
        b = bn.bias.float() if bn.bias is not None else None
        y = F.batch_norm(y.float(), bn.running_mean, bn.running_var, w, b, bn.training, bn.momentum or 0.0, bn.eps).to(y.dtype)
    else:
        y = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training, bn.momentum or 0.0, bn.eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y
