"""Conv -> BatchNorm -> ReLU (+ residual) blocks for the ResNets.

Data runs channels-last (NHWC).  The convolution itself is cuDNN's implicit GEMM
through ``F.conv2d`` for now -- 1x1 convolutions on CUDA are routed to the
tcgen05 GEMM (an NHWC 1x1 conv *is* ``[N*H*W, Cin] @ W[Cout, Cin]^T``) -- and the
BatchNorm-apply + ReLU + residual-add tail is one fused pass
(``csrc/bn_act.cu``) instead of three ATen kernels.

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
    return F.conv2d(x, conv.weight, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups)


def conv_bn_act(x, conv: nn.Conv2d, bn: nn.BatchNorm2d, relu: bool = True, residual=None):
    y = _conv(x, conv)
    if bn.training:
        if bn.num_batches_tracked is not None:
            bn.num_batches_tracked.add_(1)
    y = F.batch_norm(y, bn.running_mean, bn.running_var, bn.weight, bn.bias, bn.training, bn.momentum or 0.0, bn.eps)
    if residual is not None:
        y = y + residual
    return F.relu(y) if relu else y
