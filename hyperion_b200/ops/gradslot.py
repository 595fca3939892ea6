"""Gradient slots: let a backward kernel write a parameter's gradient straight into its place in the flat
gradient buffer (``parallel.flat.FlatParams``).

Without this, every parameter costs an extra elementwise pass per step: the op returns a fresh dW tensor and
autograd's AccumulateGrad adds it into the ``.grad`` view of the flat buffer (read dW + read-modify-write grad).
With it, ``FlatParams.zero_grad`` leaves ``param.grad = None``; the op asks :func:`claim` for the parameter's slot,
makes the wgrad GEMM / column-sum / norm-reduce kernel store there, and returns that view -- AccumulateGrad then
adopts the returned tensor as ``param.grad`` without touching memory (it steals a gradient nobody else references
when ``.grad`` is undefined).  A post-accumulate hook installed by ``FlatParams`` checks the result: whatever ended
up in ``param.grad`` that is *not* the slot (an op that does not know about slots, a second use of a tied
parameter that made autograd sum out of place) is copied into the slot and ``.grad`` re-pointed, so the flat
buffer is always the truth the optimizer / collectives read.

A slot can be claimed once per zero_grad (tied weights: the second use returns an ordinary tensor, autograd
adds); with gradient accumulation (``.grad`` already defined) nothing is claimed and autograd accumulates in
place into the slot as before.

The reference has no counterpart (it trains through torch.optim.AdamW on per-parameter tensors).
"""
from __future__ import annotations

import torch

ENABLED = True


class GradSlot:
    __slots__ = ("flat", "spec", "claimed")

    def __init__(self, flat, spec):
        self.flat, self.spec, self.claimed = flat, spec, False

    def view(self) -> torch.Tensor:
        from ..parallel.flat import shaped_view

        return shaped_view(self.flat.grad, self.spec)


def slot_of(param) -> GradSlot | None:
    """The parameter's slot (attached by FlatParams), or None."""
    return getattr(param, "_hyp_slot", None)


def claim(slot: GradSlot | None, dtype: torch.dtype | None = None) -> torch.Tensor | None:
    """A fresh view of the slot's flat-gradient range if a kernel may write this step's gradient there, else None."""
    if slot is None or not ENABLED or slot.claimed or slot.spec.param.grad is not None:
        return None
    if dtype is not None and slot.flat.grad.dtype != dtype:
        return None
    slot.claimed = True
    return slot.view()
