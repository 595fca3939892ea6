"""Mixed-precision policies.

The reference uses ``torch.cuda.amp.autocast()`` (fp16) + ``GradScaler`` in the DDP trainers
(02_development/distributed_utils.py:163, 175-180) and FSDP ``MixedPrecision(bf16 x3)``
(:320-324).  Here a policy converts the *module* to the 16-bit compute dtype once; fp32 master
weights live in :class:`hyperion_b200.optim.FusedAdamW`, and norm / softmax / loss math is fp32
inside the kernels, so no per-op autocast casts remain on the hot path.

``DynamicLossScaler`` is the GradScaler equivalent with all state on the device: ``scale``,
``growth_tracker`` and ``found_inf`` are tensors; unscale + inf check happen inside the gradient
all-reduce kernel (multi-GPU) or the AdamW kernel (1 GPU); ``update()`` is one one-thread kernel (``hyp_scaler_update``).
``state_dict`` uses GradScaler's keys so ``{"model_state_dict", "scaler_state_dict"}`` checkpoints
(02_development/mixed_precision.ipynb c5 L7-15) keep their layout.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass
class Precision:
    name: str
    dtype: torch.dtype
    needs_scaler: bool


POLICIES = {
    "fp32": Precision("fp32", torch.float32, False),
    "fp16": Precision("fp16", torch.float16, True),
    "bf16": Precision("bf16", torch.bfloat16, False),
}


def policy(name: str) -> Precision:
    return POLICIES[name]


class DynamicLossScaler:
    def __init__(self, device, init_scale: float = 65536.0, growth_factor: float = 2.0, backoff_factor: float = 0.5,
                 growth_interval: int = 2000, enabled: bool = True):
        self.enabled = enabled
        self.growth_factor, self.backoff_factor, self.growth_interval = growth_factor, backoff_factor, growth_interval
        self.scale = torch.full((1,), init_scale if enabled else 1.0, device=device, dtype=torch.float32)
        self.inv_scale = self.scale.reciprocal()
        self.growth_tracker = torch.zeros(1, device=device, dtype=torch.float32)
        self.found_inf = torch.zeros(1, device=device, dtype=torch.float32)

    def scale_loss(self, loss: torch.Tensor) -> torch.Tensor:
        return loss * self.scale.to(loss.dtype).reshape(()) if self.enabled else loss

    def _native(self) -> bool:
        from .ops import _lib

        return self.scale.is_cuda and _lib.available() and hasattr(_lib.raw(), "hyp_scaler_update")

    def note_overflow(self, sumsq_flag: torch.Tensor) -> None:
        """``sumsq_flag``: fp32[2] from the norm kernel; element 1 (or a non-finite sum) marks an overflow."""
        if self._native() and sumsq_flag.dtype == torch.float32 and sumsq_flag.is_contiguous():
            from .ops import _lib

            _lib.call("hyp_scaler_note", _lib.ptr(self.found_inf), _lib.ptr(sumsq_flag), _lib.stream_ptr(self.scale.device))
            return
        bad = (sumsq_flag[1] != 0) | ~torch.isfinite(sumsq_flag[0])
        self.found_inf.copy_(torch.maximum(self.found_inf, bad.float().reshape(1)))

    @torch.no_grad()
    def update(self) -> None:
        """GradScaler.update(): back off on overflow, grow after ``growth_interval`` clean steps (device-side)."""
        if not self.enabled:
            return
        if self._native():
            from .ops import _lib

            _lib.call("hyp_scaler_update", _lib.ptr(self.scale), _lib.ptr(self.inv_scale), _lib.ptr(self.growth_tracker),
                      _lib.ptr(self.found_inf), float(self.growth_factor), float(self.backoff_factor),
                      float(self.growth_interval), _lib.stream_ptr(self.scale.device))
            return
        overflow = self.found_inf != 0
        grown = self.growth_tracker + 1
        do_grow = (~overflow) & (grown >= self.growth_interval)
        self.scale.copy_(torch.where(overflow, self.scale * self.backoff_factor,
                                     torch.where(do_grow, self.scale * self.growth_factor, self.scale)))
        self.growth_tracker.copy_(torch.where(overflow | do_grow, torch.zeros_like(grown), grown))
        self.inv_scale.copy_(self.scale.reciprocal())
        self.found_inf.zero_()

    def state_dict(self) -> dict:
        return {"scale": float(self.scale.item()), "growth_factor": self.growth_factor,
                "backoff_factor": self.backoff_factor, "growth_interval": self.growth_interval,
                "_growth_tracker": int(self.growth_tracker.item())}

    def load_state_dict(self, sd: dict) -> None:
        self.scale.fill_(sd["scale"])
        self.inv_scale.copy_(self.scale.reciprocal())
        self.growth_tracker.fill_(sd.get("_growth_tracker", 0))
