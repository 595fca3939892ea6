"""FusedAdamW: AdamW over :class:`FlatParams` in one kernel launch (``csrc/optim.cu``).

* fp32 master weights + exp_avg / exp_avg_sq live here; the module's parameters are the 16-bit
  compute copy, rewritten by the same kernel (so AMP needs no per-op weight casts).
* ``max_norm`` (global-norm clipping), ``inv_scale`` (AMP unscale) and ``found_inf`` (skip on
  overflow) are device scalars read by the kernel -- the step never syncs the host.
* Without the native library (CPU tests) the identical update runs as torch ops.

Reference: ``torch.optim.AdamW(model.parameters(), lr=...)`` + ``GradScaler.step`` +
``clip_grad_norm_`` (02_development/distributed_utils.py:161, 179, 231, 334, 351-352, 503, 522-523).
"""
from __future__ import annotations

import torch

from .ops import _lib, native
from .parallel.flat import FlatParams


class FusedAdamW:
    def __init__(self, flat: FlatParams, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, max_norm: float = 0.0, shard: tuple[int, int] | None = None):
        """``shard=(start, end)``: keep optimizer state only for that element range (FSDP)."""
        self.flat = flat
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = lr, betas, eps, weight_decay, max_norm
        self.step_count = 0
        self.start, self.end = shard if shard is not None else (0, flat.numel)
        n = self.end - self.start
        dev = flat.device
        self.master = flat.data[self.start:self.end].to(torch.float32).clone() if flat.dtype != torch.float32 else None
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        self._norm_buf = torch.zeros(2, device=dev, dtype=torch.float32)
        self.step_t = torch.zeros(1, device=dev, dtype=torch.float32)  # device-side step count (CUDA-graph safe)

    # -- norms ---------------------------------------------------------------------------------
    def grad_sumsq(self, grad: torch.Tensor | None = None) -> torch.Tensor:
        """fp32[2] = (sum of squares, non-finite flag) of the (local) gradient, computed on device."""
        g = grad if grad is not None else self.flat.grad[self.start:self.end]
        self._norm_buf.zero_()
        if g.is_cuda and _lib.available():
            native.sumsq(g, self._norm_buf)
        else:
            gf = g.float()
            self._norm_buf[0] = (gf * gf).sum()
            self._norm_buf[1] = (~torch.isfinite(gf)).any().float()
        return self._norm_buf

    # -- step ----------------------------------------------------------------------------------
    @torch.no_grad()
    def step(self, grad: torch.Tensor | None = None, sumsq: torch.Tensor | None = None,
             inv_scale: torch.Tensor | None = None, found_inf: torch.Tensor | None = None) -> None:
        self.step_count += 1
        self.step_t += 1
        g = grad if grad is not None else self.flat.grad[self.start:self.end]
        p16 = self.flat.data[self.start:self.end]
        b1, b2 = self.betas
        if self.max_norm > 0 and sumsq is None:
            sumsq = self.grad_sumsq(g)
        if g.is_cuda and _lib.available():
            if self.master is not None:
                native.adamw_step(self.master, g, self.exp_avg, self.exp_avg_sq, p16, lr=self.lr, beta1=b1, beta2=b2,
                                  eps=self.eps, weight_decay=self.weight_decay, step=self.step_count,
                                  max_norm=self.max_norm, sumsq=sumsq, inv_scale=inv_scale, found_inf=found_inf,
                                  step_t=self.step_t)
            else:
                native.adamw_step(p16, g, self.exp_avg, self.exp_avg_sq, None, lr=self.lr, beta1=b1, beta2=b2,
                                  eps=self.eps, weight_decay=self.weight_decay, step=self.step_count,
                                  max_norm=self.max_norm, sumsq=sumsq, inv_scale=inv_scale, found_inf=found_inf,
                                  step_t=self.step_t)
            return
        # ---- torch formulation of the same update (CPU / oracle) ----
        if found_inf is not None and float(found_inf.reshape(-1)[0]) != 0.0:
            return
        gf = g.float()
        scale = float(inv_scale.reshape(-1)[0]) if inv_scale is not None else 1.0
        if self.max_norm > 0:
            norm = float(sumsq[0]) ** 0.5 * scale
            coef = self.max_norm / (norm + 1e-6)
            if coef < 1.0:
                scale *= coef
        gf = gf * scale
        p = self.master if self.master is not None else p16
        self.exp_avg.mul_(b1).add_(gf, alpha=1 - b1)
        self.exp_avg_sq.mul_(b2).addcmul_(gf, gf, value=1 - b2)
        c1 = 1 - b1 ** self.step_count
        c2 = 1 - b2 ** self.step_count
        denom = (self.exp_avg_sq.sqrt() / (c2 ** 0.5)).add_(self.eps)
        p.mul_(1 - self.lr * self.weight_decay).addcdiv_(self.exp_avg, denom, value=-self.lr / c1)
        if self.master is not None:
            p16.copy_(p.to(p16.dtype))

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.flat.zero_grad()

    # -- checkpointing ---------------------------------------------------------------------------
    @torch.no_grad()
    def refresh_master(self, sd: dict | None = None) -> None:
        """Re-seed the fp32 master from the module's (just loaded) parameters; fp32 values in ``sd`` (an unwrapped
        state_dict) overlay it so nothing is lost through the 16-bit copy.  Call after ANY weights-only load: otherwise
        the next step writes the stale master back over the loaded weights."""
        if self.master is None:
            return
        self.master.copy_(self.flat.data[self.start:self.end].to(torch.float32))
        if sd is not None and self.start == 0 and self.end == self.flat.numel:
            from .parallel.flat import shaped_view

            for sp in self.flat.specs:
                v = sd.get(sp.name)
                if v is not None and tuple(v.shape) == tuple(sp.param.shape):
                    shaped_view(self.master, sp).copy_(v.to(device=self.master.device, dtype=torch.float32))

    def master_state_overlay(self) -> dict:
        """{parameter name: fp32 master view} for checkpoint export (true fp32 weights, as the reference's AMP keeps)."""
        if self.master is None or self.start != 0 or self.end != self.flat.numel:
            return {}
        from .parallel.flat import shaped_view

        return {sp.name: shaped_view(self.master, sp) for sp in self.flat.specs}

    def state_dict(self) -> dict:
        return {"step": int(self.step_t.item()), "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "master": self.master,
                "lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay}

    def load_state_dict(self, sd: dict) -> None:
        self.step_count = int(sd["step"])
        self.step_t.fill_(float(self.step_count))
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        if self.master is not None and sd.get("master") is not None:
            self.master.copy_(sd["master"])
            self.flat.data[self.start:self.end].copy_(self.master.to(self.flat.dtype))
