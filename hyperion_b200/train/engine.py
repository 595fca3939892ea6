"""Training engine shared by every trainer: precision policy + parallel wrapper + fused optimizer.

    eng = build_engine(model, strategy="ddp", precision="bf16", lr=2e-4, device=dev)
    loss = eng.model.loss(x, y); eng.backward_and_step(loss)

One ``backward_and_step`` = scale loss -> backward (communication kernels fire from autograd hooks and
overlap) -> fused AdamW (unscale / clip / skip-on-overflow read from device scalars) -> scaler update ->
zero the flat gradient.  Nothing in it synchronises the host.
"""
from __future__ import annotations

import contextlib

import torch
import torch.nn as nn

from ..amp import DynamicLossScaler, policy
from ..optim import FusedAdamW
from ..parallel.ddp import HyperionDDP
from ..parallel.flat import FlatParams
from ..parallel.fsdp import HyperionFSDP
from ..parallel.runtime import world
from ..utils import nvtx


def cast_trainable(model: nn.Module, dtype: torch.dtype) -> nn.Module:
    """Move floating-point parameters and buffers to the compute dtype (BatchNorm statistics stay fp32)."""
    if dtype == torch.float32:
        return model
    for m in model.modules():
        for name, p in m.named_parameters(recurse=False):
            if p.is_floating_point():
                p.data = p.data.to(dtype)
        for name, b in m.named_buffers(recurse=False):
            if b.is_floating_point() and not isinstance(m, nn.modules.batchnorm._BatchNorm):
                setattr(m, name, b.to(dtype))
    return model


class Engine:
    def __init__(self, model, wrapped, flat, optimizer, scaler, strategy: str, precision, clip: float):
        self.raw_model, self.model, self.flat, self.optimizer, self.scaler = model, wrapped, flat, optimizer, scaler
        self.strategy, self.precision, self.clip = strategy, precision, clip
        self.rank, self.world_size = world()

    def backward_and_step(self, loss: torch.Tensor) -> None:
        scaler = self.scaler
        with nvtx.range("hyperion/backward+comm"):   # reduce-scatter / all-reduce kernels fire from autograd hooks in here
            (scaler.scale_loss(loss) if scaler is not None else loss).backward()
        from ..ops import rng
        from ..ops.conv import flush_bn_counters

        flush_bn_counters()
        if self.strategy == "fsdp":
            with nvtx.range("hyperion/optimizer(fsdp)"):
                self.model.step()
            rng.advance(self.flat.device)   # next step draws new dropout masks (device-side: replays of a graph advance too)
            return
        with nvtx.range("hyperion/optimizer"):
            self._optimizer_phase(scaler)
        rng.advance(self.flat.device)

    def _optimizer_phase(self, scaler) -> None:
        inv_scale = found_inf = sumsq = None
        fused_ddp = self.strategy == "ddp" and getattr(self.model, "fused", False)
        if scaler is not None and scaler.enabled:
            if fused_ddp:
                found_inf = self.model.symm.found_inf  # raised on every rank by the all-reduce kernel; grads already unscaled
                scaler.found_inf = found_inf
            else:
                sumsq = self.optimizer.grad_sumsq()
                scaler.note_overflow(sumsq)
                inv_scale, found_inf = scaler.inv_scale, scaler.found_inf
        if self.clip > 0 and sumsq is None:
            sumsq = self.optimizer.grad_sumsq()
        self.optimizer.step(sumsq=sumsq if self.clip > 0 else None, inv_scale=inv_scale, found_inf=found_inf)
        if scaler is not None:
            scaler.update()
        self.optimizer.zero_grad()
        symm = getattr(self.model, "symm", None)
        if symm is not None:
            symm.end_step(getattr(self.model, "comm_stream", None))

    def state_for_checkpoint(self):
        return self.optimizer if self.strategy != "fsdp" else None


def build_engine(model: nn.Module, *, strategy: str = "single", precision: str = "bf16", lr: float = 1e-3,
                 betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01, clip: float = 0.0, device=None,
                 fused_comm: bool | None = None, bucket_cap_mb: float = 25.0, fsdp_units=None,
                 fsdp_min_num_params: int = 100_000, broadcast_buffers: bool = False,
                 reshard_after_forward: bool | None = None) -> Engine:
    rank, W = world()
    pol = policy(precision)
    device = torch.device(device) if device is not None else next(model.parameters()).device
    if device.type != "cuda" and precision == "fp16" and not torch.cuda.is_available():
        pol = policy("fp32")          # no fp16 arithmetic worth the name on a CPU-only machine
    model = model.to(device)
    cast_trainable(model, pol.dtype)
    scaler = DynamicLossScaler(device, enabled=True) if pol.needs_scaler else None
    if strategy == "fsdp":
        wrapped = HyperionFSDP(model, units=fsdp_units, min_num_params=fsdp_min_num_params, lr=lr, betas=betas, eps=eps,
                               weight_decay=weight_decay, max_norm=clip, fused=fused_comm, loss_scaler=scaler,
                               reshard_after_forward=reshard_after_forward)
        return Engine(model, wrapped, wrapped.flat, None, scaler, strategy, pol, clip)
    if strategy == "ddp" and W > 1:
        wrapped = HyperionDDP(model, bucket_cap_mb=bucket_cap_mb, fused=fused_comm, loss_scaler=scaler,
                              broadcast_buffers=broadcast_buffers)
        flat = wrapped.flat
    else:
        wrapped = model
        flat = FlatParams(model.named_parameters(), device=device)
        strategy = "single" if strategy != "ddp" else "ddp"
    opt = FusedAdamW(flat, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, max_norm=clip)
    return Engine(model, wrapped, flat, opt, scaler, strategy, pol, clip)


class GraphedStep:
    """One whole training step (forward, backward, communication kernels, optimizer, scaler update) captured in a
    CUDA graph and replayed with a single launch.

    A GPT-2-medium step is ~1000 kernel launches; enqueueing them from Python takes about as long as the GPU needs
    to run them, so the step is launch-bound once the kernels are fast.  Capturing removes that: every launch made
    through the C ABI lands on the capturing stream, tensors come from the graph's private pool (fixed addresses, so
    the TMA descriptors baked into the launches stay valid), and everything that varies per step lives in device
    memory (AdamW step count, loss scale, found_inf, communication epochs).

        step = GraphedStep(eng, lambda batch: eng.model.loss(batch[0]), example_batch)
        loss = step(batch)            # copies the batch into the static input, replays, returns the static loss tensor

    ``warmup`` eager steps run first (they are real optimizer steps).  Shapes are fixed by ``example_batch``.
    ``StepRunner`` (train/common.py) is the wrapper the trainers use: it captures lazily, counts the warm-up steps as
    training steps on their own batches and falls back to eager for a ragged last batch.
    """

    def __init__(self, eng: Engine, loss_fn, example_batch, warmup: int = 2):
        assert torch.cuda.is_available()
        self.eng = eng
        self.static_batch = tuple(t.clone() for t in example_batch)
        dev = self.static_batch[0].device
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side) if warmup > 0 else contextlib.nullcontext():
            for _ in range(warmup):
                out = loss_fn(self.static_batch)
                eng.backward_and_step(out[0] if isinstance(out, (tuple, list)) else out)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        # events recorded by the eager warm-up steps (pending parameter gathers) are complete after the synchronize;
        # waiting on them from the capturing stream would be a dependency on uncaptured work
        if hasattr(eng.model, "_gather_events"):
            eng.model._gather_events = [None] * len(eng.model._gather_events)
        from ..ops import _lib

        n0 = _lib.launch_count()
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph, capture_error_mode="thread_local"):  # other threads (NCCL watchdog) may poll events
            out = loss_fn(self.static_batch)
            # ``loss_fn`` may return ``loss`` or ``(loss, *extras)`` (e.g. a correct-prediction count): the extras are
            # static graph outputs too and come back from every replay
            self.loss = out[0] if isinstance(out, (tuple, list)) else out
            eng.backward_and_step(self.loss)
        if isinstance(out, (tuple, list)):
            self.loss = tuple(t.detach() for t in out)
        else:
            self.loss = self.loss.detach()
        self.kernels_per_replay = _lib.launch_count() - n0  # library kernels recorded in the graph
        self._add_launches = _lib.add_launches

    def __call__(self, batch):
        for dst, src in zip(self.static_batch, batch):
            dst.copy_(src, non_blocking=True)
        self.graph.replay()
        self._add_launches(self.kernels_per_replay)
        return self.loss
