"""HyperionDDP: data parallelism with an own bucketed gradient all-reduce.

What torch's ``DistributedDataParallel`` does for the reference (02_development/
distributed_utils.py:159, 229, 475 -- Reducer buckets of 25 MiB, first 1 MiB, fp32 NCCL
all-reduce, param broadcast at construction, BN-buffer broadcast each forward) is done here as:

* trainable parameters and their gradients are views of flat buffers (``FlatParams``) laid out in
  gradient-ready order; a bucket is a contiguous range of the flat gradient;
* ``register_post_accumulate_grad_hook`` counts a bucket down; when complete it is reduced on a side
  stream while backward continues:
    - **fused path** (CUDA, world > 1, symmetric memory): ``hyp_ddp_allreduce`` (csrc/comm.cu) --
      one kernel that reads every peer's bucket over NVLink (NVLS ``multimem.ld_reduce`` when a
      multicast mapping exists, P2P loads otherwise), multiplies by ``1/(world*loss_scale)``,
      checks finiteness (sets ``found_inf`` on every rank), and writes the averaged gradient back
      in the comm dtype -- AMP unscale + inf-check + cast + all-reduce in one pass (K-C1);
    - **baseline path** (gloo/CPU, or ``fused=False``): ``dist.all_reduce`` per bucket (async), the
      stock behaviour, kept as the comparison baseline and the CPU test path;
* a callback queued on the autograd engine joins the side stream at the end of backward.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from .flat import FlatParams
from .runtime import world


class HyperionDDP(nn.Module):
    def __init__(self, module: nn.Module, bucket_cap_mb: float = 25.0, first_bucket_mb: float = 1.0,
                 broadcast_buffers: bool = False, fused: bool | None = None, process_group=None,
                 loss_scaler=None, sync_params: bool = True):
        super().__init__()
        self.module = module
        self.rank, self.world_size = world()
        self.group = process_group
        # torch DDP re-broadcasts rank 0's BatchNorm statistics before every forward (K-C1c).  Here that is opt-in:
        # by default the buffers are synchronised when a checkpoint is taken (``sync_buffers()``), which keeps NCCL
        # out of the step (and out of a captured CUDA graph).
        self.broadcast_buffers = broadcast_buffers
        self.loss_scaler = loss_scaler
        p0 = next(p for p in module.parameters() if p.requires_grad)
        self.device = p0.device
        want_fused = fused if fused is not None else True
        self.fused = bool(want_fused and self.device.type == "cuda" and self.world_size > 1)
        self.symm = None
        allocator = None
        if self.fused:
            from . import symm

            if symm.supported():
                self.symm = symm.SymmetricArena(self.device, group=process_group)
                allocator = self.symm.alloc
            else:
                self.fused = False
        self.flat = FlatParams(module.named_parameters(), device=self.device, allocator=allocator, pad_to=max(1, self.world_size) * 64)
        if self.fused:
            self.symm.rendezvous()
        self.buckets = self.flat.buckets(int(bucket_cap_mb * 2 ** 20), int(first_bucket_mb * 2 ** 20))
        self._spec_bucket = {}
        for b, (_, _, idxs) in enumerate(self.buckets):
            for i in idxs:
                self._spec_bucket[i] = b
        self._pending = [0] * len(self.buckets)
        self._handles = []
        self._callback_queued = False
        self._require_sync = True
        self.comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        if sync_params and self.world_size > 1:
            self._broadcast_params()
        for i, s in enumerate(self.flat.specs):
            s.param.register_post_accumulate_grad_hook(self._make_hook(i))
        self._reset_pending()

    # ------------------------------------------------------------------ construction-time sync (K-C1b)
    def _broadcast_params(self) -> None:
        dist.broadcast(self.flat.data, src=0, group=self.group)
        for b in self.module.buffers():
            dist.broadcast(b, src=0, group=self.group)

    def _reset_pending(self) -> None:
        for b, (_, _, idxs) in enumerate(self.buckets):
            self._pending[b] = len(idxs)

    # ------------------------------------------------------------------ hooks
    def _make_hook(self, spec_idx: int):
        def hook(_param):
            if not self._require_sync or self.world_size == 1:
                return
            if not self._callback_queued:
                torch.autograd.Variable._execution_engine.queue_callback(self._finalize)
                self._callback_queued = True
            b = self._spec_bucket[spec_idx]
            self._pending[b] -= 1
            if self._pending[b] == 0:
                self._reduce_bucket(b)
        return hook

    def _reduce_bucket(self, b: int) -> None:
        start, end, _ = self.buckets[b]
        if self.fused:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                inv = self.loss_scaler.inv_scale if (self.loss_scaler is not None and self.loss_scaler.enabled) else None
                found = self.loss_scaler.found_inf if self.loss_scaler is not None else None
                self.symm.allreduce_avg_(self.flat.grad, start, end, inv_scale=inv, found_inf=found)
        else:
            view = self.flat.grad[start:end]
            h = dist.all_reduce(view, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._handles.append((h, view))

    def _finalize(self) -> None:
        """Runs once at the end of backward: join the communication with the compute stream.

        A bucket whose counter did not reach zero holds a parameter that received no gradient this step (unused branch).
        The model is the same on every rank, so the same buckets are short everywhere: they are reduced here anyway (the
        missing slots are zeros from ``zero_grad``), which keeps the ranks' collective sequences identical instead of
        silently skipping the bucket (torch DDP raises in that situation; see ``unused`` for the names)."""
        self._callback_queued = False
        self.unused = []
        for b, (_, _, idxs) in enumerate(self.buckets):
            if 0 < self._pending[b] < len(idxs) or (self._pending[b] == len(idxs) and self._require_sync):
                self.unused += [self.flat.specs[i].name for i in idxs if self.flat.specs[i].param.grad is None]
                self._reduce_bucket(b)
        if self.fused:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        else:
            for h, view in self._handles:
                h.wait()
                view.div_(self.world_size)
            self._handles.clear()
        self._reset_pending()

    # ------------------------------------------------------------------ module API
    def forward(self, *args, **kwargs):
        if self.broadcast_buffers and self.world_size > 1 and self.training:
            self._sync_buffers()
        return self.module(*args, **kwargs)

    def sync_buffers(self) -> None:
        """Collective: make every rank's buffers equal to rank 0's (call before saving)."""
        if self.world_size > 1:
            self._sync_buffers()

    def _sync_buffers(self) -> None:
        """K-C1c: rank 0's BatchNorm statistics win, as with torch DDP ``broadcast_buffers=True``."""
        bufs = [b for b in self.module.buffers() if b.numel() > 0]
        if not bufs:
            return
        floats = [b for b in bufs if b.is_floating_point()]
        if floats:
            flat = torch.cat([b.reshape(-1).float() for b in floats])
            dist.broadcast(flat, src=0, group=self.group)
            off = 0
            for b in floats:
                b.copy_(flat[off:off + b.numel()].view_as(b).to(b.dtype))
                off += b.numel()
        for b in bufs:
            if not b.is_floating_point():
                dist.broadcast(b, src=0, group=self.group)

    class _NoSync:
        def __init__(self, ddp):
            self.ddp = ddp

        def __enter__(self):
            self.prev = self.ddp._require_sync
            self.ddp._require_sync = False

        def __exit__(self, *a):
            self.ddp._require_sync = self.prev

    def no_sync(self):
        """Gradient accumulation: skip the all-reduce inside this context."""
        return HyperionDDP._NoSync(self)

    def state_dict(self, *args, **kwargs):
        """Keys carry no ``module.`` prefix -- the reference saves ``model.module.state_dict()``."""
        return self.module.state_dict(*args, **kwargs)

    def load_state_dict(self, sd, *args, **kwargs):
        return self.module.load_state_dict(sd, *args, **kwargs)

    def loss(self, *args, **kwargs):
        """Forward through ``module.loss`` (fused LM-head + cross-entropy) with DDP's buffer sync."""
        if self.broadcast_buffers and self.world_size > 1 and self.training:
            self._sync_buffers()
        return self.module.loss(*args, **kwargs)
