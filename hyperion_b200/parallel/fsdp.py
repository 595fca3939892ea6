"""HyperionFSDP: sharded data parallelism (ZeRO-3 style state sharding) with own communication kernels.

What FSDP1 ``FULL_SHARD`` + ``MixedPrecision(bf16 x3)`` does for the reference (02_development/
distributed_utils.py:318-332, 488-499) is re-designed for a GPU with 180 GB of HBM3e:

* every trainable parameter lives in one flat 16-bit buffer; the buffer is cut into *units* (one per wrapped
  module: size-based policy ``min_num_params=100_000`` or one unit per decoder layer, like the reference);
* what is sharded 1/W per rank: the fp32 master weights, AdamW ``exp_avg`` / ``exp_avg_sq`` and the reduced
  gradient -- 12 of the 16 bytes/parameter of mixed-precision Adam state.  The 16-bit compute copy of a unit
  stays resident after it is gathered (it is 2 bytes/parameter; 13.5 GB for Llama-2-7B), so parameters are
  gathered **once per step** instead of twice (FSDP1 re-gathers in backward);
* gradient path (K-C2): when a unit's gradients are complete, one kernel on the comm stream makes rank r pull
  slice r of that unit's gradient from every peer over NVLink (NVLS ``multimem.ld_reduce`` when available),
  average in fp32 and -- without clipping -- apply AdamW to the fp32 master shard and emit the new 16-bit shard
  in the same pass.  With global-norm clipping (the reference clips at 1.0) the kernel emits the fp32 reduced
  shard and its sum of squares instead; after one scalar all-reduce a single AdamW launch finishes the step
  (two-phase, exact global norm -- the reference computes a per-rank local norm);
* parameter path (K-C3): after the optimizer step the all-gathers of all units are queued on the comm stream
  in forward order; each unit's forward waits only for its own event, so gathers overlap earlier compute;
* world size 1 degrades to NO_SHARD (no communication kernels at all), as FSDP1 does;
* gloo / CPU: the same algorithm with ``dist.all_reduce`` / ``dist.all_gather_into_tensor`` (test path).
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import _lib, native
from .flat import FlatParams
from .runtime import world


def size_based_units(module: nn.Module, min_num_params: int = 100_000) -> list[nn.Module]:
    """Innermost modules holding >= ``min_num_params`` parameters (mirrors size_based_auto_wrap_policy)."""
    units = []

    def visit(m: nn.Module) -> int:
        total = sum(p.numel() for p in m.parameters(recurse=False))
        for c in m.children():
            total += visit(c)
        if m is not module and total >= min_num_params:
            units.append(m)
            return 0
        return total

    visit(module)
    return units


def class_based_units(module: nn.Module, cls) -> list[nn.Module]:
    return [m for m in module.modules() if isinstance(m, cls)]


class _AgHandle:
    """Attached to the first-consumed weight of a GEMM-first unit (``weight._hyp_ag``): while ``pending`` the
    unit's parameters are stale locally, and the next GEMM on that weight gathers the whole unit from the peers'
    shards *inside* the GEMM kernel (K-C3, ``hyp_gemm_ag``)."""

    def __init__(self, fsdp: "HyperionFSDP", ui: int):
        self.fsdp, self.ui = fsdp, ui
        self.pending = False
        self.arrived = torch.zeros(fsdp.world_size, device=fsdp.device, dtype=torch.int32)
        self.done = torch.zeros(1, device=fsdp.device, dtype=torch.int32)  # chunks per slice gathered by past launches

    def run(self, x2, w2, out, bias=None, aux=None, act=None, preact_out=None) -> None:
        f = self.fsdp
        s, e, _ = f.units[self.ui]
        a, b = f._shard_range(self.ui)
        nch = ((b - a) * f.shard16.element_size() + 32767) // 32768
        f.arena.gemm_allgather(x2, w2, out, shard_buf=f.shard16, shard_start=a, per_elems=b - a,
                               unit_dst=f.flat.data[s:e], arrived=self.arrived, target=nch, target_base=self.done,
                               bias=bias, aux=aux, act=act, preact_out=preact_out)
        self.done += nch  # device-side, so a captured graph replays with advancing targets
        self.pending = False


class HyperionFSDP(nn.Module):
    def __init__(self, module: nn.Module, units: list[nn.Module] | None = None, min_num_params: int = 100_000,
                 lr: float = 1e-4, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.01,
                 max_norm: float = 0.0, fused: bool | None = None, process_group=None, loss_scaler=None,
                 sync_params: bool = True, fuse_allgather_gemm: bool = True, reshard_after_forward: bool | None = None):
        super().__init__()
        self.reshard = False
        self.module = module
        self.rank, self.world_size = world()
        self.group = process_group
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = lr, betas, eps, weight_decay, max_norm
        self.loss_scaler = loss_scaler
        p0 = next(p for p in module.parameters() if p.requires_grad)
        self.device = p0.device
        W = self.world_size
        want_fused = fused if fused is not None else True
        self.fused = bool(want_fused and self.device.type == "cuda" and W > 1)
        self.arena = None
        if self.fused:
            from . import symm

            if symm.supported():
                self.arena = symm.SymmetricArena(self.device, group=process_group)
            else:
                self.fused = False
        # ---- units
        unit_modules = units if units is not None else size_based_units(module, min_num_params)
        owner: dict[str, int] = {}
        for ui, um in enumerate(unit_modules):
            prefix = next((n for n, m in module.named_modules() if m is um), None)
            for pn, _ in um.named_parameters():
                owner[f"{prefix}.{pn}" if prefix else pn] = ui
        self.unit_modules = unit_modules
        galloc = self.arena.alloc if self.fused else None
        self.flat = FlatParams(module.named_parameters(), device=self.device, unit_key=lambda n: owner.get(n, -1),
                               unit_align=W * 128, grad_allocator=galloc)
        self.units = [(s, e, k) for s, e, k in self.flat.units]
        N = self.flat.numel
        self.shard_numel = N // W
        dt = self.flat.dtype
        if sync_params and W > 1:
            dist.broadcast(self.flat.data, src=0, group=self.group)
            for b in module.buffers():
                dist.broadcast(b, src=0, group=self.group)
        # ---- sharded state
        alloc16 = self.arena.alloc if self.fused else (lambda n, d, dev: torch.zeros(n, dtype=d, device=dev))
        self.shard16 = alloc16(self.shard_numel, dt, self.device)
        if self.fused:
            self.arena.rendezvous()
        self._copy_local_shards(self.flat.data, self.shard16)
        self.master = self.shard16.to(torch.float32).clone() if (W > 1 or dt != torch.float32) else None
        self.exp_avg = torch.zeros(self.shard_numel, device=self.device, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(self.shard_numel, device=self.device, dtype=torch.float32)
        self.gshard = torch.zeros(self.shard_numel, device=self.device, dtype=torch.float32) if (W > 1) else None
        # (sum of squares, non-finite flag); lives in symmetric memory when fused so the global-norm reduction is
        # one tiny launch of the all-reduce kernel instead of an NCCL call (K-C4)
        self.sumsq = (self.arena.alloc(4, torch.float32, self.device) if self.fused else
                      torch.zeros(4, device=self.device, dtype=torch.float32))
        self.step_count = 0
        self.step_t = torch.ones(1, device=self.device, dtype=torch.float32)  # device-side number of the step in flight
        # ---- schedule
        self.comm_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None
        self._spec_unit = {}
        for ui, (s, e, _) in enumerate(self.units):
            for i, sp in enumerate(self.flat.specs):
                if s <= sp.offset < e:
                    self._spec_unit[i] = ui
        self._unit_nparams = [0] * len(self.units)
        for i, ui in self._spec_unit.items():
            self._unit_nparams[ui] += 1
        self._pending = list(self._unit_nparams)
        self._gather_events: list = [None] * len(self.units)
        self._callback_queued = False
        self._require_sync = True
        self._fired: set = set()
        self._learned = False
        self._key_to_units: dict[int, list[int]] = {}
        for ui, (_, _, k) in enumerate(self.units):
            self._key_to_units.setdefault(k, []).append(ui)
        for i, sp in enumerate(self.flat.specs):
            sp.param.register_post_accumulate_grad_hook(self._make_grad_hook(i))
        for k, um in enumerate(unit_modules):
            um.register_forward_pre_hook(self._make_fwd_hook(k))
        # ---- K-C3: units whose first consumer is a GEMM on a known weight gather inside that GEMM
        self._ag_handles: dict[int, _AgHandle] = {}
        import os

        # HYPERION_FSDP_AG=plain gathers every unit with the stand-alone kernel (queued at the start of the forward, so it
        # overlaps the forward just the same) and leaves the GEMMs on the faster CTA-pair kernel
        fuse_allgather_gemm = fuse_allgather_gemm and os.environ.get("HYPERION_FSDP_AG", "fused") == "fused"
        self._gathers_due = False
        if self.fused and fuse_allgather_gemm and self.flat.dtype in (torch.bfloat16, torch.float16):
            from ..models.transformer import MultiheadSelfAttention

            flat_ids = {id(sp.param) for sp in self.flat.specs}
            for k, um in enumerate(unit_modules):
                target = None
                if isinstance(um, nn.Linear):
                    target = um.weight
                elif isinstance(um, MultiheadSelfAttention):
                    target = um.in_proj_weight
                uis = self._key_to_units.get(k, [])
                if target is not None and id(target) in flat_ids and len(uis) == 1 and target.shape[1] % 8 == 0:
                    h = _AgHandle(self, uis[0])
                    target._hyp_ag = h
                    self._ag_handles[uis[0]] = h

    # ------------------------------------------------------------------ shard <-> flat index helpers
    def _shard_range(self, ui: int) -> tuple[int, int]:
        s, e, _ = self.units[ui]
        W = self.world_size
        return s // W, e // W

    def _copy_local_shards(self, flat_src: torch.Tensor, shard_dst: torch.Tensor) -> None:
        W, r = self.world_size, self.rank
        for ui, (s, e, _) in enumerate(self.units):
            per = (e - s) // W
            a, b = self._shard_range(ui)
            shard_dst[a:b].copy_(flat_src[s + r * per: s + (r + 1) * per].to(shard_dst.dtype))

    # ------------------------------------------------------------------ forward: wait for gathered units
    def _make_fwd_hook(self, key: int):
        def hook(_m, _args):
            self._fired.add(key)
            if self.comm_stream is None:
                return
            for ui in self._key_to_units.get(key, ()):
                ev = self._gather_events[ui]
                if ev is not None:
                    torch.cuda.current_stream(self.device).wait_event(ev)
        return hook

    def _wait_units_at_start(self) -> None:
        # Root parameters (embeddings, head, final norm) and units whose module object is never *called* (the
        # models read e.g. ``linear1.weight`` directly inside a fused op) cannot rely on a forward-pre hook:
        # they are waited for here.  Which units those are is learned during the first forward, when every
        # parameter is still resident from construction and no gather is pending.
        keys = [-1] + ([k for k in range(len(self.unit_modules)) if k not in self._fired] if self._learned else
                       list(range(len(self.unit_modules))))
        cur = torch.cuda.current_stream(self.device)
        for k in keys:
            for ui in self._key_to_units.get(k, ()):
                ev = self._gather_events[ui]
                if ev is not None:
                    cur.wait_event(ev)

    def _start_of_forward(self) -> None:
        # the parameter gathers of the previous optimizer step are issued here, in forward order on the communication
        # stream, so they run under this forward (each unit's first use waits for its own event); inside a captured step
        # this keeps them in the same graph as their consumers instead of serialising them at the end of the step
        if self._gathers_due:
            self._gathers_due = False
            self._queue_gathers()

    def forward(self, *args, **kwargs):
        self._start_of_forward()
        if self.comm_stream is not None:
            self._wait_units_at_start()
        out = self.module(*args, **kwargs)
        self._learned = True
        return out

    def loss(self, *args, **kwargs):
        """Forward through ``module.loss`` (fused LM-head + CE) with the same gather waits as ``forward``."""
        self._start_of_forward()
        if self.comm_stream is not None:
            self._wait_units_at_start()
        out = self.module.loss(*args, **kwargs)
        self._learned = True
        return out

    # ------------------------------------------------------------------ backward: per-unit reduce-scatter
    def _make_grad_hook(self, spec_idx: int):
        def hook(_p):
            if self.world_size == 1 or not self._require_sync:
                return
            if not self._callback_queued:
                torch.autograd.Variable._execution_engine.queue_callback(self._end_of_backward)
                self._callback_queued = True
            ui = self._spec_unit[spec_idx]
            self._pending[ui] -= 1
            if self._pending[ui] == 0:
                self._reduce_unit(ui)
        return hook

    def _adam_args(self, a: int, b: int) -> dict:
        return dict(master=self.master[a:b], m=self.exp_avg[a:b], v=self.exp_avg_sq[a:b], p16=self.shard16[a:b],
                    lr=self.lr, betas=self.betas, eps=self.eps, weight_decay=self.weight_decay, step=self.step_count + 1,
                    step_t=self.step_t)

    def _unscale(self) -> float:
        return 1.0 / self.world_size

    def _reduce_unit(self, ui: int) -> None:
        s, e, _ = self.units[ui]
        W = self.world_size
        per = (e - s) // W
        a, b = self._shard_range(ui)
        if self.fused:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                if self.max_norm > 0 or self.loss_scaler is not None:
                    self.arena.reduce_scatter(self.flat.grad, s, per, gshard=self.gshard[a:b], sumsq=self.sumsq)
                else:
                    self.arena.reduce_scatter(self.flat.grad, s, per, adam=self._adam_args(a, b))
        else:
            view = self.flat.grad[s:e]
            dist.all_reduce(view, group=self.group)
            r = self.rank
            self.gshard[a:b].copy_(view[r * per:(r + 1) * per].float() / W)

    def _end_of_backward(self) -> None:
        self._callback_queued = False
        # units holding a parameter that received no gradient this step never reach pending == 0: reduce them here (zeros
        # in the missing slots), on every rank alike, so neither a stale gshard is re-applied nor a unit left un-updated
        self.unused = []
        for ui, n in enumerate(self._unit_nparams):
            if self._pending[ui] > 0 and n > 0:
                self.unused += [self.flat.specs[i].name for i, u in self._spec_unit.items()
                                if u == ui and self.flat.specs[i].param.grad is None]
                self._reduce_unit(ui)
        if self.fused:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        self._pending = list(self._unit_nparams)

    # ------------------------------------------------------------------ optimizer step
    @torch.no_grad()
    def step(self) -> None:
        """Finish the optimizer step (call after ``loss.backward()``), then queue the next all-gathers."""
        W = self.world_size
        native_ok = self.device.type == "cuda" and _lib.available()
        two_phase = (W == 1) or (not self.fused) or self.max_norm > 0 or self.loss_scaler is not None
        inv_scale = self.loss_scaler.inv_scale if (self.loss_scaler is not None and self.loss_scaler.enabled) else None
        found_inf = None
        if two_phase:
            if W == 1:
                g = self.flat.grad
                sumsq = None
                if self.max_norm > 0 or self.loss_scaler is not None:
                    self.sumsq.zero_()
                    if native_ok:
                        native.sumsq(g, self.sumsq)
                    else:
                        gf = g.float()
                        self.sumsq[0] = (gf * gf).sum()
                        self.sumsq[1] = (~torch.isfinite(gf)).any().float()
                    sumsq = self.sumsq
            else:
                g = self.gshard
                if not self.fused:
                    self.sumsq.zero_()
                    self.sumsq[0] = (g * g).sum()
                    self.sumsq[1] = (~torch.isfinite(g)).any().float()
                if self.max_norm > 0 or self.loss_scaler is not None:
                    if self.fused:  # scalar all-reduce on the comm stream: exact GLOBAL norm, no NCCL in the step
                        with torch.cuda.stream(self.comm_stream):
                            self.arena.allreduce_avg_(self.sumsq, 0, 4, extra_scale=1.0)
                        torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
                    else:
                        dist.all_reduce(self.sumsq, group=self.group)
                sumsq = self.sumsq
            if self.loss_scaler is not None:
                self.loss_scaler.note_overflow(self.sumsq)
                found_inf = self.loss_scaler.found_inf
            self.step_count += 1
            p16 = self.shard16 if W > 1 else self.flat.data
            master = self.master if self.master is not None else p16
            if native_ok:
                native.adamw_step(master, g, self.exp_avg, self.exp_avg_sq, p16 if self.master is not None else None,
                                  lr=self.lr, beta1=self.betas[0], beta2=self.betas[1], eps=self.eps,
                                  weight_decay=self.weight_decay, step=self.step_count, max_norm=self.max_norm,
                                  sumsq=sumsq if self.max_norm > 0 else None, inv_scale=inv_scale, found_inf=found_inf,
                                  step_t=self.step_t)
            else:
                self._adamw_torch(master, g, p16, sumsq, inv_scale, found_inf)
            self.sumsq.zero_()
        else:
            self.step_count += 1  # AdamW already applied inside the reduce-scatter kernels
        self.step_t += 1
        if self.loss_scaler is not None:
            self.loss_scaler.update()
        self.flat.zero_grad()  # also leaves param.grad undefined so the next backward's kernels write into the slots
        if W > 1:
            self._gathers_due = True   # issued at the start of the next forward (see _start_of_forward)
            if self.fused and torch.cuda.is_current_stream_capturing():
                torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)  # re-join before the capture ends
        if self.arena is not None:
            self.arena.end_step(self.comm_stream)

    def _adamw_torch(self, master, g, p16, sumsq, inv_scale, found_inf) -> None:
        if found_inf is not None and float(found_inf.reshape(-1)[0]) != 0.0:
            return
        gf = g.float()
        scale = float(inv_scale.reshape(-1)[0]) if inv_scale is not None else 1.0
        if self.max_norm > 0 and sumsq is not None:
            coef = self.max_norm / (float(sumsq[0]) ** 0.5 * scale + 1e-6)
            if coef < 1.0:
                scale *= coef
        gf = gf * scale
        b1, b2 = self.betas
        self.exp_avg.mul_(b1).add_(gf, alpha=1 - b1)
        self.exp_avg_sq.mul_(b2).addcmul_(gf, gf, value=1 - b2)
        c1, c2 = 1 - b1 ** self.step_count, 1 - b2 ** self.step_count
        denom = (self.exp_avg_sq.sqrt() / (c2 ** 0.5)).add_(self.eps)
        master.mul_(1 - self.lr * self.weight_decay).addcdiv_(self.exp_avg, denom, value=-self.lr / c1)
        if p16 is not master:
            p16.copy_(master.to(p16.dtype))

    def _queue_gathers(self, force_plain: bool = False) -> None:
        """Gather every unit's new 16-bit shard into the resident flat buffer, in forward order.  ``force_plain`` also
        gathers the GEMM-gathered units with the stand-alone kernel (``sync_params``)."""
        W = self.world_size
        order = sorted(range(len(self.units)), key=lambda ui: (self.units[ui][2] != -1, -self.units[ui][0]))
        if self.fused:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                for ui in order:
                    if ui in self._ag_handles and not force_plain:  # gathered later, inside the unit's first GEMM (K-C3)
                        self._ag_handles[ui].pending = True
                        self._gather_events[ui] = None
                        continue
                    s, e, _ = self.units[ui]
                    a, b = self._shard_range(ui)
                    self.arena.allgather(self.shard16, a, b - a, self.flat.data[s:e])
                    gev = torch.cuda.Event()
                    gev.record(self.comm_stream)
                    self._gather_events[ui] = gev
                    if ui in self._ag_handles:
                        self._ag_handles[ui].pending = False
        else:
            for ui in order:
                s, e, _ = self.units[ui]
                a, b = self._shard_range(ui)
                dist.all_gather_into_tensor(self.flat.data[s:e], self.shard16[a:b].contiguous(), group=self.group)

    def zero_grad(self, set_to_none: bool = False) -> None:
        self.flat.zero_grad()

    def sync_params(self) -> None:
        """Make the resident 16-bit parameters current NOW (the gathers after an optimizer step are otherwise issued
        lazily at the start of the next forward): call before reading ``module.parameters()`` directly."""
        if self._gathers_due:
            self._gathers_due = False
            self._queue_gathers(force_plain=True)
        elif any(h.pending for h in self._ag_handles.values()):
            self._gather_pending_ag_units()
        if self.comm_stream is not None:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)

    def _gather_pending_ag_units(self) -> None:
        """Stand-alone gather of the units that were waiting for their first GEMM (collective: every rank calls)."""
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self.comm_stream.wait_event(ev)
        with torch.cuda.stream(self.comm_stream):
            for ui, h in sorted(self._ag_handles.items()):
                if h.pending:
                    s, e, _ = self.units[ui]
                    a, b = self._shard_range(ui)
                    self.arena.allgather(self.shard16, a, b - a, self.flat.data[s:e])
                    h.pending = False

    # ------------------------------------------------------------------ checkpoints (K-C6 / K-C7)
    @torch.no_grad()
    def gather_full_state_dict(self) -> dict:
        """Collective (ALL ranks must call -- the reference calls it on rank 0 only and hangs, A-3): the fp32
        master shards are all-gathered and re-assembled into an unwrapped, unsharded state_dict."""
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        W = self.world_size
        if W == 1 or self.master is None:
            return {k: v.detach().clone() for k, v in self.module.state_dict().items()}
        gathered = torch.empty(W * self.shard_numel, device=self.device, dtype=torch.float32)
        dist.all_gather_into_tensor(gathered, self.master, group=self.group)
        full = torch.empty(self.flat.numel, device=self.device, dtype=torch.float32)
        for ui, (s, e, _) in enumerate(self.units):
            per = (e - s) // W
            a, _b = self._shard_range(ui)
            for q in range(W):
                full[s + q * per: s + (q + 1) * per] = gathered[q * self.shard_numel + a: q * self.shard_numel + a + per]
        sd = {k: v.detach().clone() for k, v in self.module.state_dict().items()}
        by_ptr = {sp.param.data_ptr(): sp for sp in self.flat.specs}
        for name, p in self.module.named_parameters():
            sp = by_ptr.get(p.data_ptr())
            if sp is not None and name in sd:
                sd[name] = full[sp.offset: sp.offset + sp.numel].view(p.shape).clone()
        for k, v in self.module.state_dict().items():  # tied parameters share storage
            sp = by_ptr.get(v.data_ptr())
            if sp is not None:
                sd[k] = full[sp.offset: sp.offset + sp.numel].view(v.shape).clone()
        return sd

    def local_shard_state(self) -> dict:
        return {"shard16": self.shard16.detach().cpu(), "master": None if self.master is None else self.master.cpu(),
                "units": self.units, "rank": self.rank, "world_size": self.world_size}

    def state_dict(self, *a, **kw):
        return self.module.state_dict(*a, **kw)

    def load_state_dict(self, sd, strict: bool = True, *a, **kw):
        """Load an unwrapped full state_dict: the resident 16-bit copy, THEN the local 16-bit shard and the fp32 master
        shard (taken from the fp32 values in ``sd`` where present, so no precision is lost through the 16-bit copy)."""
        res = self.module.load_state_dict(sd, strict=strict, *a, **kw)
        self.refresh_shards(sd)
        return res

    @torch.no_grad()
    def refresh_shards(self, sd: dict | None = None) -> None:
        """Re-derive ``shard16`` / ``master`` from the resident parameters (after any weights-only load)."""
        full32 = self.flat.data.to(torch.float32)
        if sd is not None:
            for sp in self.flat.specs:
                v = sd.get(sp.name)
                if v is not None and tuple(v.shape) == tuple(sp.param.shape):
                    from .flat import shaped_view

                    shaped_view(full32, sp).copy_(v.to(device=self.device, dtype=torch.float32))
        self._copy_local_shards(self.flat.data, self.shard16)
        if self.master is not None:
            self._copy_local_shards(full32, self.master)
        self._gathers_due = False
        for h in self._ag_handles.values():
            h.pending = False

    def optimizer_state_dict(self) -> dict:
        """This rank's shard of the optimizer state (fp32 master, exp_avg, exp_avg_sq) plus the step count."""
        # the device-side counter is the truth: a CUDA-graph replay advances it without running the Python above
        return {"step": int(self.step_t.item()) - 1, "exp_avg": self.exp_avg, "exp_avg_sq": self.exp_avg_sq, "master": self.master,
                "rank": self.rank, "world_size": self.world_size, "shard_numel": self.shard_numel}

    @torch.no_grad()
    def load_optimizer_state_dict(self, sd: dict) -> None:
        if int(sd.get("world_size", self.world_size)) != self.world_size or int(sd.get("shard_numel", self.shard_numel)) != self.shard_numel:
            raise ValueError(f"optimizer shard was saved with world_size={sd.get('world_size')} / shard_numel={sd.get('shard_numel')}; "
                             f"this engine has {self.world_size} / {self.shard_numel}")
        self.step_count = int(sd["step"])
        self.step_t.fill_(float(self.step_count + 1))
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        if self.master is not None and sd.get("master") is not None:
            self.master.copy_(sd["master"])
            self.shard16.copy_(self.master.to(self.shard16.dtype))
            if self.world_size > 1:
                self._gathers_due = True
