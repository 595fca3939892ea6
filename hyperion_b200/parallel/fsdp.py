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
* ``reshard_after_forward=True`` (``run_distributed.py --reshard``, ``HYPERION_FSDP_RESHARD=1``) gives the reference's
  ``ShardingStrategy.FULL_SHARD`` memory behaviour instead: a unit's gathered parameters live in one of two slot
  buffers only while the unit computes; they are gathered before its forward (the next unit is prefetched into the
  other slot), dropped, and gathered again before its backward -- twice the all-gather traffic, parameter memory =
  1/W shard + two unit-sized slots + the root unit.  Frozen parameters inside such units (the LoRA base weights) are
  sharded and gathered the same way, so "LoRA under FSDP" shards the 13.5 GB base and not only the adapters.  Units whose
  module is a leaf that the model code reads without calling it (``nn.Linear`` / ``nn.Embedding`` / norms) and the root
  parameters stay resident, like FSDP1's root unit;
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


# HYPERION_COMM_POISON=1: the race-detection build of the parameter path.  compute-sanitizer is closed on the pod this
# project is developed on, so ordering bugs between the communication stream and the compute stream are hunted this way
# instead: every gather destination is filled with NaN on the communication stream right before the gather kernel, so a
# kernel that reads parameters before its gather event has fired -- or a slot that is reused while its previous owner is
# still computing -- turns the loss into NaN, which the parity tests assert against (tests/test_gpu_multi.py).
import os as _os_mod

_POISON = _os_mod.environ.get("HYPERION_COMM_POISON", "0") == "1"


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
        import os as _os

        if reshard_after_forward is None:
            reshard_after_forward = _os.environ.get("HYPERION_FSDP_RESHARD", "0") == "1"
        self.reshard = bool(reshard_after_forward)
        self.module = module
        self.rank, self.world_size = world()
        self.group = process_group
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = lr, betas, eps, weight_decay, max_norm
        self.loss_scaler = loss_scaler
        p0 = next(p for p in module.parameters() if p.requires_grad)
        self.device = p0.device
        W = self.world_size
        self.reshard = self.reshard and W > 1
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
        import os

        # K-C3 on composite units (a transformer block, a decoder layer): the unit is gathered by warps riding inside the
        # GEMM that consumes its first weight matrix.  What the block reads BEFORE that GEMM -- its norm scale / shift
        # vectors -- would still be stale, so those few KB are carved out into the root unit (gathered up front).
        want_ag = (fuse_allgather_gemm and os.environ.get("HYPERION_FSDP_AG", "fused") == "fused" and self.fused
                   and not self.reshard and os.environ.get("HYPERION_FSDP_AG_BLOCKS", "0") == "1")
        want_ag = want_ag or os.environ.get("HYPERION_FSDP_AG_CARVE") == "1"   # layout-only switch for the CPU unit test
        owner: dict[str, int] = {}
        carved_names: set = set()
        for ui, um in enumerate(unit_modules):
            prefix = next((n for n, m in module.named_modules() if m is um), None)
            carve = set()
            if want_ag and not isinstance(um, self._LEAF_TYPES):
                for sn, sm in um.named_modules():
                    if sm is not um and "norm" in type(sm).__name__.lower():
                        carve |= {f"{sn}.{pn}" for pn, _ in sm.named_parameters(recurse=False)}
            for pn, _ in um.named_parameters():
                if pn not in carve:
                    owner[f"{prefix}.{pn}" if prefix else pn] = ui
                else:
                    carved_names.add(f"{prefix}.{pn}" if prefix else pn)
        self.unit_modules = unit_modules
        galloc = self.arena.alloc if self.fused else None
        # units are runs of consecutive parameters with the same key: the carved-out vectors go to the front of the list (=
        # the tail of the reversed flat layout, next to the embeddings) so that they do not cut a block's run in two
        named = list(module.named_parameters())
        named = [np_ for np_ in named if np_[0] in carved_names] + [np_ for np_ in named if np_[0] not in carved_names]
        self.flat = FlatParams(named, device=self.device, unit_key=lambda n: owner.get(n, -1),
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
        self._gathers_due = False
        self._rs_units: list[int] = []
        if self.reshard:
            self._setup_reshard(sync_params)
        # ---- K-C3: units whose first consumer is a GEMM on a known weight gather inside that GEMM
        self._ag_handles: dict[int, _AgHandle] = {}

        # HYPERION_FSDP_AG=plain gathers every unit with the stand-alone kernel (queued at the start of the forward, so it
        # overlaps the forward just the same) and leaves the GEMMs on the faster CTA-pair kernel
        fuse_allgather_gemm = (fuse_allgather_gemm and os.environ.get("HYPERION_FSDP_AG", "fused") == "fused"
                               and not self.reshard)
        if self.fused and fuse_allgather_gemm and self.flat.dtype in (torch.bfloat16, torch.float16):
            from ..models.transformer import MultiheadSelfAttention

            flat_ids = {id(sp.param) for sp in self.flat.specs}
            # HYPERION_FSDP_AG_BLOCKS=1: composite units (GPT-2 Block, Llama decoder layer) gather inside their first GEMM
            # too.  Off by default: with resident parameters the stand-alone gathers already hide completely under the
            # forward, and the gather GEMM runs on the one-CTA kernel -- GPT-2 medium at N=2: 28.1 ms/step fused vs 27.4
            # plain (profiles/bench/r2); leaf units (nn.Linear / attention in-projection) keep the fused path.
            blocks_ok = os.environ.get("HYPERION_FSDP_AG_BLOCKS", "0") == "1"

            def first_gemm_weight(um):
                """The weight of the first GEMM a unit's forward runs (registration order == call order in the models here)."""
                for sm in um.modules():
                    if isinstance(sm, MultiheadSelfAttention):
                        return sm.in_proj_weight
                    if isinstance(sm, nn.Linear):
                        return sm.weight
                    if sm is not um and not blocks_ok:
                        return None
                return None

            for k, um in enumerate(unit_modules):
                target = first_gemm_weight(um) if (blocks_ok or isinstance(um, (nn.Linear, MultiheadSelfAttention))) else None
                uis = self._key_to_units.get(k, [])
                if target is not None and id(target) in flat_ids and len(uis) == 1 and target.shape[1] % 8 == 0:
                    h = _AgHandle(self, uis[0])
                    target._hyp_ag = h
                    self._ag_handles[uis[0]] = h
                    if not isinstance(um, (nn.Linear, MultiheadSelfAttention)):
                        um.register_forward_hook(self._make_ag_check_hook(h, k))

    # ------------------------------------------------------------------ shard <-> flat index helpers
    def _shard_range(self, ui: int) -> tuple[int, int]:
        s, e, _ = self.units[ui]
        W = self.world_size
        return s // W, e // W

    def _copy_local_shards(self, flat_src: torch.Tensor, shard_dst: torch.Tensor) -> None:
        """``flat_src`` is a full-size flat buffer (flat offsets)."""
        W, r = self.world_size, self.rank
        for ui, (s, e, _) in enumerate(self.units):
            per = (e - s) // W
            a, b = self._shard_range(ui)
            shard_dst[a:b].copy_(flat_src[s + r * per: s + (r + 1) * per].to(shard_dst.dtype))

    # ------------------------------------------------------------------ FULL_SHARD mode (reshard_after_forward)
    _LEAF_TYPES = (nn.Linear, nn.Embedding, nn.LayerNorm, nn.Conv2d, nn.modules.batchnorm._BatchNorm)

    @staticmethod
    def _view_at(buf: torch.Tensor, off: int, numel: int, shape, channels_last: bool = False) -> torch.Tensor:
        flat = buf[off:off + numel]
        if channels_last:
            K, C, R, S = shape
            return flat.view(K, R, S, C).permute(0, 3, 1, 2)
        return flat.view(shape)

    def _dst(self, ui: int) -> torch.Tensor:
        """Where unit ``ui``'s full 16-bit parameters live when gathered."""
        s, e, _ = self.units[ui]
        if not self.reshard:
            return self.flat.data[s:e]
        buf, off = self._unit_base[ui]
        return buf[off:off + (e - s)]

    @torch.no_grad()
    def _setup_reshard(self, sync_params: bool) -> None:
        W, r, dev, dt = self.world_size, self.rank, self.device, self.flat.dtype
        K = 2
        # forward order of the units: root first, then decreasing flat offset (the flat buffer is in reverse registration order)
        order = sorted(range(len(self.units)), key=lambda ui: (self.units[ui][2] != -1, -self.units[ui][0]))
        self._rs_units = [ui for ui in order if self.units[ui][2] != -1
                          and not isinstance(self.unit_modules[self.units[ui][2]], self._LEAF_TYPES)]
        rs = set(self._rs_units)
        self._rs_pos = {ui: i for i, ui in enumerate(self._rs_units)}
        self._slot_of = {ui: i % K for i, ui in enumerate(self._rs_units)}
        self._slot_owner: list = [None] * K
        max_unit = max((self.units[ui][1] - self.units[ui][0] for ui in self._rs_units), default=0)
        self._slots = [torch.empty(max_unit, dtype=dt, device=dev) for _ in range(K)] if max_unit else []
        # resident buffer: root + leaf units, compacted
        self._unit_base: dict[int, tuple[torch.Tensor, int]] = {}
        roff = 0
        for ui in range(len(self.units)):
            if ui not in rs:
                self._unit_base[ui] = (None, roff)
                roff += self.units[ui][1] - self.units[ui][0]
        self._resident = torch.empty(max(roff, 1), dtype=dt, device=dev)
        for ui in range(len(self.units)):
            s, e, _ = self.units[ui]
            if ui in rs:
                self._unit_base[ui] = (self._slots[self._slot_of[ui]], 0)
            else:
                off = self._unit_base[ui][1]
                self._unit_base[ui] = (self._resident, off)
                self._resident[off:off + (e - s)].copy_(self.flat.data[s:e])
        for i, sp in enumerate(self.flat.specs):
            ui = self._spec_unit[i]
            buf, off = self._unit_base[ui]
            sp.param.data = self._view_at(buf, off + sp.offset - self.units[ui][0], sp.numel, sp.param.shape, sp.channels_last)
        self.flat.data = torch.empty(0, dtype=dt, device=dev)   # the full-size 16-bit copy is gone
        # ---- frozen parameters inside reshardable units (LoRA base weights): sharded too
        self._frozen: dict[int, list] = {}       # ui -> [(name, param, offset in unit, numel)]
        self._frozen_range: dict[int, tuple[int, int, int]] = {}  # ui -> (shard start, per-rank elems, unit elems)
        names = {id(p_): n for n, p_ in self.module.named_parameters()}
        seen: set = set()
        foff = 0
        max_f = 0
        for ui in self._rs_units:
            um = self.unit_modules[self.units[ui][2]]
            lst, off = [], 0
            for p_ in um.parameters():
                if p_.requires_grad or id(p_) in seen or p_.dtype != dt or p_.device != dev:
                    continue
                seen.add(id(p_))
                lst.append((names.get(id(p_), ""), p_, off, p_.numel()))
                off += (p_.numel() + 127) // 128 * 128
            if not lst:
                continue
            size = (off + W * 128 - 1) // (W * 128) * (W * 128)
            self._frozen[ui] = lst
            self._frozen_range[ui] = (foff, size // W, size)
            foff += size // W
            max_f = max(max_f, size)
        self.fshard16 = None
        self._fslots = []
        if self._frozen:
            alloc16 = self.arena.alloc if self.fused else (lambda n, d, dv: torch.zeros(n, dtype=d, device=dv))
            self.fshard16 = alloc16(foff, dt, dev)
            tmp = torch.empty(max_f, dtype=dt, device=dev)
            for ui, lst in self._frozen.items():
                fa, per, size = self._frozen_range[ui]
                tmp.zero_()
                for _n, p_, off, n in lst:
                    tmp[off:off + n].copy_(p_.data.reshape(-1))
                if sync_params:
                    dist.broadcast(tmp[:size], src=0, group=self.group)
                self.fshard16[fa:fa + per].copy_(tmp[r * per:(r + 1) * per])
            del tmp
            self._fslots = [torch.empty(max_f, dtype=dt, device=dev) for _ in range(K)]
            for ui, lst in self._frozen.items():
                slot = self._fslots[self._slot_of[ui]]
                for _n, p_, off, n in lst:
                    p_.data = slot[off:off + n].view(p_.shape)
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)
            torch.cuda.empty_cache()
        for ui in self._rs_units:
            um = self.unit_modules[self.units[ui][2]]
            um.register_forward_pre_hook(self._make_rs_hook(ui))
            um.register_full_backward_pre_hook(self._make_rs_bwd_hook(ui))

    def _gather_unit(self, ui: int) -> None:
        """All-gather unit ``ui`` (trainable and frozen parts) into its slot; ordered after everything already enqueued on
        the compute stream, which includes the last use of the slot's previous owner."""
        a, b = self._shard_range(ui)
        dst = self._dst(ui)
        fr = self._frozen_range.get(ui)
        if self.fused:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            self.comm_stream.wait_event(ev)
            with torch.cuda.stream(self.comm_stream):
                if _POISON:
                    dst.fill_(float("nan"))   # debug build: a consumer that runs ahead of the gather event computes NaNs
                self.arena.allgather(self.shard16, a, b - a, dst)
                if fr is not None:
                    self.arena.allgather(self.fshard16, fr[0], fr[1], self._fslots[self._slot_of[ui]][:fr[2]])
                gev = torch.cuda.Event()
                gev.record(self.comm_stream)
            self._gather_events[ui] = gev
        else:
            dist.all_gather_into_tensor(dst, self.shard16[a:b].contiguous(), group=self.group)
            if fr is not None:
                dist.all_gather_into_tensor(self._fslots[self._slot_of[ui]][:fr[2]], self.fshard16[fr[0]:fr[0] + fr[1]].contiguous(),
                                            group=self.group)
        self.gathers_issued = getattr(self, "gathers_issued", 0) + 1

    def _ensure_unit(self, ui: int, wait: bool) -> None:
        slot = self._slot_of[ui]
        if self._slot_owner[slot] != ui:
            self._slot_owner[slot] = ui
            self._gather_unit(ui)
        if wait and self.fused:
            ev = self._gather_events[ui]
            if ev is not None:
                torch.cuda.current_stream(self.device).wait_event(ev)

    def _rs_touch(self, ui: int, backward: bool) -> None:
        self._ensure_unit(ui, wait=True)
        nxt = self._rs_pos[ui] + (-1 if backward else 1)
        if 0 <= nxt < len(self._rs_units):
            self._ensure_unit(self._rs_units[nxt], wait=False)   # prefetch into the other slot

    def _make_rs_hook(self, ui: int):
        def hook(_m, _args):
            # a forward that runs inside backward is an activation-checkpoint recompute: prefetch in backward direction
            self._rs_touch(ui, backward=torch._C._current_graph_task_id() != -1)
        return hook

    def _make_rs_bwd_hook(self, ui: int):
        def hook(_m, _grad_out):
            self._rs_touch(ui, backward=True)
        return hook

    def _make_ag_check_hook(self, handle: "_AgHandle", key: int):
        def hook(_m, _args, _out):
            # the unit's first GEMM must have run (and gathered the unit) inside this forward; anything else would have
            # computed with last step's weights
            if handle.pending:
                raise RuntimeError(f"FSDP unit {key}: its forward finished without running the GEMM that gathers it "
                                   "(HYPERION_FSDP_AG_BLOCKS=0 or HYPERION_FSDP_AG=plain selects stand-alone gathers)")
        return hook

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
            if self._ag_handles and self.comm_stream is not None:
                # Co-residency invariant: a GEMM with gather warps (one CTA per SM, spinning on its peers' flags) must never
                # share the GPU with a communication kernel of this engine (512 threads x 128 registers = a whole SM each, also
                # spinning on peers): neither fits beside the other, and across two ranks the waits could form a cycle.  So
                # every stand-alone gather of this step is complete before the first forward kernel is issued; during the
                # forward the comm stream stays idle (all remaining units gather inside GEMMs), and in backward the
                # reduce-scatters only ever run beside ordinary kernels, which never wait for a peer.
                torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)

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
        from ..utils import nvtx

        nvtx.mark(f"hyperion/comm/reduce_scatter unit {ui}")
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
        if self.reshard:
            self._slot_owner = [None] * len(self._slot_owner)   # the slots hold pre-update parameters now
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
                    if ui in self._slot_of_or_empty():   # FULL_SHARD unit: gathered on demand into its slot
                        continue
                    if ui in self._ag_handles and not force_plain:  # gathered later, inside the unit's first GEMM (K-C3)
                        self._ag_handles[ui].pending = True
                        self._gather_events[ui] = None
                        continue
                    s, e, _ = self.units[ui]
                    a, b = self._shard_range(ui)
                    if _POISON:
                        self._dst(ui).fill_(float("nan"))
                    self.arena.allgather(self.shard16, a, b - a, self._dst(ui))
                    gev = torch.cuda.Event()
                    gev.record(self.comm_stream)
                    self._gather_events[ui] = gev
                    if ui in self._ag_handles:
                        self._ag_handles[ui].pending = False
        else:
            for ui in order:
                if ui in self._slot_of_or_empty():
                    continue
                a, b = self._shard_range(ui)
                dist.all_gather_into_tensor(self._dst(ui), self.shard16[a:b].contiguous(), group=self.group)

    def _slot_of_or_empty(self) -> dict:
        return self._slot_of if self.reshard else {}

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
                    self.arena.allgather(self.shard16, a, b - a, self._dst(ui))
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
        by_id = {id(sp.param): sp for sp in self.flat.specs}
        for name, p in self.module.named_parameters(remove_duplicate=False):   # tied parameters appear under every name
            sp = by_id.get(id(p))
            if sp is not None and name in sd:
                from .flat import shaped_view

                sd[name] = shaped_view(full, sp).clone()
        if self.reshard and self._frozen:     # sharded frozen parameters (their slots hold whichever unit ran last)
            F = self.fshard16.numel()
            gf = torch.empty(W * F, device=self.device, dtype=self.fshard16.dtype)
            dist.all_gather_into_tensor(gf, self.fshard16.contiguous(), group=self.group)
            for ui, lst in self._frozen.items():
                fa, per, size = self._frozen_range[ui]
                unit = torch.cat([gf[q * F + fa: q * F + fa + per] for q in range(W)])
                for name, p_, off, n in lst:
                    if name in sd:
                        sd[name] = unit[off:off + n].view(p_.shape).clone()
        return sd

    def local_shard_state(self) -> dict:
        return {"shard16": self.shard16.detach().cpu(), "master": None if self.master is None else self.master.cpu(),
                "units": self.units, "rank": self.rank, "world_size": self.world_size}

    def state_dict(self, *a, **kw):
        return self.module.state_dict(*a, **kw)

    def load_state_dict(self, sd, strict: bool = True, *a, **kw):
        """Load an unwrapped full state_dict: the resident 16-bit copy, THEN the local 16-bit shard and the fp32 master
        shard (taken from the fp32 values in ``sd`` where present, so no precision is lost through the 16-bit copy)."""
        if self.reshard:
            return self._load_state_dict_resharded(sd, strict)
        res = self.module.load_state_dict(sd, strict=strict, *a, **kw)
        self.refresh_shards(sd)
        return res

    @torch.no_grad()
    def _load_state_dict_resharded(self, sd, strict: bool):
        from .flat import shaped_view

        W, r = self.world_size, self.rank
        full32 = torch.zeros(self.flat.numel, device=self.device, dtype=torch.float32)
        handled, missing = set(), []
        for sp in self.flat.specs:
            v = sd.get(sp.name)
            if v is None:
                missing.append(sp.name)
                continue
            shaped_view(full32, sp).copy_(v.to(device=self.device, dtype=torch.float32))
            handled.add(sp.name)
        self._copy_local_shards(full32, self.shard16)
        if self.master is not None:
            self._copy_local_shards(full32, self.master)
        for ui in range(len(self.units)):
            if ui not in self._slot_of:
                s, e, _ = self.units[ui]
                self._dst(ui).copy_(full32[s:e].to(self.flat.dtype))
        del full32
        for ui, lst in self._frozen.items():
            fa, per, size = self._frozen_range[ui]
            tmp = torch.zeros(size, device=self.device, dtype=self.fshard16.dtype)
            for name, p_, off, n in lst:
                v = sd.get(name)
                if v is None:
                    missing.append(name)
                    continue
                tmp[off:off + n].copy_(v.reshape(-1).to(device=self.device, dtype=tmp.dtype))
                handled.add(name)
            self.fshard16[fa:fa + per].copy_(tmp[r * per:(r + 1) * per])
        if strict and missing:
            raise KeyError(f"state_dict is missing {missing[:5]}{'...' if len(missing) > 5 else ''}")
        tied = {n for n, p_ in self.module.named_parameters(remove_duplicate=False)
                if id(p_) in {id(sp.param) for sp in self.flat.specs}}
        rest = {k: v for k, v in sd.items() if k not in handled and k not in tied}
        res = self.module.load_state_dict(rest, strict=False)
        self._slot_owner = [None] * len(self._slot_owner)
        self._gathers_due = False
        return res

    @torch.no_grad()
    def refresh_shards(self, sd: dict | None = None) -> None:
        """Re-derive ``shard16`` / ``master`` from the resident parameters (after any weights-only load)."""
        if self.reshard:
            if sd is None:
                raise RuntimeError("FULL_SHARD mode keeps no resident copy to refresh from: pass the state_dict")
            self._load_state_dict_resharded(sd, strict=False)
            return
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
