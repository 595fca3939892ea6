"""Flat parameter storage.

Trainable parameters of a module are re-pointed at views of ONE flat buffer (``data``), and
their ``.grad`` at views of a second one (``grad``), so that

* AdamW is a single streaming kernel over the flat buffers (``optim.FusedAdamW``),
* a DDP bucket or an FSDP shard is a contiguous byte range that a communication kernel can
  address on every peer with the same offset (the buffers can live in symmetric memory).

Order: parameters are laid out in REVERSE registration order, which is (approximately) the order
their gradients become ready in backward -- so bucket k of the flat gradient is complete before
bucket k+1 and can be reduced while backward is still running.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.nn as nn

from ..ops.gradslot import GradSlot

ALIGN_BYTES = 256


@dataclass
class ParamSpec:
    name: str
    param: nn.Parameter
    offset: int   # element offset into the flat buffers
    numel: int
    channels_last: bool = False   # 4-D conv weight kept in [K, R, S, C] memory order (what the NHWC conv kernels read)


def shaped_view(buf: torch.Tensor, s: "ParamSpec") -> torch.Tensor:
    """The parameter-shaped view of ``s``'s range of a flat buffer, in the parameter's memory format."""
    flat = buf[s.offset:s.offset + s.numel]
    if s.channels_last:
        K, C, R, S = s.param.shape
        return flat.view(K, R, S, C).permute(0, 3, 1, 2)
    return flat.view(s.param.shape)


class FlatParams:
    def __init__(self, named_params, dtype: torch.dtype | None = None, device=None, reverse: bool = True,
                 pad_to: int = 1, allocator=None, unit_key=None, unit_align: int = 1, grad_allocator=None):
        """``named_params``: iterable of (name, Parameter); only ``requires_grad`` ones are taken.
        ``allocator(numel, dtype, device)`` lets callers place the buffers in symmetric memory."""
        items = [(n, p) for n, p in named_params if p.requires_grad]
        seen, uniq = set(), []
        for n, p in items:  # tied parameters appear once
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append((n, p))
        if reverse:
            uniq = uniq[::-1]
        if not uniq:
            raise ValueError("no trainable parameters")
        self.dtype = dtype or uniq[0][1].dtype
        self.device = device or uniq[0][1].device
        es = torch.empty((), dtype=self.dtype).element_size()
        align = ALIGN_BYTES // es
        self.specs: list[ParamSpec] = []
        # ``unit_key(name)`` groups consecutive parameters into units (FSDP); a unit starts at a multiple of
        # ``unit_align`` elements so it splits evenly over the ranks in 16-byte multiples.
        self.units: list[list] = []  # [start, end, key]
        off = 0
        prev_key = object()
        for n, p in uniq:
            key = unit_key(n) if unit_key is not None else None
            if unit_key is not None and key != prev_key:
                off = (off + unit_align - 1) // unit_align * unit_align
                if self.units:
                    self.units[-1][1] = off
                self.units.append([off, off, key])
                prev_key = key
            cl = p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last) and not p.is_contiguous()
            self.specs.append(ParamSpec(n, p, off, p.numel(), cl))
            off += (p.numel() + align - 1) // align * align
        pad_to = max(pad_to, unit_align)
        self.numel = (off + pad_to - 1) // pad_to * pad_to
        if self.units:
            self.units[-1][1] = self.numel
        alloc = allocator or (lambda n, dt, dev: torch.zeros(n, dtype=dt, device=dev))
        galloc = grad_allocator or alloc
        self.data = alloc(self.numel, self.dtype, self.device)
        self.grad = galloc(self.numel, self.dtype, self.device)
        self.data.zero_()
        self.grad.zero_()
        self.stray_copies = 0  # gradients that arrived outside their slot and had to be copied in (see ops/gradslot.py)
        self.slots: list[GradSlot] = []
        with torch.no_grad():
            for s in self.specs:
                view = shaped_view(self.data, s)
                view.copy_(s.param.data.to(self.dtype))
                s.param.data = view
                s.param.grad = None
                slot = GradSlot(self, s)
                s.param._hyp_slot = slot
                self.slots.append(slot)
                # registered before any DDP / FSDP hook, so those see the gradient in the flat buffer
                s.param.register_post_accumulate_grad_hook(self._make_settle_hook(slot))

    def _make_settle_hook(self, slot: GradSlot):
        es = torch.empty((), dtype=self.dtype).element_size()

        def settle(param):
            g = param.grad
            if g is None:
                return
            if g.data_ptr() != self.grad.data_ptr() + slot.spec.offset * es or g.dtype != self.grad.dtype:
                view = slot.view()
                with torch.no_grad():
                    view.copy_(g)
                param.grad = view
                self.stray_copies += 1

        return settle

    def grad_view(self, i: int) -> torch.Tensor:
        """View of parameter ``i``'s range of the flat gradient (what ``param.grad`` is after a backward)."""
        return self.slots[i].view()

    def zero_grad(self) -> None:
        """Zero the flat gradient and leave every ``param.grad`` undefined, so the next backward's kernels can store
        straight into the slots (ops/gradslot.py); after a backward ``param.grad`` is the slot view again."""
        self.grad.zero_()
        for slot in self.slots:
            slot.claimed = False
            slot.spec.param.grad = None

    def rebind(self, data: torch.Tensor | None = None, grad: torch.Tensor | None = None) -> None:
        """Point parameters / grads at new flat buffers of the same size (contents are copied)."""
        with torch.no_grad():
            if data is not None:
                data.copy_(self.data)
                self.data = data
                for s in self.specs:
                    s.param.data = shaped_view(data, s)
            if grad is not None:
                grad.copy_(self.grad)
                self.grad = grad
                for s in self.specs:
                    if s.param.grad is not None:
                        s.param.grad = shaped_view(grad, s)

    def buckets(self, cap_bytes: int, first_cap_bytes: int | None = None) -> list[tuple[int, int, list[int]]]:
        """Contiguous (start, end, [spec indices]) element ranges of at most ``cap_bytes`` (whole parameters)."""
        es = self.data.element_size()
        out, cur_start, cur_idx = [], 0, []
        cap = first_cap_bytes or cap_bytes
        for i, s in enumerate(self.specs):
            end = self.specs[i + 1].offset if i + 1 < len(self.specs) else self.numel
            if cur_idx and (end - cur_start) * es > cap:
                out.append((cur_start, s.offset, cur_idx))
                cur_start, cur_idx, cap = s.offset, [], cap_bytes
            cur_idx.append(i)
        out.append((cur_start, self.numel, cur_idx))
        return out
