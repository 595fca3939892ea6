"""Peer-memory (symmetric) buffers for the communication kernels.

A :class:`SymmetricArena` owns a few buffers that every rank allocates with the same size and maps
from every peer:

* provider 1 -- ``torch.distributed._symmetric_memory`` (CUDA VMM; gives per-peer pointers and, when
  the fabric supports NVLS, a multicast pointer for ``multimem.*``);
* provider 2 -- plain CUDA IPC (``cudaIpcGetMemHandle`` / ``cudaIpcOpenMemHandle`` through
  ``libhyperion_b200.so``), used when provider 1 cannot rendezvous on a box; P2P only, no multicast.

Besides the data buffers the arena holds one small flag region (per-block barrier slots + the
``found_inf`` word the kernels raise on every rank) and a launch ``epoch`` that is advanced
identically on all ranks (every rank issues the same sequence of communication kernels).

The kernels are in ``csrc/comm.cu``; NCCL is only the control plane (rendezvous, cold-path
broadcasts, scalar all-reduces).  SURVEY.md section 5.8.
"""
from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from ..ops import _lib
from ..ops._lib import DT


def supported() -> bool:
    return torch.cuda.is_available() and _lib.available() and dist.is_available() and dist.is_initialized()


_SIGS_DECLARED = False


def _declare():
    global _SIGS_DECLARED
    if _SIGS_DECLARED:
        return
    lib = _lib.require()
    _lib.declare("hyp_comm_allreduce", "pplii" + "lli" + "i" + "p" + "pf" + "ip")
    _lib.declare("hyp_comm_reduce_scatter", "pplii" + "lli" + "i" + "p" + "f" + "pp" + "i" + "pppp" + "i" + "fffff" + "i" + "p" + "ip")
    _lib.declare("hyp_comm_allgather", "ppii" + "ll" + "p" + "i" + "p" + "p" + "ip")
    _lib.declare("hyp_gemm_ag", "ppp" + "iii" + "lll" + "ii" + "pi" + "pli" + "p" + "if" + "pp" + "iii" + "p" + "pl" + "pi" + "p" + "p")
    _lib.declare("hyp_ipc_alloc", "lpp")
    _lib.declare("hyp_ipc_open", "pp")
    _lib.declare("hyp_ipc_close", "l")
    _lib.declare("hyp_ipc_free", "l")
    for n in ("hyp_comm_flag_words", "hyp_comm_inf_word", "hyp_comm_max_blocks"):
        getattr(lib, n).restype = ctypes.c_int
    _SIGS_DECLARED = True


class SymmetricBuffer:
    """One symmetric allocation: the local tensor plus this rank's view of every peer's base address."""

    def __init__(self, tensor: torch.Tensor):
        self.tensor = tensor
        self.peer_ptrs: list[int] = []
        self.multicast_ptr: int = 0
        self._c_ptrs = None

    def c_ptrs(self):
        if self._c_ptrs is None:
            self._c_ptrs = (ctypes.c_longlong * len(self.peer_ptrs))(*self.peer_ptrs)
        return self._c_ptrs


class SymmetricArena:
    def __init__(self, device, group=None, provider: str | None = None):
        _declare()
        self.device = torch.device(device)
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world_size = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.provider = provider or os.environ.get("HYPERION_SYMM_PROVIDER", "auto")
        self.buffers: list[SymmetricBuffer] = []
        self._by_ptr: dict[int, SymmetricBuffer] = {}
        self._ready = False
        self.epoch = 0          # offset of the next kernel's epoch within the current step (host side)
        self.epoch_base = None  # device uint32: epoch of the step's first kernel; advanced by end_step()
        lib = _lib.require()
        self.flag_words = lib.hyp_comm_flag_words()
        self.inf_word = lib.hyp_comm_inf_word()
        self.max_blocks = int(os.environ.get("HYPERION_COMM_BLOCKS", "48"))
        self.use_multicast = os.environ.get("HYPERION_NVLS", "1") == "1"
        self._ipc_allocs: list[int] = []
        self._ipc_opened: list[int] = []
        self._symm_ok = None
        self.flags = self.alloc(self.flag_words, torch.int32, self.device)
        self.epoch_base = torch.ones(1, device=self.device, dtype=torch.int32)       # kernels on the comm stream
        self.epoch_base_main = torch.ones(1, device=self.device, dtype=torch.int32)  # gemm_ag kernels (main stream)
        self.epoch_main = 0

    # -- allocation ------------------------------------------------------------------------------
    def _try_symm_alloc(self, numel, dtype):
        import torch.distributed._symmetric_memory as symm_mem

        return symm_mem.empty(numel, dtype=dtype, device=self.device)

    def _ipc_alloc(self, numel, dtype):
        es = torch.empty((), dtype=dtype).element_size()
        nbytes = (numel * es + 255) // 256 * 256
        out_ptr = ctypes.c_longlong(0)
        handle = (ctypes.c_ubyte * 64)()
        _lib.call("hyp_ipc_alloc", nbytes, ctypes.addressof(out_ptr), ctypes.addressof(handle))
        self._ipc_allocs.append(out_ptr.value)
        storage = torch._C._construct_storage_from_data_pointer(out_ptr.value, self.device, nbytes)
        t = torch.empty(0, dtype=dtype, device=self.device).set_(storage, 0, (numel,), (1,))
        t._hyp_ipc_handle = bytes(handle)
        return t

    def alloc(self, numel: int, dtype: torch.dtype, device=None) -> torch.Tensor:
        """Allocate a zeroed symmetric tensor (call on every rank with the same arguments, before rendezvous)."""
        assert not self._ready or True
        t = None
        if self.provider in ("auto", "torch") and self._symm_ok is not False:
            try:
                t = self._try_symm_alloc(numel, dtype)
                self._symm_ok = True
            except Exception as e:  # pragma: no cover - depends on the box
                if self.provider == "torch":
                    raise
                self._symm_ok = False
                self._symm_err = repr(e)
        if t is None:
            t = self._ipc_alloc(numel, dtype)
        t.zero_()
        sb = SymmetricBuffer(t)
        self.buffers.append(sb)
        self._by_ptr[t.data_ptr()] = sb
        if self._ready:
            self._rendezvous_one(sb)
        return t

    # -- rendezvous ------------------------------------------------------------------------------
    def _rendezvous_one(self, sb: SymmetricBuffer) -> None:
        t = sb.tensor
        if hasattr(t, "_hyp_ipc_handle"):
            handles = [None] * self.world_size
            dist.all_gather_object(handles, t._hyp_ipc_handle, group=self.group)
            ptrs = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    ptrs.append(t.data_ptr())
                else:
                    out_ptr = ctypes.c_longlong(0)
                    hb = (ctypes.c_ubyte * 64).from_buffer_copy(h)
                    _lib.call("hyp_ipc_open", ctypes.addressof(hb), ctypes.addressof(out_ptr))
                    self._ipc_opened.append(out_ptr.value)
                    ptrs.append(out_ptr.value)
            sb.peer_ptrs, sb.multicast_ptr = ptrs, 0
        else:
            import torch.distributed._symmetric_memory as symm_mem

            try:
                symm_mem.enable_symm_mem_for_group(self.group.group_name)
            except Exception:
                pass
            hdl = symm_mem.rendezvous(t, self.group)
            sb.handle = hdl
            base = [int(p) for p in hdl.buffer_ptrs]
            off = t.data_ptr() - base[self.rank] if base[self.rank] else 0
            sb.peer_ptrs = [p + off for p in base]
            mc = int(hdl.multicast_ptr) if getattr(hdl, "multicast_ptr", 0) else 0
            sb.multicast_ptr = (mc + off) if (mc and self.use_multicast) else 0

    def rendezvous(self) -> None:
        """Collective: exchange mappings for every buffer allocated so far."""
        torch.cuda.synchronize(self.device)
        for sb in self.buffers:
            if not sb.peer_ptrs:
                self._rendezvous_one(sb)
        self._flags_sb = self._by_ptr[self.flags.data_ptr()]
        self.found_inf = self.flags.view(torch.float32)[self.inf_word:self.inf_word + 1]
        self._ready = True
        dist.barrier(group=self.group)

    def lookup(self, tensor: torch.Tensor) -> tuple[SymmetricBuffer, int]:
        """(buffer, byte offset) of a tensor that is a view into one of the arena's buffers."""
        p = tensor.data_ptr()
        for sb in self.buffers:
            b = sb.tensor.data_ptr()
            if b <= p < b + sb.tensor.numel() * sb.tensor.element_size():
                return sb, p - b
        raise ValueError("tensor is not inside a symmetric buffer of this arena")

    @property
    def has_multicast(self) -> bool:
        return any(sb.multicast_ptr for sb in self.buffers)

    def _next_epoch(self, n: int = 4) -> int:
        """Epoch offset of the next kernel.  The kernel adds the device word ``epoch_base`` to it, so a captured
        CUDA graph (which bakes the offset) stays in lock-step with the other ranks when it is replayed."""
        e = self.epoch
        self.epoch += n
        return e

    def _next_epoch_main(self, n: int = 2) -> int:
        e = self.epoch_main
        self.epoch_main += n
        return e

    def end_step(self, comm_stream=None) -> None:
        """Advance the device epoch bases past every kernel launched since the last call; call it at the same point
        of the step on every rank.  Each base is only touched on the stream whose kernels read it (``comm_stream``
        for the collectives, the current stream for the all-gather-fused GEMMs), so stream order keeps it
        consistent; during graph capture all streams are joined and the current stream is used for both."""
        if self.epoch:
            if comm_stream is not None and not torch.cuda.is_current_stream_capturing():
                with torch.cuda.stream(comm_stream):
                    self.epoch_base += self.epoch
            else:
                self.epoch_base += self.epoch
            self.epoch = 0
        if self.epoch_main:
            self.epoch_base_main += self.epoch_main
            self.epoch_main = 0

    # -- kernels ---------------------------------------------------------------------------------
    def allreduce_avg_(self, flat: torch.Tensor, start: int, end: int, inv_scale=None, found_inf=None,
                       extra_scale: float | None = None) -> None:
        """In-place average of ``flat[start:end]`` over all ranks (K-C1).  ``inv_scale`` (device fp32[1])
        folds the AMP unscale in; non-finite results raise ``self.found_inf`` on every rank."""
        sb, off = self.lookup(flat)
        es = flat.element_size()
        scale = (1.0 / self.world_size) if extra_scale is None else extra_scale
        _lib.call("hyp_comm_allreduce", ctypes.addressof(sb.c_ptrs()), ctypes.addressof(self._flags_sb.c_ptrs()),
                  sb.multicast_ptr, self.rank, self.world_size, off + start * es, end - start, DT[flat.dtype],
                  self._next_epoch(), _lib.ptr(self.epoch_base), _lib.ptr(inv_scale), float(scale), self.max_blocks,
                  _lib.stream_ptr(self.device))

    def reduce_scatter(self, grad_flat: torch.Tensor, start: int, per_elems: int, *, gshard=None, sumsq=None,
                       adam=None, scale: float | None = None) -> None:
        """K-C2: rank r reduces (averages) ``grad_flat[start + r*per : start + (r+1)*per]`` across ranks.
        ``adam`` = dict(master, m, v, p16, lr, betas, eps, weight_decay, step) applies AdamW in the same pass."""
        sb, off = self.lookup(grad_flat)
        es = grad_flat.element_size()
        scale = (1.0 / self.world_size) if scale is None else scale
        if adam is not None:
            p16 = adam.get("p16")
            args = (1, _lib.ptr(adam["master"]), _lib.ptr(adam["m"]), _lib.ptr(adam["v"]), _lib.ptr(p16),
                    DT[p16.dtype] if p16 is not None else 2, float(adam["lr"]), float(adam["betas"][0]),
                    float(adam["betas"][1]), float(adam["eps"]), float(adam["weight_decay"]), int(adam["step"]),
                    _lib.ptr(adam.get("step_t")))
        else:
            args = (0, None, None, None, None, 2, 0.0, 0.9, 0.999, 1e-8, 0.0, 1, None)
        _lib.call("hyp_comm_reduce_scatter", ctypes.addressof(sb.c_ptrs()), ctypes.addressof(self._flags_sb.c_ptrs()),
                  sb.multicast_ptr, self.rank, self.world_size, off + start * es, per_elems, DT[grad_flat.dtype],
                  self._next_epoch(), _lib.ptr(self.epoch_base), float(scale), _lib.ptr(gshard), _lib.ptr(sumsq), *args,
                  self.max_blocks,
                  _lib.stream_ptr(self.device))

    def allgather(self, shard_buf: torch.Tensor, shard_start: int, per_elems: int, dst: torch.Tensor,
                  chunk_flags: torch.Tensor | None = None) -> None:
        """K-C3 (gather part): ``dst[q*per:(q+1)*per] = rank q's shard_buf[shard_start : shard_start+per]``."""
        sb, off = self.lookup(shard_buf)
        es = shard_buf.element_size()
        assert dst.numel() >= per_elems * self.world_size and dst.dtype == shard_buf.dtype
        _lib.call("hyp_comm_allgather", ctypes.addressof(sb.c_ptrs()), ctypes.addressof(self._flags_sb.c_ptrs()),
                  self.rank, self.world_size, off + shard_start * es, per_elems * es, _lib.ptr(dst), self._next_epoch(),
                  _lib.ptr(self.epoch_base), _lib.ptr(chunk_flags), self.max_blocks, _lib.stream_ptr(self.device))

    def gemm_allgather(self, x2: torch.Tensor, w: torch.Tensor, out: torch.Tensor, *, shard_buf: torch.Tensor,
                       shard_start: int, per_elems: int, unit_dst: torch.Tensor, arrived: torch.Tensor, target: int,
                       target_base: torch.Tensor | None = None, bias=None, aux=None, act=None, preact_out=None,
                       alpha: float = 1.0) -> None:
        """K-C3 fused: ``out = epilogue(x2 @ w.T)`` where ``w`` (and ``bias``) are views into ``unit_dst`` (the local,
        full-size unit buffer) whose W slices are pulled from every rank's ``shard_buf[shard_start:+per_elems]``
        by gather warps inside the GEMM kernel."""
        from ..ops.native import ACT, _ld

        sb, off = self.lookup(shard_buf)
        es = shard_buf.element_size()
        srcs = (ctypes.c_longlong * self.world_size)(*[p + off + shard_start * es for p in sb.peer_ptrs])
        M, K = x2.shape
        N = w.shape[0]
        _lib.call("hyp_gemm_ag", _lib.ptr(x2), _lib.ptr(w), _lib.ptr(out), M, N, K, _ld(x2), _ld(w), _ld(out), DT[x2.dtype],
                  DT[out.dtype], _lib.ptr(bias), DT[bias.dtype] if bias is not None else 0, _lib.ptr(aux),
                  _ld(aux) if aux is not None else 0, DT[aux.dtype] if aux is not None else 0, _lib.ptr(preact_out), ACT[act],
                  float(alpha), ctypes.addressof(srcs), ctypes.addressof(self._flags_sb.c_ptrs()), self.rank, self.world_size,
                  self._next_epoch_main(), _lib.ptr(self.epoch_base_main), _lib.ptr(unit_dst), per_elems * es, _lib.ptr(arrived),
                  int(target), _lib.ptr(target_base), _lib.stream_ptr(self.device))

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        for p in self._ipc_opened:
            try:
                _lib.call("hyp_ipc_close", p)
            except RuntimeError:
                pass
        self._ipc_opened.clear()
