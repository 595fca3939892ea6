"""Thin functional wrappers (no autograd) over the C ABI: tensors in, tensors out.

Every function here launches sm_100a kernels from ``libhyperion_b200.so`` on the
current CUDA stream.  The autograd-aware ops live in the sibling modules and
are built from these.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import DT, call, ptr, stream_ptr

ACT = {None: 0, "none": 0, "relu": 1, "gelu": 2, "gelu_erf": 2, "gelu_tanh": 3, "silu": 4}


def _ld(t: torch.Tensor) -> int:
    assert t.dim() == 2 and t.stride(1) == 1, "expected a row-major 2-D tensor (unit inner stride)"
    return t.stride(0) if t.size(0) > 1 else max(t.stride(0), t.size(1))


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn_major: bool = False, b_mn_major: bool = False,
         out: torch.Tensor | None = None, out_dtype: torch.dtype | None = None, bias: torch.Tensor | None = None,
         aux: torch.Tensor | None = None, act: str | None = None, mode: int = 0, alpha: float = 1.0,
         beta: float = 0.0, splits: int = 1, preact_out: torch.Tensor | None = None,
         colsum: torch.Tensor | None = None) -> torch.Tensor:
    """D = epilogue(alpha * A @ B^T) on the tcgen05 GEMM.

    ``colsum`` (``mode=1`` only): fp32 ``[N]`` accumulator; the epilogue adds the column sums of the rounded output to it
    (bias gradient of the layer whose pre-activation gradient this GEMM produces).

    ``a`` is [M, K] (or [K, M] when ``a_mn_major``); ``b`` is [N, K] -- the
    ``nn.Linear.weight`` layout -- (or [K, N] when ``b_mn_major``).
    """
    assert a.is_cuda and b.is_cuda and a.dtype == b.dtype
    if a.dtype == torch.float32:  # tf32 path is K-major only
        if a_mn_major:
            a, a_mn_major = a.t().contiguous(), False
        if b_mn_major:
            b, b_mn_major = b.t().contiguous(), False
    if a_mn_major:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn_major:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, f"reduction dims differ: {K} vs {Kb}"
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=out_dtype or a.dtype)
        if splits > 1:
            out.zero_()
    assert out.shape == (M, N)
    if preact_out is not None:
        assert preact_out.shape == out.shape and preact_out.dtype == out.dtype and _ld(preact_out) == _ld(out)
    if colsum is not None:
        assert mode == 1 and preact_out is None and colsum.dtype == torch.float32 and colsum.numel() == N and colsum.is_contiguous()
        preact_out = colsum   # the C ABI reuses the pre-activation argument for it in mode 1 (include/hyperion_b200.h)
    call("hyp_gemm", ptr(a), ptr(b), ptr(out), M, N, K, _ld(a), _ld(b), _ld(out), DT[a.dtype], DT[out.dtype],
         int(a_mn_major), int(b_mn_major), ptr(bias), DT[bias.dtype] if bias is not None else 0, ptr(aux),
         _ld(aux) if aux is not None else 0, DT[aux.dtype] if aux is not None else 0, ptr(preact_out), ACT[act], mode,
         float(alpha),
         float(beta), int(splits), stream_ptr())
    return out


def gemm_dualk(a: torch.Tensor, b: torch.Tensor, a2: torch.Tensor, b2: torch.Tensor, *, b_mn_major: bool = False,
               out: torch.Tensor | None = None, bias: torch.Tensor | None = None, aux: torch.Tensor | None = None,
               alpha: float = 1.0) -> torch.Tensor:
    """``D = alpha * (a @ b^T + a2 @ b2^T) + bias + aux`` in one launch (second pair K-major, K2 % 64 == 0)."""
    M, K = a.shape
    N = b.shape[1] if b_mn_major else b.shape[0]
    K2 = a2.shape[1]
    assert a2.shape == (M, K2) and b2.shape == (N, K2) and K2 % 64 == 0 and a2.stride(1) == 1 and b2.stride(1) == 1
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=a.dtype)
    call("hyp_gemm_dualk", ptr(a), ptr(b), ptr(out), M, N, K, _ld(a), _ld(b), _ld(out), DT[a.dtype], DT[out.dtype], 0,
         int(b_mn_major), ptr(a2), ptr(b2), K2, _ld(a2), _ld(b2), ptr(bias), DT[bias.dtype] if bias is not None else 0,
         ptr(aux), _ld(aux) if aux is not None else 0, DT[aux.dtype] if aux is not None else 0, float(alpha), stream_ptr())
    return out


def stream_add(x: torch.Tensor, y: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(x)
    call("hyp_stream_add", ptr(x), ptr(y), ptr(out), x.numel(), DT[x.dtype], stream_ptr())
    return out


def stream_copy(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty_like(x)
    call("hyp_stream_copy", ptr(x), ptr(out), x.numel() * x.element_size(), stream_ptr())
    return out


def colsum(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """fp32 column sums of a 2-D tensor (accumulates into ``out`` when given)."""
    M, N = x.shape
    if out is None:
        out = torch.zeros(N, device=x.device, dtype=torch.float32)
    call("hyp_colsum", ptr(x), ptr(out), M, N, _ld(x), DT[x.dtype], stream_ptr())
    return out


_colsum_ws: dict = {}


def colsum_cast(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """Column sums of a 16-bit 2-D tensor, accumulated in fp32 and rounded to x's dtype, in one launch (bias
    gradients).  The fp32 accumulator is a per-device workspace the kernel leaves zeroed."""
    M, N = x.shape
    ws = _colsum_ws.get(x.device)
    if ws is None or ws[0].numel() < N:
        ws = (torch.zeros(max(N, 65536), device=x.device, dtype=torch.float32),
              torch.zeros(1, device=x.device, dtype=torch.int32))
        _colsum_ws[x.device] = ws
    if out is None:
        out = torch.empty(N, device=x.device, dtype=x.dtype)
    assert out.dtype == x.dtype and out.numel() == N and out.is_contiguous()
    call("hyp_colsum_cast", ptr(x), ptr(out), M, N, _ld(x), DT[x.dtype], ptr(ws[0]), ptr(ws[1]), stream_ptr())
    return out


def swiglu_fwd(gate: torch.Tensor, up: torch.Tensor) -> torch.Tensor:
    out = torch.empty_like(gate)
    call("hyp_swiglu_fwd", ptr(gate), ptr(up), ptr(out), gate.numel(), DT[gate.dtype], stream_ptr())
    return out


def swiglu_bwd(gate, up, dout):
    dgate, dup = torch.empty_like(gate), torch.empty_like(up)
    call("hyp_swiglu_bwd", ptr(gate), ptr(up), ptr(dout), ptr(dgate), ptr(dup), gate.numel(), DT[gate.dtype],
         stream_ptr())
    return dgate, dup


def cast(x: torch.Tensor, dtype: torch.dtype, scale: float = 1.0, out: torch.Tensor | None = None,
         scale_t: torch.Tensor | None = None) -> torch.Tensor:
    """``out = (x * scale * scale_t).to(dtype)``; ``scale_t`` is an fp32 device scalar read by the kernel."""
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=dtype)
    assert out.numel() == x.numel() and out.dtype == dtype and out.is_contiguous() and x.is_contiguous()
    if scale_t is not None:
        assert scale_t.dtype == torch.float32 and scale_t.numel() == 1
        call("hyp_cast_dscale", ptr(x), ptr(out), x.numel(), DT[x.dtype], DT[dtype], float(scale), ptr(scale_t), stream_ptr())
    else:
        call("hyp_cast", ptr(x), ptr(out), x.numel(), DT[x.dtype], DT[dtype], float(scale), stream_ptr())
    return out


def norm_fwd(x, gamma, beta=None, residual=None, eps: float = 1e-5, rms: bool = False, want_sum: bool = False,
             drop=None):
    """``drop = (rng_state, salt, p)``: ``z = dropout(x) + residual`` inside the same pass (hyp_norm_fwd_drop)."""
    d = x.shape[-1]
    rows = x.numel() // d
    y = torch.empty_like(x)
    z = torch.empty_like(x) if (want_sum or residual is not None or drop is not None) else None
    mean = torch.empty(rows, device=x.device, dtype=torch.float32)
    rstd = torch.empty(rows, device=x.device, dtype=torch.float32)
    if drop is not None:
        st, salt, p = drop
        call("hyp_norm_fwd_drop", ptr(x), ptr(residual), ptr(gamma), ptr(beta), ptr(y), ptr(z), ptr(mean), ptr(rstd), rows, d,
             float(eps), int(rms), DT[x.dtype], ptr(st), int(salt), float(p), stream_ptr())
    else:
        call("hyp_norm_fwd", ptr(x), ptr(residual), ptr(gamma), ptr(beta), ptr(y), ptr(z), ptr(mean), ptr(rstd), rows, d,
             float(eps), int(rms), DT[x.dtype], stream_ptr())
    return y, (z if z is not None else x), mean, rstd


def dropout(x: torch.Tensor, rng_state: torch.Tensor, salt: int, p: float, out: torch.Tensor | None = None) -> torch.Tensor:
    """``x * keep * 1/(1-p_eff)`` with the (row, col) Philox mask of common/philox.cuh over ``x`` viewed as
    ``[-1, last dim]`` (contiguous); the same call on a gradient is the backward."""
    assert x.is_contiguous()
    cols = x.shape[-1]
    out = torch.empty_like(x) if out is None else out
    call("hyp_dropout", ptr(x), ptr(out), x.numel() // cols, cols, DT[x.dtype], ptr(rng_state), int(salt), float(p), stream_ptr())
    return out


_scratch: dict = {}


def _get_scratch(device, nbytes: int) -> torch.Tensor:
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 1 << 20), device=device, dtype=torch.uint8)
        _scratch[key] = buf
    return buf


_norm_ws: dict = {}


def norm_bwd(dy, z, gamma, mean, rstd, rms: bool = False, need_beta: bool = True, dres=None, dgamma_out=None,
             dbeta_out=None, dz_colsum_out=None):
    """dz (+ ``dres``: the skip-connection gradient, added in the same pass), dgamma, dbeta -- one kernel.
    ``dgamma_out`` / ``dbeta_out``: contiguous tensors of gamma's dtype (e.g. flat-gradient slots) that receive the
    parameter gradients directly; otherwise they come back as fresh fp32 tensors.  ``dz_colsum_out`` ([d], same dtype
    as the parameter gradients, d <= 2048) additionally receives the column sums of dz."""
    d = z.shape[-1]
    rows = z.numel() // d
    dz = torch.empty_like(z)
    want_beta = need_beta and not rms
    # parameter gradients are written in ONE dtype -- the slots' when given, else fp32
    pdt = dgamma_out.dtype if dgamma_out is not None else (dbeta_out.dtype if dbeta_out is not None else torch.float32)
    dgamma = dgamma_out if dgamma_out is not None else torch.empty(d, device=z.device, dtype=pdt)
    dbeta = (dbeta_out if dbeta_out is not None else torch.empty(d, device=z.device, dtype=pdt)) if want_beta else None
    assert dgamma.dtype == pdt and (dbeta is None or dbeta.dtype == pdt)
    if dres is not None:
        assert dres.shape == z.shape and dres.dtype == z.dtype and dres.is_contiguous()
    ws = _norm_ws.get(z.device)   # persistent accumulator the kernel leaves zeroed
    if ws is None:
        ws = torch.zeros(3 * 8192 + 8, device=z.device, dtype=torch.float32)
        _norm_ws[z.device] = ws
    if dz_colsum_out is not None:
        assert dz_colsum_out.dtype == pdt and dz_colsum_out.numel() == d and d <= 2048
    call("hyp_norm_bwd_fused", ptr(dy), ptr(z), ptr(gamma), ptr(mean), ptr(rstd), ptr(dres), ptr(dz), ptr(dgamma),
         ptr(dbeta), ptr(dz_colsum_out), DT[pdt], ptr(ws), ws.numel() * 4, rows, d, int(rms), DT[z.dtype], stream_ptr())
    return dz, dgamma, dbeta


def norm_bwd_two_pass(dy, z, gamma, mean, rstd, rms: bool = False, need_beta: bool = True):
    """The deterministic two-kernel backward (``hyp_norm_bwd``: per-block partial sums in caller scratch, then a
    column reduce); fp32 dgamma / dbeta."""
    d = z.shape[-1]
    rows = z.numel() // d
    dz = torch.empty_like(z)
    dgamma = torch.empty(d, device=z.device, dtype=torch.float32)
    dbeta = torch.empty(d, device=z.device, dtype=torch.float32) if (need_beta and not rms) else None
    nbytes = _lib.require().hyp_norm_bwd_scratch_bytes(rows, d)
    scratch = _get_scratch(z.device, nbytes)
    call("hyp_norm_bwd", ptr(dy), ptr(z), ptr(gamma), ptr(mean), ptr(rstd), ptr(dz), ptr(dgamma), ptr(dbeta),
         ptr(scratch), scratch.numel(), rows, d, int(rms), DT[z.dtype], 0, stream_ptr())
    return dz, dgamma, dbeta


def adamw_step(p, g, m, v, p_copy, *, lr, beta1, beta2, eps, weight_decay, step, max_norm=0.0, sumsq=None,
               inv_scale=None, found_inf=None, step_t=None) -> None:
    call("hyp_adamw", ptr(p), ptr(g), ptr(m), ptr(v), ptr(p_copy), p.numel(), DT[g.dtype],
         DT[p_copy.dtype] if p_copy is not None else 2, float(lr), float(beta1), float(beta2), float(eps),
         float(weight_decay), int(step), float(max_norm), ptr(sumsq), ptr(inv_scale), ptr(found_inf), ptr(step_t),
         stream_ptr())


def sumsq(x: torch.Tensor, out2: torch.Tensor | None = None) -> torch.Tensor:
    if out2 is None:
        out2 = torch.zeros(2, device=x.device, dtype=torch.float32)
    call("hyp_sumsq", ptr(x), x.numel(), DT[x.dtype], ptr(out2), stream_ptr())
    return out2


def ce_fwd_bwd(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int, sum_count: torch.Tensor,
               gscale_ptr: torch.Tensor | None = None, gscale: float = 1.0, write_grad: bool = True,
               n_classes: int | None = None, row_lse: torch.Tensor | None = None) -> None:
    rows = logits.shape[0]
    V = n_classes if n_classes is not None else logits.shape[1]
    call("hyp_ce_fwd_bwd", ptr(logits), _ld(logits), ptr(targets), rows, V, int(ignore_index), ptr(sum_count),
         ptr(row_lse), ptr(gscale_ptr), float(gscale), int(write_grad), DT[logits.dtype], stream_ptr())


def ce_finalize(sum_count, loss_out=None, inv_count_out=None) -> None:
    call("hyp_ce_finalize", ptr(sum_count), ptr(loss_out), ptr(inv_count_out), stream_ptr())


def embedding_fwd(table: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    vocab, d = table.shape
    out = torch.empty((*idx.shape, d), device=table.device, dtype=table.dtype)
    call("hyp_embedding_fwd", ptr(table), ptr(idx), ptr(out), idx.numel(), d, vocab, DT[table.dtype], stream_ptr())
    return out


def embedding_bwd(dout: torch.Tensor, idx: torch.Tensor, dtable: torch.Tensor, padding_idx: int = -1) -> None:
    vocab, d = dtable.shape
    assert dtable.dtype == torch.float32
    call("hyp_embedding_bwd", ptr(dout), ptr(idx), ptr(dtable), idx.numel(), d, vocab, int(padding_idx),
         DT[dout.dtype], stream_ptr())


def rope_(x: torch.Tensor, theta: float = 10000.0, positions: torch.Tensor | None = None, inverse: bool = False,
          layout: str = "bshd") -> torch.Tensor:
    """In-place rotary embedding on a 4-D tensor (``bshd`` or ``bhsd``) with unit stride on the last dim."""
    assert x.dim() == 4 and x.stride(3) == 1
    if layout == "bshd":
        B, S, H, D = x.shape
        sb, ss, sh = x.stride(0), x.stride(1), x.stride(2)
    else:
        B, H, S, D = x.shape
        sb, sh, ss = x.stride(0), x.stride(1), x.stride(2)
    call("hyp_rope", ptr(x), B, S, H, D, sb, ss, sh, ptr(positions), float(theta), int(inverse), DT[x.dtype],
         stream_ptr())
    return x
