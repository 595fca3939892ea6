"""Autograd-aware ops.  On a CUDA tensor they run the hand-written sm_100a kernels
(``libhyperion_b200.so``); on a CPU tensor they run the plain PyTorch formulation
of the same maths, which is both the oracle of the GPU tests and the code path of
the gloo/CPU configuration (BASELINE config 1).

There is no silent GPU fallback: a CUDA tensor with the native library missing
raises (see :func:`_native`).
"""
from __future__ import annotations

import os

import torch
import torch.nn.functional as F

from . import _lib, gradslot, native, rng

_SUPPORTED = (torch.float32, torch.float16, torch.bfloat16)

# fp32 tensors: "tf32" runs them on the tensor cores (tcgen05 kind::tf32, 10-bit mantissa products, fp32 accumulation);
# "ieee" keeps true fp32 arithmetic by leaving fp32 tensors to the library path with TF32 off -- what the reference's fp32
# baseline benchmark computes (baseline_performance.ipynb c0 L92-101).  HYPERION_FP32 / set_fp32_math() select it.
_FP32_MATH = os.environ.get("HYPERION_FP32", "tf32")


def set_fp32_math(mode: str) -> None:
    global _FP32_MATH
    assert mode in ("tf32", "ieee")
    _FP32_MATH = mode
    if mode == "ieee":
        torch.backends.cuda.matmul.allow_tf32 = False


def fp32_math() -> str:
    return _FP32_MATH


def _native(*tensors) -> bool:
    """True -> run the sm_100a kernels.  CUDA tensors without the library are an error."""
    t0 = next(t for t in tensors if t is not None)
    if not t0.is_cuda:
        return False
    if not _lib.available():
        if torch.cuda.is_available() and _lib.raw() is None:
            raise RuntimeError("CUDA tensor given but libhyperion_b200.so is not built; run `python -m hyperion_b200.build`")
        return False
    if t0.dtype == torch.float32 and _FP32_MATH == "ieee":
        return False
    return t0.dtype in _SUPPORTED


def _ensure_gathered(*weights) -> None:
    """Library fallback paths (CPU, misaligned K, disabled native library) read the weights directly: if one of them belongs
    to an FSDP unit that was waiting to be gathered inside its first GEMM, gather it now with the stand-alone kernel --
    otherwise the fallback would silently compute with last step's parameters."""
    for w in weights:
        h = getattr(w, "_hyp_ag", None) if w is not None else None
        if h is not None and h.pending:
            h.fsdp.sync_params()
            return


def _as2d(t: torch.Tensor) -> torch.Tensor:
    """2-D row-major view usable as a TMA operand (16-byte aligned rows), copying only if needed."""
    t2 = t.reshape(-1, t.shape[-1])
    es = t2.element_size()
    ok = t2.stride(1) == 1 and (t2.stride(0) * es) % 16 == 0 and t2.data_ptr() % 16 == 0
    if ok:
        return t2
    return _padded_copy(t2)


def _padded_empty(rows: int, cols: int, dtype, device) -> torch.Tensor:
    """[rows, cols] view of a buffer whose leading dimension is padded to 16 bytes (128 bytes for wide ragged rows,
    so that the GEMM can address the tensor MN-major through its rank-3 tensor map)."""
    per = 16 // torch.empty((), dtype=dtype).element_size()
    if cols > 256 and cols % 64 != 0:
        per = 64
    ld = (cols + per - 1) // per * per
    if ld == cols:
        return torch.empty((rows, cols), dtype=dtype, device=device)
    return torch.empty((rows, ld), dtype=dtype, device=device)[:, :cols]


def _padded_copy(t2: torch.Tensor) -> torch.Tensor:
    out = _padded_empty(t2.shape[0], t2.shape[1], t2.dtype, t2.device)
    out.copy_(t2)
    return out


_SM_COUNT = 148
_CE_LOGITS_BUDGET = 2 << 30  # bytes of logits linear_cross_entropy may hold at once


def _ce_l2_rows(vocab: int, es: int) -> int:
    """Rows per slice of the logits in linear_cross_entropy; 0 (default) = one slice.  ``HYPERION_CE_L2_ROWS=n`` processes
    the head in n-row slices so a slice stays in the 126 MB L2 between the GEMM that writes it, the CE kernel and the
    dgrad / bias-sum passes.  Measured on the reference LM (4064 tokens x 50257, fp16; step 1.55 ms unsliced): 1024-row
    slices 1.71 ms, 256-row slices 2.32 ms -- the small-M GEMMs and the extra launches cost more than the HBM traffic they
    save, so slicing stays off."""
    v = os.environ.get("HYPERION_CE_L2_ROWS")
    return int(v) if v is not None else 0


def _wgrad(dy2: torch.Tensor, x2: torch.Tensor, out_dtype, slot=None) -> torch.Tensor:
    """dW[N, K] = dy2[M, N]^T @ x2[M, K] (both operands MN-major).  When the output has few tiles (N*K is small next
    to the token count M) the reduction is split across SMs: split-K partial tiles are reduced with vector fp32
    atomics, then rounded to the parameter dtype.  ``slot``: the weight's gradient slot (ops/gradslot.py) -- when it
    can be claimed the result is stored straight into the flat gradient buffer."""
    N, K = dy2.shape[1], x2.shape[1]
    splits = 1
    if min(N, K) >= 256:
        # 256 x 256 tiles on CTA pairs: split only when the tiles leave more than half of the pairs idle (measured on
        # GPT-2-medium shapes: 48 tiles run as fast unsplit, 16 tiles are 1.7x faster with 4 splits)
        tiles, pairs = ((N + 255) // 256) * ((K + 255) // 256), _SM_COUNT // 2
        if 2 * tiles <= pairs and dy2.shape[0] >= 2048:
            # up to 4 splits at GPT-2 sizes (tuned there); a handful of tiles (the 256-wide layers of the reference LM,
            # 1-8 tiles over 4064 tokens) were latency-bound at ~17 us whatever their size: give each split ~512 tokens
            cap = 4 if tiles >= 9 else 8
            splits = max(1, min(cap, pairs // tiles, dy2.shape[0] // (1024 if tiles >= 9 else 512)))
    else:
        tiles = ((N + 127) // 128) * ((K + 255) // 256 if K > 128 else 1)
        if tiles < 120 and dy2.shape[0] >= 2048:
            splits = min(8, max(1, (2 * _SM_COUNT) // tiles), dy2.shape[0] // 512)
    es = torch.empty((), dtype=out_dtype).element_size()
    dst = gradslot.claim(slot, out_dtype) if (K * es) % 16 == 0 else None
    if dst is not None:
        dst = dst.view(N, K)
    if splits > 1 and (K * 4) % 16 == 0:
        acc = torch.zeros((N, K), device=dy2.device, dtype=torch.float32)
        native.gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=acc, splits=splits)
        if out_dtype == torch.float32:
            return acc if dst is None else dst.copy_(acc)
        return native.cast(acc, out_dtype, out=dst)
    out = dst if dst is not None else _padded_empty(N, K, out_dtype, dy2.device)
    native.gemm(dy2, x2, a_mn_major=True, b_mn_major=True, out=out)
    return out if out.is_contiguous() else out.contiguous()


def _dgrad(dh2: torch.Tensor, w2: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """dx[M, K] = dh2[M, N] @ w2[N, K].  When the output has few tiles and the reduction dimension N is long (the
    vocabulary projection of a narrow model: 4064 x 256 outputs reduced over 50257 columns ran on 32 of 74 CTA pairs),
    the reduction is split across SMs into an fp32 accumulator and rounded once."""
    M, N = dh2.shape
    K = w2.shape[1]
    bn = 256 if K >= 256 else 128
    tiles, pairs = ((M + 255) // 256) * ((K + bn - 1) // bn), _SM_COUNT // 2
    splits = 1
    if 2 * tiles <= pairs and N >= 8192 and (K * 4) % 16 == 0:
        splits = max(1, min(8, pairs // tiles, N // 4096))
    if splits > 1:
        acc = torch.zeros((M, K), device=dh2.device, dtype=torch.float32)
        native.gemm(dh2, w2, b_mn_major=True, out=acc, splits=splits)
        return native.cast(acc, dh2.dtype, out=out)
    if out is None:
        out = _padded_empty(M, K, dh2.dtype, dh2.device)
    return native.gemm(dh2, w2, b_mn_major=True, out=out)


def _bias_grad(dy2: torch.Tensor, dtype, slot=None, dy=None) -> torch.Tensor:
    """db[N] = column sums of dy2 (fp32 accumulation), rounded into the bias's gradient slot when it can be claimed.
    ``dy``: the gradient tensor as autograd delivered it -- when the norm backward that produced it attached its column
    sums (``_hyp_colsum``), they are used instead of another pass over the tensor."""
    N = dy2.shape[1]
    cs = getattr(dy, "_hyp_colsum", None) if dy is not None else None
    if cs is not None and cs.numel() == N and cs.dtype == dtype and dy.shape[-1] == N:
        dst = gradslot.claim(slot, dtype)
        return cs if dst is None else dst.copy_(cs)
    if dtype == dy2.dtype and dtype != torch.float32 and N % 8 == 0 and dy2.stride(0) % 8 == 0 and dy2.data_ptr() % 16 == 0:
        return native.colsum_cast(dy2, out=gradslot.claim(slot, dtype))  # one launch: sum, round, store
    acc = native.colsum(dy2)
    if dtype == torch.float32:
        dst = gradslot.claim(slot, dtype)
        return acc if dst is None else dst.copy_(acc)
    return native.cast(acc, dtype, out=gradslot.claim(slot, dtype))


def _act_ref(x, act):
    if act in (None, "none"):
        return x
    if act == "relu":
        return F.relu(x)
    if act in ("gelu", "gelu_erf"):
        return F.gelu(x)
    if act == "gelu_tanh":
        return F.gelu(x, approximate="tanh")
    if act == "silu":
        return F.silu(x)
    raise ValueError(act)


# --------------------------------------------------------------------------- linear
class _LinearFn(torch.autograd.Function):
    """y = act(x @ W^T + b) (+ residual) on gemm_tcgen05; dgrad / wgrad use the MN-major operand paths."""

    @staticmethod
    def forward(ctx, x, w, b, residual, act, ag=None):
        x2 = _as2d(x)
        w2 = _as2d(w)
        M, N = x2.shape[0], w2.shape[0]
        out = _padded_empty(M, N, x.dtype, x.device)
        res2 = _as2d(residual) if residual is not None else None
        pre = _padded_empty(M, N, x.dtype, x.device) if act not in (None, "none", "relu") else None
        if ag is not None and ag.pending:
            ag.run(x2, w2, out, bias=b, aux=res2, act=act, preact_out=pre)  # K-C3: the unit is gathered inside this GEMM
        else:
            native.gemm(x2, w2, out=out, bias=b, aux=res2, act=act, preact_out=pre)
        ctx.act = act
        ctx.has_bias = b is not None
        ctx.has_res = residual is not None
        ctx.x_shape = x.shape
        ctx.slots = (gradslot.slot_of(w), gradslot.slot_of(b) if b is not None else None)
        # relu'(h) is recoverable from the output only when no residual was added
        if act == "relu":
            if residual is not None:
                raise NotImplementedError("relu + residual in one linear epilogue is not differentiable from the output")
            ctx.save_for_backward(x2, w2, out)
        else:
            ctx.save_for_backward(x2, w2, pre)
        return out.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, w2, h = ctx.saved_tensors
        dy2 = _as2d(dy)
        act = ctx.act
        if act not in (None, "none"):
            # dh = dy * act'(h)   (one elementwise pass; the fused two-layer MLP op avoids it)
            dh = _dact_mul(dy2, h, act)
        else:
            dh = dy2
        dx = dw = db = dres = None
        if ctx.needs_input_grad[0]:
            dx = _dgrad(dh, w2).reshape(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            dw = _wgrad(dh, x2, w2.dtype, ctx.slots[0])
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = _bias_grad(dh, dy.dtype, ctx.slots[1], dy if dh is dy2 else None)
        if ctx.has_res and ctx.needs_input_grad[3]:
            dres = dy
        return dx, dw, db, dres, None, None


def _dact_mul(dy2, h, act):
    """dy * act'(h) elementwise (torch ops; only used by the un-fused single linear+activation)."""
    hf = h.float()
    if act == "relu":
        g = (hf > 0).float()
    elif act in ("gelu", "gelu_erf"):
        g = 0.5 * (1 + torch.erf(hf * 0.7071067811865476)) + hf * torch.exp(-0.5 * hf * hf) * 0.3989422804014327
    elif act == "gelu_tanh":
        u = 0.7978845608028654 * (hf + 0.044715 * hf ** 3)
        t = torch.tanh(u)
        g = 0.5 * (1 + t) + 0.5 * hf * (1 - t * t) * 0.7978845608028654 * (1 + 3 * 0.044715 * hf * hf)
    elif act == "silu":
        s = torch.sigmoid(hf)
        g = s * (1 + hf * (1 - s))
    else:
        raise ValueError(act)
    return (dy2.float() * g).to(dy2.dtype)


def linear(x, w, b=None, act: str | None = None, residual=None):
    """``act(x @ w.T + b) + residual`` with the bias / activation / residual fused into the GEMM epilogue."""
    if _native(x, w) and x.shape[-1] % (16 // x.element_size()) == 0:
        return _LinearFn.apply(x, w, b, residual, act, getattr(w, "_hyp_ag", None))
    _ensure_gathered(w)
    y = _act_ref(F.linear(x, w, b), act)
    return y if residual is None else y + residual


# --------------------------------------------------------------------------- fused MLP
class _MlpFn(torch.autograd.Function):
    """y = dropout(act(x W1^T + b1)) W2^T + b2 (+ residual); backward fuses act' into the dgrad GEMM epilogue and
    re-applies the regenerated dropout mask in place (mask and act' are both element-wise, so the order is free)."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2, residual, act, ag1=None, ag2=None, dropout_p=0.0):
        x2 = _as2d(x)
        M = x2.shape[0]
        h = torch.empty((M, w1.shape[0]), dtype=x.dtype, device=x.device)
        a = torch.empty_like(h)
        if ag1 is not None and ag1.pending:
            ag1.run(x2, w1, a, bias=b1, act=act, preact_out=h)
        else:
            native.gemm(x2, w1, out=a, bias=b1, act=act, preact_out=h)
        ctx.drop = None
        if dropout_p > 0.0:
            ctx.drop = (rng.state(x.device), rng.next_salt(), float(dropout_p))
            native.dropout(a, *ctx.drop, out=a)      # in place: only the dropped activation is needed again (wgrad of W2)
        y = torch.empty((M, w2.shape[0]), dtype=x.dtype, device=x.device)
        res2 = _as2d(residual) if residual is not None else None
        if ag2 is not None and ag2.pending:
            ag2.run(a, w2, y, bias=b2, aux=res2)
        else:
            native.gemm(a, w2, out=y, bias=b2, aux=res2)
        ctx.save_for_backward(x2, w1, w2, h, a)
        ctx.act = act
        ctx.flags = (b1 is not None, b2 is not None, residual is not None)
        ctx.slots = tuple(gradslot.slot_of(t) if t is not None else None for t in (w1, b1, w2, b2))
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], w2.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, w1, w2, h, a = ctx.saved_tensors
        dy2 = _as2d(dy)
        has_b1, has_b2, has_res = ctx.flags
        # dh = (dy @ W2) * act'(h): activation derivative applied in the GEMM epilogue
        # the bias gradient of the first layer = column sums of dh: produced by the same epilogue when dh is final there
        # (no dropout to re-apply afterwards) and the shape takes the staged epilogue for every tile
        N1 = w2.shape[1]
        fuse_db1 = (has_b1 and ctx.drop is None and dy2.dtype != torch.float32 and N1 % 256 == 0
                    and os.environ.get("HYPERION_GEMM_COLSUM", "1") == "1")
        cs1 = torch.zeros(N1, device=dy2.device, dtype=torch.float32) if fuse_db1 else None
        dh = native.gemm(dy2, w2, b_mn_major=True, aux=h, act=ctx.act, mode=1, colsum=cs1)
        if ctx.drop is not None:
            native.dropout(dh, *ctx.drop, out=dh)
        s_w1, s_b1, s_w2, s_b2 = ctx.slots
        dw2 = _wgrad(dy2, a, w2.dtype, s_w2)
        dw1 = _wgrad(dh, x2, w1.dtype, s_w1)
        dx = native.gemm(dh, w1, b_mn_major=True).reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        if fuse_db1:
            db1 = native.cast(cs1, dy.dtype, out=gradslot.claim(s_b1, dy.dtype))
        else:
            db1 = _bias_grad(dh, dy.dtype, s_b1) if has_b1 else None
        db2 = _bias_grad(dy2, dy.dtype, s_b2, dy) if has_b2 else None
        return dx, dw1, db1, dw2, db2, (dy if has_res else None), None, None, None, None


def mlp(x, w1, b1, w2, b2, act: str = "relu", residual=None, dropout_p: float = 0.0):
    """Two-layer feed-forward block (the FFN of every transformer in the model zoo); ``dropout_p`` drops the hidden
    activation (``dropout`` inside ``nn.TransformerEncoderLayer``'s FFN) without leaving the fused op."""
    es = x.element_size()
    if _native(x, w1) and x.shape[-1] % (16 // es) == 0 and w1.shape[0] % (16 // es) == 0 and w2.shape[0] % (16 // es) == 0:
        if dropout_p > 0.0 and not native_dropout_enabled():
            h = F.dropout(linear(x, w1, b1, act=act), dropout_p, True)
            return linear(h, w2, b2, residual=residual)
        return _MlpFn.apply(x, w1, b1, w2, b2, residual, act, getattr(w1, "_hyp_ag", None), getattr(w2, "_hyp_ag", None),
                            dropout_p)
    _ensure_gathered(w1, w2)
    h = _act_ref(F.linear(x, w1, b1), act)
    if dropout_p > 0.0:
        h = F.dropout(h, dropout_p, True)
    y = F.linear(h, w2, b2)
    return y if residual is None else y + residual


# --------------------------------------------------------------------------- norms
def _norm_param_grads(ctx, dy, dres):
    """Shared backward of the norm ops: dz (+ dres), dgamma / dbeta stored into their gradient slots when claimable."""
    z, gamma, mean, rstd = ctx.saved_tensors
    s_g, s_b = ctx.slots
    g_out = gradslot.claim(s_g, gamma.dtype)
    b_out = gradslot.claim(s_b, gamma.dtype) if ctx.has_beta else None
    # dz is usually the output gradient of the linear layer that fed this norm: hand its column sums (= that layer's
    # bias gradient) along with the tensor so the layer does not have to re-read dz (see _bias_grad)
    cs = None
    if g_out is not None and g_out.dtype == z.dtype and z.shape[-1] <= 2048:
        cs = torch.empty(z.shape[-1], device=z.device, dtype=z.dtype)
    dz, dgamma, dbeta = native.norm_bwd(dy.contiguous(), z, gamma, mean, rstd, rms=ctx.rms, need_beta=ctx.has_beta,
                                        dres=dres.contiguous() if dres is not None else None,
                                        dgamma_out=g_out, dbeta_out=b_out, dz_colsum_out=cs)
    if cs is not None:
        dz._hyp_colsum = cs
    dgamma = dgamma if dgamma.dtype == gamma.dtype else dgamma.to(gamma.dtype)
    if dbeta is not None and dbeta.dtype != gamma.dtype:
        dbeta = dbeta.to(gamma.dtype)
    return dz, dgamma, dbeta


class _NormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, eps, rms, dropout_p=0.0):
        xc = x.contiguous()
        rc = residual.contiguous() if residual is not None else None
        ctx.drop = None
        if dropout_p > 0.0:
            # z = dropout(x) + residual in the norm kernel's own pass; the mask is regenerated in backward
            ctx.drop = (rng.state(x.device), rng.next_salt(), float(dropout_p))
        y, z, mean, rstd = native.norm_fwd(xc, gamma, beta, residual=rc, eps=eps, rms=rms, drop=ctx.drop)
        ctx.save_for_backward(z, gamma, mean, rstd)
        ctx.rms = rms
        ctx.has_beta = beta is not None
        ctx.has_res = residual is not None
        ctx.slots = (gradslot.slot_of(gamma), gradslot.slot_of(beta) if beta is not None else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        dz, dgamma, dbeta = _norm_param_grads(ctx, dy, None)
        dx = dz
        if ctx.drop is not None:
            st, salt, p = ctx.drop
            dx = native.dropout(dz.contiguous(), st, salt, p)
        return dx, dgamma, dbeta, dz if ctx.has_res else None, None, None, None


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p):
        ctx.drop = (rng.state(x.device), rng.next_salt(), float(p))
        return native.dropout(x.contiguous(), *ctx.drop)

    @staticmethod
    def backward(ctx, dy):
        return native.dropout(dy.contiguous(), *ctx.drop), None


def native_dropout_enabled() -> bool:
    """The Philox dropout kernels (attention probabilities, residual branches, FFN hidden, LoRA input) are the default;
    ``HYPERION_DROPOUT=torch`` selects the library path (F.dropout / SDPA) for A/B runs."""
    import os

    return os.environ.get("HYPERION_DROPOUT", "native") == "native"


def dropout(x, p: float, training: bool = True):
    """``F.dropout`` on the native Philox kernel (mask regenerated in backward, nothing stored).  The effective drop
    probability is ``round(256 p) / 256`` and survivors are scaled accordingly (unbiased)."""
    if not training or p <= 0.0:
        return x
    if _native(x) and native_dropout_enabled() and x.dim() >= 1 and x.numel() > 0:
        return _DropoutFn.apply(x, p)
    return F.dropout(x, p, True)


class _NormForkFn(torch.autograd.Function):
    """(norm(x), x): the second output is x itself, handed on to the skip connection of a pre-norm block.  Owning the
    fork lets the backward add the skip-connection gradient to dz inside the norm-backward kernel instead of leaving
    the sum of the two branches to a separate autograd elementwise pass."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps, rms):
        xc = x.contiguous()
        y, _, mean, rstd = native.norm_fwd(xc, gamma, beta, residual=None, eps=eps, rms=rms)
        ctx.save_for_backward(xc, gamma, mean, rstd)
        ctx.rms = rms
        ctx.has_beta = beta is not None
        ctx.slots = (gradslot.slot_of(gamma), gradslot.slot_of(beta) if beta is not None else None)
        return y, x.view_as(x)

    @staticmethod
    def backward(ctx, dy, dskip):
        if dy is None:  # only the skip branch was used
            return dskip, None, None, None, None
        dz, dgamma, dbeta = _norm_param_grads(ctx, dy, dskip)
        return dz, dgamma, dbeta, None, None


def layer_norm(x, gamma, beta, eps: float = 1e-5, residual=None, dropout_p: float = 0.0):
    """LayerNorm(dropout(x) + residual) in one pass: the post-norm residual add -- and, in training, the dropout of the
    residual branch (``dropout_p``) -- are fused into the norm kernel."""
    d = x.shape[-1]
    if _native(x, gamma) and d % 8 == 0 and d <= 8192:
        if dropout_p > 0.0 and not native_dropout_enabled():
            x, dropout_p = F.dropout(x, dropout_p, True), 0.0
        return _NormFn.apply(x, gamma, beta, residual, eps, False, dropout_p)
    if dropout_p > 0.0:
        x = F.dropout(x, dropout_p, True)
    z = x if residual is None else x + residual
    return F.layer_norm(z, (d,), gamma, beta, eps)


def layer_norm_fork(x, gamma, beta, eps: float = 1e-5):
    """``(LayerNorm(x), x)`` for a pre-norm residual block: feed the second value to the skip connection
    (``residual=`` of the block's last linear) so the backward sums both branches inside the norm kernel."""
    d = x.shape[-1]
    if _native(x, gamma) and d % 8 == 0 and d <= 8192:
        return _NormForkFn.apply(x, gamma, beta, eps, False)
    return F.layer_norm(x, (d,), gamma, beta, eps), x


def rms_norm_fork(x, gamma, eps: float = 1e-6):
    """RMSNorm flavour of :func:`layer_norm_fork`."""
    d = x.shape[-1]
    if _native(x, gamma) and d % 8 == 0 and d <= 8192:
        return _NormForkFn.apply(x, gamma, None, eps, True)
    return rms_norm(x, gamma, eps), x


def rms_norm(x, gamma, eps: float = 1e-6, residual=None):
    d = x.shape[-1]
    if _native(x, gamma) and d % 8 == 0 and d <= 8192:
        return _NormFn.apply(x, gamma, None, residual, eps, True, 0.0)
    z = x if residual is None else x + residual
    zf = z.float()
    return (zf * torch.rsqrt(zf.pow(2).mean(-1, keepdim=True) + eps)).to(x.dtype) * gamma


# --------------------------------------------------------------------------- embedding
class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, idx, table, padding_idx):
        ctx.save_for_backward(idx)
        ctx.shape = table.shape
        ctx.dtype = table.dtype
        ctx.padding_idx = -1 if padding_idx is None else padding_idx
        ctx.slot = gradslot.slot_of(table)
        return native.embedding_fwd(table, idx.contiguous())

    @staticmethod
    def backward(ctx, dout):
        (idx,) = ctx.saved_tensors
        dt = torch.zeros(ctx.shape, device=dout.device, dtype=torch.float32)
        native.embedding_bwd(dout.contiguous(), idx.contiguous(), dt, ctx.padding_idx)
        if ctx.dtype == torch.float32:
            return None, dt, None
        return None, native.cast(dt, ctx.dtype, out=gradslot.claim(ctx.slot, ctx.dtype)), None


def embedding(idx, table, padding_idx=None):
    if _native(table) and (table.shape[1] * table.element_size()) % 16 == 0:
        return _EmbeddingFn.apply(idx, table, padding_idx)
    return F.embedding(idx, table, padding_idx)


# --------------------------------------------------------------------------- linear + cross-entropy
class _LinearCEFn(torch.autograd.Function):
    """mean CE(x @ W^T + b, targets) with the loss AND all three gradients produced in the forward.

    The vocabulary projection writes 16-bit logits, the fused CE kernel turns them into their gradient in place, and the
    dgrad / wgrad GEMMs consume that gradient right away.  The [tokens, vocab] matrix is held while that happens (up to
    ``_CE_LOGITS_BUDGET``; beyond it the rows are processed in chunks with an fp32 wgrad accumulator) -- see the comment
    in ``forward`` for why the recompute variant that avoids it is not used, and for the L2-resident row slices.
    """

    @staticmethod
    def forward(ctx, x, w, b, targets, ignore_index, chunk_rows):
        x2 = _as2d(x)
        w2 = _as2d(w)
        M, K = x2.shape
        V = w2.shape[0]
        dev = x.device
        tgt = targets.reshape(-1).contiguous()
        sum_count = torch.zeros(2, device=dev, dtype=torch.float32)
        inv_cnt = (tgt != ignore_index).sum().clamp(min=1).to(torch.float32).reciprocal().reshape(1)
        dx = torch.empty_like(x2)
        # The logits matrix (tokens x vocab, 16-bit) IS held when it is affordable (<= 2 GiB; 0.8 GB for GPT-2 medium at 8 x
        # 1024 tokens): the head's wgrad is then a single GEMM with a 16-bit output instead of fp32 accumulation across
        # chunks (which reads and writes V x K fp32 per chunk).  A recompute design (online LSE in the forward epilogue,
        # logits rebuilt in backward) would save that memory but costs a second vocabulary GEMM (0.6 ms at GPT-2-medium
        # size) to save ~0.25 ms of HBM traffic, so it is not what runs.  Optional L2-sized row slices: see _ce_l2_rows.
        single = M * V * x.element_size() <= _CE_LOGITS_BUDGET
        if single:
            chunk_rows = M
        dw = None if single else torch.zeros((V, K), device=dev, dtype=torch.float32)
        db = torch.zeros(V, device=dev, dtype=torch.float32) if b is not None else None
        ag = getattr(w, "_hyp_ag", None)
        l2_rows = _ce_l2_rows(V, x.element_size())
        for r0 in range(0, M, chunk_rows):
            r1 = min(M, r0 + chunk_rows)
            logits = _padded_empty(r1 - r0, V, x.dtype, dev)
            step = l2_rows if (single and 0 < l2_rows < r1 - r0) else (r1 - r0)
            for q0 in range(r0, r1, step):
                q1 = min(r1, q0 + step)
                lg = logits[q0 - r0:q1 - r0]
                if ag is not None and ag.pending:
                    ag.run(x2[q0:q1], w2, lg, bias=b)  # first slice gathers the head's unit inside its GEMM
                else:
                    native.gemm(x2[q0:q1], w2, out=lg, bias=b)
                native.ce_fwd_bwd(lg, tgt[q0:q1], ignore_index, sum_count, gscale_ptr=inv_cnt, n_classes=V)
                _dgrad(lg, w2, out=dx[q0:q1])
                if db is not None:
                    native.colsum(lg, out=db)
            if single:
                dw = _wgrad(logits, x2, w.dtype)
            else:
                native.gemm(logits, x2[r0:r1], a_mn_major=True, b_mn_major=True, out=dw, beta=1.0)
        loss = torch.empty((), device=dev, dtype=torch.float32)
        native.ce_finalize(sum_count, loss)
        ctx.save_for_backward(dx, dw, db)
        ctx.x_shape, ctx.w_dtype = x.shape, w.dtype
        ctx.slots = (gradslot.slot_of(w), gradslot.slot_of(b) if b is not None else None)
        return loss

    @staticmethod
    def backward(ctx, dloss):
        dx, dw, db = ctx.saved_tensors
        s = dloss.to(torch.float32).reshape(1)
        gx = (dx * s.to(dx.dtype)).reshape(ctx.x_shape)
        # scale by the incoming gradient and round in one pass, into the head's gradient slot when it is free
        gw = native.cast(dw, ctx.w_dtype, scale_t=s, out=gradslot.claim(ctx.slots[0], ctx.w_dtype))
        gb = native.cast(db, ctx.w_dtype, scale_t=s, out=gradslot.claim(ctx.slots[1], ctx.w_dtype)) if db is not None else None
        return gx, gw, gb, None, None, None


def linear_cross_entropy(x, w, b, targets, ignore_index: int = -100, chunk_rows: int = 4096):
    """Mean cross-entropy of ``x @ w.T + b`` against ``targets`` (fp32 scalar)."""
    if _native(x, w) and x.shape[-1] % (16 // x.element_size()) == 0:
        return _LinearCEFn.apply(x, w, b, targets, ignore_index, chunk_rows)
    _ensure_gathered(w)
    logits = F.linear(x, w, b)
    return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), targets.reshape(-1), ignore_index=ignore_index)


class _CEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, targets, ignore_index):
        l2 = _padded_copy(logits.reshape(-1, logits.shape[-1]))
        tgt = targets.reshape(-1).contiguous()
        sc = torch.zeros(2, device=logits.device, dtype=torch.float32)
        inv_cnt = (tgt != ignore_index).sum().clamp(min=1).to(torch.float32).reciprocal().reshape(1)
        native.ce_fwd_bwd(l2, tgt, ignore_index, sc, gscale_ptr=inv_cnt, n_classes=logits.shape[-1])
        loss = torch.empty((), device=logits.device, dtype=torch.float32)
        native.ce_finalize(sc, loss)
        ctx.save_for_backward(l2)
        ctx.shape = logits.shape
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (g,) = ctx.saved_tensors
        return (g * dloss.to(g.dtype)).reshape(ctx.shape), None, None


def cross_entropy(logits, targets, ignore_index: int = -100):
    if _native(logits):
        return _CEFn.apply(logits, targets, ignore_index)
    return F.cross_entropy(logits.reshape(-1, logits.shape[-1]).float(), targets.reshape(-1), ignore_index=ignore_index)


# --------------------------------------------------------------------------- rope / swiglu
class _RopeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, theta, positions, layout):
        ctx.theta, ctx.layout = theta, layout
        ctx.save_for_backward(positions) if positions is not None else None
        ctx.has_pos = positions is not None
        y = x.contiguous().clone()
        return native.rope_(y, theta, positions, inverse=False, layout=layout)

    @staticmethod
    def backward(ctx, dy):
        pos = ctx.saved_tensors[0] if ctx.has_pos else None
        g = dy.contiguous().clone()
        return native.rope_(g, ctx.theta, pos, inverse=True, layout=ctx.layout), None, None, None


def rope(x, theta: float = 10000.0, positions=None, layout: str = "bshd"):
    """Rotate-half rotary position embedding on ``[B,S,H,D]`` (``bshd``) or ``[B,H,S,D]`` (``bhsd``)."""
    if _native(x):
        return _RopeFn.apply(x, theta, positions, layout)
    D = x.shape[-1]
    S = x.shape[1] if layout == "bshd" else x.shape[2]
    inv_freq = 1.0 / (theta ** (torch.arange(0, D, 2, device=x.device, dtype=torch.float32) / D))
    if positions is None:
        pos = torch.arange(S, device=x.device, dtype=torch.float32)[None, :].expand(x.shape[0], S)
    else:
        pos = positions.float()
    ang = pos[..., None] * inv_freq  # [B,S,D/2]
    cos = torch.cat([ang, ang], -1).cos()
    sin = torch.cat([ang, ang], -1).sin()
    if layout == "bshd":
        cos, sin = cos[:, :, None, :], sin[:, :, None, :]
    else:
        cos, sin = cos[:, None, :, :], sin[:, None, :, :]
    xf = x.float()
    rot = torch.cat([-xf[..., D // 2:], xf[..., : D // 2]], -1)
    return (xf * cos + rot * sin).to(x.dtype)


class _SwigluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, gate, up):
        g, u = gate.contiguous(), up.contiguous()
        ctx.save_for_backward(g, u)
        return native.swiglu_fwd(g, u)

    @staticmethod
    def backward(ctx, dout):
        g, u = ctx.saved_tensors
        return native.swiglu_bwd(g, u, dout.contiguous())


def swiglu(gate, up):
    if _native(gate) and gate.dtype != torch.float32 and gate.numel() % 8 == 0:
        return _SwigluFn.apply(gate, up)
    return F.silu(gate) * up


# --------------------------------------------------------------------------- attention
def attention(q, k, v, causal: bool = False, key_padding_mask=None, dropout_p: float = 0.0, scale: float | None = None):
    """softmax(q k^T * scale + mask) v for ``[B, H, S, D]`` tensors.

    key_padding_mask: bool ``[B, S_k]``, True = attend.  Dispatches to the
    tcgen05 flash kernel (:mod:`hyperion_b200.ops.attn`) when it applies.
    """
    from . import attn as _attn

    return _attn.attention(q, k, v, causal=causal, key_padding_mask=key_padding_mask, dropout_p=dropout_p, scale=scale)


# --------------------------------------------------------------------------- LoRA
def _lora_pads(A: torch.Tensor, B: torch.Tensor):
    """Zero-padded 64-wide copies of the adapter matrices for the dual-K GEMM: A_pad [64, in], At_pad [in, 64],
    B_pad [out, 64].  The buffers are cached on the parameter (zeros outside the rank-r block) and refreshed by copy."""
    r = A.shape[0]
    pads = getattr(A, "_hyp_pads", None)
    if pads is None or pads[0].shape[1] != A.shape[1] or pads[2].shape[0] != B.shape[0] or pads[0].dtype != A.dtype:
        pads = (torch.zeros(64, A.shape[1], device=A.device, dtype=A.dtype),
                torch.zeros(A.shape[1], 64, device=A.device, dtype=A.dtype),
                torch.zeros(B.shape[0], 64, device=A.device, dtype=A.dtype))
        try:
            A._hyp_pads = pads
        except AttributeError:
            pass
    with torch.no_grad():
        pads[0][:r].copy_(A)
        pads[1][:, :r].copy_(A.t())
        pads[2][:, :r].copy_(B)
    return pads


class _LoraLinearFn(torch.autograd.Function):
    """y = x W^T + b + scaling * (xd A^T) B^T (+ residual); W is usually frozen.

    With enough rows for the pair kernel (M >= 256) the base projection and the rank-r update are ONE dual-K GEMM
    (``hyp_gemm_dualk``): t = s * xd A^T is computed first (zero-padded to 64 columns), then the reduction of the main
    GEMM simply continues over (t, B) into the same TMEM accumulator; the backward does the same for
    dx = dy W + dt A.  Otherwise: three launches of the plain GEMM (beta-accumulate)."""

    @staticmethod
    def forward(ctx, x, xd, w, b, A, B, scaling, residual):
        x2, xd2 = _as2d(x), _as2d(xd)
        same = xd is x
        M, N, r = x2.shape[0], w.shape[0], A.shape[0]
        res2 = _as2d(residual) if residual is not None else None
        dual = M >= 256 and N > 128 and r <= 64 and x.dtype != torch.float32 and w.dtype == x.dtype
        if dual:
            A_pad, At_pad, B_pad = _lora_pads(A, B)
            t_pad = native.gemm(xd2, A_pad, alpha=scaling)          # [M, 64], columns >= r are zero
            y = native.gemm_dualk(x2, w, t_pad, B_pad, bias=b, aux=res2)
            t = t_pad[:, :r]
            ctx.pads = (At_pad, B_pad)
        else:
            t = native.gemm(xd2, A, alpha=scaling)  # [M, r], scaling folded in
            y = torch.empty((M, N), dtype=x.dtype, device=x.device)
            native.gemm(x2, w, out=y, bias=b, aux=res2)
            native.gemm(t, B, out=y, beta=1.0)  # rank-r update accumulated into the same output tile
            ctx.pads = None
        ctx.save_for_backward(x2, xd2, w, A, B, t)
        ctx.scaling, ctx.same = scaling, same
        ctx.has_bias, ctx.has_res = b is not None, residual is not None
        ctx.slots = (gradslot.slot_of(A), gradslot.slot_of(B))
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy):
        x2, xd2, w, A, B, t = ctx.saved_tensors
        dy2 = _as2d(dy)
        need_x, need_xd, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1], ctx.needs_input_grad[2]
        r = A.shape[0]
        if ctx.pads is not None:
            At_pad, B_pad = ctx.pads
            dt_pad = native.gemm(dy2, B_pad, b_mn_major=True, alpha=ctx.scaling)   # [M, 64]
            dt = dt_pad[:, :r]
        else:
            dt_pad = None
            dt = native.gemm(dy2, B, b_mn_major=True, alpha=ctx.scaling)        # [M, r]
        s_A, s_B = ctx.slots
        dB = native.gemm(dy2, t, a_mn_major=True, b_mn_major=True, out=gradslot.claim(s_B, B.dtype))    # [out, r]
        dA = native.gemm(dt, xd2, a_mn_major=True, b_mn_major=True, out=gradslot.claim(s_A, A.dtype))   # [r, in]
        dx = dxd = None
        if need_x or (ctx.same and need_xd):
            if ctx.same and dt_pad is not None and w.shape[1] > 128:
                dx = native.gemm_dualk(dy2, w, dt_pad, At_pad, b_mn_major=True)   # dy W + dt A in one launch
            else:
                dx = native.gemm(dy2, w, b_mn_major=True)
                if ctx.same:
                    native.gemm(dt, A, b_mn_major=True, out=dx, beta=1.0)
            dx = dx.reshape(ctx.x_shape)
        if not ctx.same and need_xd:
            dxd = native.gemm(dt, A, b_mn_major=True).reshape(ctx.x_shape)
        dw = native.gemm(dy2, x2, a_mn_major=True, b_mn_major=True) if need_w else None
        db = native.colsum(dy2).to(dy.dtype) if (ctx.has_bias and ctx.needs_input_grad[3]) else None
        return dx, dxd, dw, db, dA, dB, None, (dy if ctx.has_res else None)


def lora_linear(x, xd, w, b, A, B, scaling: float, residual=None):
    """LoRA projection. ``xd`` is ``dropout(x)`` (pass ``x`` itself when dropout is off)."""
    per = 16 // x.element_size()
    if _native(x, w) and x.shape[-1] % per == 0 and w.shape[0] % per == 0 and A.shape[0] % per == 0:
        return _LoraLinearFn.apply(x, xd, w, b, A, B, scaling, residual)
    y = F.linear(x, w, b) + scaling * F.linear(F.linear(xd, A), B)
    return y if residual is None else y + residual


def attention_qkvpacked(qkv, causal: bool = False, key_padding_mask=None, dropout_p: float = 0.0, scale: float | None = None):
    """Attention over the packed projection output ``qkv[B, S, 3, H, D]`` -> ``[B, H, S, D]`` (no q/k/v copies; the
    backward fills one packed gradient buffer)."""
    from . import attn as _attn

    return _attn.attention_qkvpacked(qkv, causal=causal, key_padding_mask=key_padding_mask, dropout_p=dropout_p, scale=scale)
