"""GPU numerics tests: every sm_100a kernel vs a plain PyTorch fp32 reference of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def nat():
    from hyperion_b200.ops import _lib, native

    assert _lib.available(), "native library must be loadable on a GPU box"
    assert _lib.require().hyp_device_supported() == 1
    return native


def _rel_err(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


def _assert_gemm_close(d, ref, K, out_dtype):
    """Frobenius error AND element-wise / row-wise bounds: a handful of wrong rows or one wrong tile edge in a large output
    would pass a norm-only check.  |d - ref| <= K-dependent accumulation bound + output rounding, per element."""
    d, ref = d.float(), ref.float()
    assert _rel_err(d, ref) < 1e-2, _rel_err(d, ref)
    eps_out = {torch.bfloat16: 2.0 ** -8, torch.float16: 2.0 ** -11, torch.float32: 2.0 ** -22}[out_dtype]
    tol = eps_out * ref.abs() + 1e-3 * math.sqrt(K) + 2e-2
    bad = (d - ref).abs() > tol
    assert not bad.any(), f"{int(bad.sum())} elements off, worst {float((d - ref).abs().max())} at {bad.nonzero()[0].tolist()}"
    row_err = (d - ref).norm(dim=1) / (ref.norm(dim=1) + 1e-6)
    assert float(row_err.max()) < 2e-2, (int(row_err.argmax()), float(row_err.max()))


GEMM_SHAPES = [
    (128, 256, 64), (256, 512, 256), (4064, 256, 256), (4064, 2048, 256), (512, 768, 3072),
    (1000, 1000, 1000 // 8 * 8), (8192, 8192, 1024), (130, 72, 200), (4096, 50264, 256),
    (257, 264, 72), (1, 256, 64), (383, 1032, 136),   # ragged in M, N and K: partial tiles on every edge
]


@pytest.mark.parametrize("M,N,K", GEMM_SHAPES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_gemm_tn(nat, M, N, K, dtype):
    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda", dtype=dtype)
    b = torch.randn(N, K, device="cuda", dtype=dtype)
    d = nat.gemm(a, b)
    ref = a.float() @ b.float().t()
    _assert_gemm_close(d, ref, K, dtype)


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 512, 256), (4064, 256, 2048), (1024, 3072, 4096), (200, 136, 328)])
def test_gemm_layouts(nat, M, N, K, a_mn, b_mn):
    torch.manual_seed(1)
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    a_arg = a.t().contiguous() if a_mn else a
    b_arg = b.t().contiguous() if b_mn else b
    d = nat.gemm(a_arg, b_arg, a_mn_major=a_mn, b_mn_major=b_mn, out_dtype=torch.float32)
    ref = a.float() @ b.float().t()
    _assert_gemm_close(d, ref, K, torch.float32)


@pytest.mark.parametrize("b_mn", [False, True])
def test_gemm_dual_k(nat, b_mn):
    """One launch, two operand pairs: D = A B^T + A2 B2^T (+ bias + residual)."""
    torch.manual_seed(9)
    M, N, K, K2 = 1000, 640, 320, 64
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    a2 = torch.randn(M, K2, device="cuda", dtype=torch.bfloat16)
    b2 = torch.randn(N, K2, device="cuda", dtype=torch.bfloat16)
    a2[:, 16:] = 0   # as LoRA uses it: rank 16 zero-padded to 64
    bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
    ref = a.float() @ b.float().t() + a2.float() @ b2.float().t()
    bb = b.t().contiguous() if b_mn else b
    assert _rel_err(nat.gemm_dualk(a, bb, a2, b2, b_mn_major=b_mn), ref) < 1e-2
    d = nat.gemm_dualk(a, bb, a2, b2, b_mn_major=b_mn, bias=bias, aux=res)
    assert _rel_err(d, ref + bias.float() + res.float()) < 1e-2


def test_gemm_tf32(nat):
    torch.manual_seed(2)
    a = torch.randn(512, 384, device="cuda")
    b = torch.randn(640, 384, device="cuda")
    d = nat.gemm(a, b)
    ref = a.double() @ b.double().t()
    assert _rel_err(d, ref) < 2e-3
    d2 = nat.gemm(a.t().contiguous(), b.t().contiguous(), a_mn_major=True, b_mn_major=True)
    assert _rel_err(d2, ref) < 2e-3


@pytest.mark.parametrize("act", [None, "relu", "gelu", "gelu_tanh", "silu"])
def test_gemm_epilogue_bias_act_residual(nat, act):
    torch.manual_seed(3)
    M, N, K = 384, 520, 256
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16) * 0.1
    bias = torch.randn(N, device="cuda", dtype=torch.bfloat16)
    res = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
    d = nat.gemm(a, b, bias=bias, aux=res, act=act)
    pre = a.float() @ b.float().t() + bias.float()
    fn = {None: lambda x: x, "relu": torch.relu, "gelu": torch.nn.functional.gelu,
          "gelu_tanh": lambda x: torch.nn.functional.gelu(x, approximate="tanh"),
          "silu": torch.nn.functional.silu}[act]
    ref = fn(pre) + res.float()
    assert _rel_err(d, ref) < 1e-2
    # activation backward epilogue: D = acc * act'(aux)
    if act is not None:
        h = torch.randn(M, N, device="cuda", dtype=torch.bfloat16)
        d2 = nat.gemm(a, b, aux=h, act=act, mode=1, out_dtype=torch.float32)
        hh = h.float().requires_grad_(True)
        fn(hh).sum().backward()
        ref2 = (a.float() @ b.float().t()) * hh.grad
        assert _rel_err(d2, ref2) < 1e-2


def test_gemm_splitk_and_beta(nat):
    torch.manual_seed(4)
    M, N, K = 256, 256, 8192
    a = torch.randn(K, M, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(K, N, device="cuda", dtype=torch.bfloat16)
    ref = a.float().t() @ b.float()
    d = nat.gemm(a, b, a_mn_major=True, b_mn_major=True, out_dtype=torch.float32, splits=8)
    assert _rel_err(d, ref) < 1e-2
    acc = torch.ones(M, N, device="cuda", dtype=torch.float32)
    nat.gemm(a, b, a_mn_major=True, b_mn_major=True, out=acc, beta=1.0)
    assert _rel_err(acc, ref + 1.0) < 1e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
def test_stream_add_copy(nat, dtype):
    n = 10_000_003 if dtype == torch.float32 else 10_000_008
    x = torch.randn(n, device="cuda").to(dtype)
    y = torch.randn(n, device="cuda").to(dtype)
    z = nat.stream_add(x, y)
    assert torch.equal(z, x + y)
    xc = x[: (n // 16) * 16]
    assert torch.equal(nat.stream_copy(xc), xc)


@pytest.mark.parametrize("M,N", [(8192, 4096), (4064, 1024), (333, 72), (5, 8)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_colsum_and_colsum_cast(nat, M, N, dtype):
    """Column sums (fp32 atomics) and the one-launch sum+round variant; the latter twice in a row, because it has to
    leave its accumulator workspace zeroed for the next call."""
    torch.manual_seed(1)
    x = torch.randn(M, N, device="cuda", dtype=dtype)
    ref = x.float().sum(0)
    assert _rel_err(nat.colsum(x), ref) < 1e-3
    for _ in range(2):
        out = nat.colsum_cast(x)
        assert out.dtype == dtype and _rel_err(out, ref) < 1e-2
    xs = x[:, : N // 16 * 8] if N >= 16 else x   # strided rows, still a multiple of 8 columns
    assert _rel_err(nat.colsum_cast(xs), xs.float().sum(0)) < 1e-2


def test_cast_with_device_scale(nat):
    x = torch.randn(1000, 264, device="cuda")
    s = torch.tensor([0.37], device="cuda")
    for dt in (torch.bfloat16, torch.float16):
        assert _rel_err(nat.cast(x, dt, scale=2.0, scale_t=s), x * 0.74) < 1e-2
        assert _rel_err(nat.cast(x[:, :7].contiguous(), dt, scale_t=s), x[:, :7] * 0.37) < 1e-2   # scalar tail path
    assert _rel_err(nat.cast(x.to(torch.bfloat16), torch.float32, scale_t=s), x.to(torch.bfloat16).float() * 0.37) < 1e-6
    xb = x.to(torch.bfloat16)
    assert _rel_err(nat.cast(xb, torch.bfloat16, scale_t=s), xb.float() * 0.37) < 1e-2   # 16-bit -> same type, vector path


@pytest.mark.parametrize("rows,d", [(4064, 256), (512, 512), (1024, 768), (6304, 768), (333, 1024), (256, 4096), (64, 8192)])
@pytest.mark.parametrize("rms", [False, True])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_norm_fwd_bwd(nat, rows, d, rms, dtype):
    torch.manual_seed(5)
    x = torch.randn(rows, d, device="cuda").to(dtype)
    r = torch.randn(rows, d, device="cuda").to(dtype)
    g = (1 + 0.1 * torch.randn(d, device="cuda")).to(dtype)
    b = (0.1 * torch.randn(d, device="cuda")).to(dtype)
    dy = torch.randn(rows, d, device="cuda").to(dtype)
    y, z, mean, rstd = nat.norm_fwd(x, g, None if rms else b, residual=r, eps=1e-5, rms=rms)
    zf = (x.float() + r.float()).requires_grad_(True)
    gf, bf = g.float().requires_grad_(True), b.float().requires_grad_(True)
    if rms:
        ref = zf * torch.rsqrt(zf.pow(2).mean(-1, keepdim=True) + 1e-5) * gf
    else:
        ref = torch.nn.functional.layer_norm(zf, (d,), gf, bf, 1e-5)
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    assert _rel_err(y, ref) < tol
    assert _rel_err(z, zf) < tol
    dz, dgamma, dbeta = nat.norm_bwd(dy, z, g, mean, rstd, rms=rms)
    # reference backward from the 16-bit-rounded z the kernel saw
    zq = z.float().detach().requires_grad_(True)
    if rms:
        ref2 = zq * torch.rsqrt(zq.pow(2).mean(-1, keepdim=True) + 1e-5) * gf
    else:
        ref2 = torch.nn.functional.layer_norm(zq, (d,), gf, bf, 1e-5)
    ref2.backward(dy.float())
    assert _rel_err(dz, zq.grad) < tol
    assert _rel_err(dgamma, gf.grad) < tol
    if not rms:
        assert _rel_err(dbeta, bf.grad) < tol
    # again (the one-kernel backward must have left its accumulator workspace zeroed), with the skip-connection
    # gradient added in-kernel and the parameter gradients written in the parameter dtype
    dres = torch.randn(rows, d, device="cuda").to(dtype)
    g_out = torch.empty(d, device="cuda", dtype=dtype)
    b_out = torch.empty(d, device="cuda", dtype=dtype)
    cs = torch.empty(d, device="cuda", dtype=dtype) if d <= 2048 else None
    dz2, dgamma2, dbeta2 = nat.norm_bwd(dy, z, g, mean, rstd, rms=rms, dres=dres, dgamma_out=g_out, dbeta_out=b_out,
                                        dz_colsum_out=cs)
    assert dgamma2.data_ptr() == g_out.data_ptr()
    if cs is not None:   # column sums of dz ride along (bias gradient of the layer feeding the norm)
        assert _rel_err(cs, (zq.grad + dres.float()).sum(0)) < (3e-2 if dtype == torch.bfloat16 else 1e-4)
    assert _rel_err(dz2, zq.grad + dres.float()) < tol
    assert _rel_err(dgamma2, gf.grad) < tol
    if not rms:
        assert _rel_err(dbeta2, bf.grad) < tol
    # the deterministic two-kernel entry point
    dz3, dgamma3, dbeta3 = nat.norm_bwd_two_pass(dy, z, g, mean, rstd, rms=rms)
    assert _rel_err(dz3, zq.grad) < tol and _rel_err(dgamma3, gf.grad) < tol
    if not rms:
        assert _rel_err(dbeta3, bf.grad) < tol


def test_adamw_matches_torch(nat):
    torch.manual_seed(6)
    n = 1_000_003
    p = torch.randn(n, device="cuda")
    ref_p = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    p16 = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for step in range(1, 4):
        g = torch.randn(n, device="cuda")
        ref_p.grad = g.clone()
        opt.step()
        nat.adamw_step(p, g, m, v, p16, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, step=step)
    assert _rel_err(p, ref_p.data) < 1e-5
    assert torch.equal(p16, p.to(torch.bfloat16))
    # clipping + unscale + found_inf
    g = torch.randn(n, device="cuda") * 64.0
    ss = nat.sumsq(g)
    inv = torch.tensor([1.0 / 64.0], device="cuda")
    p2, m2, v2 = p.clone(), m.clone(), v.clone()
    nat.adamw_step(p2, g, m2, v2, None, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0, step=4,
                   max_norm=1.0, sumsq=ss, inv_scale=inv)
    gu = g / 64.0
    gu = gu * min(1.0, 1.0 / (gu.norm().item() + 1e-6))
    m_ref = 0.9 * m + 0.1 * gu
    assert _rel_err(m2, m_ref) < 1e-4
    found = torch.ones(1, device="cuda")
    p3 = p.clone()
    nat.adamw_step(p3, g, m.clone(), v.clone(), None, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.0,
                   step=4, found_inf=found)
    assert torch.equal(p3, p)
    bad = g.clone(); bad[12345] = float("inf")
    assert nat.sumsq(bad)[1].item() == 1.0 and ss[1].item() == 0.0


@pytest.mark.parametrize("V,dtype", [(50257, torch.bfloat16), (32000, torch.float32), (1000, torch.float16)])
def test_cross_entropy(nat, V, dtype):
    torch.manual_seed(7)
    rows, pad = 515, 7
    ld = (V + 7) // 8 * 8
    buf = torch.zeros(rows, ld, device="cuda", dtype=dtype)
    logits = buf[:, :V]
    logits.copy_(torch.randn(rows, V, device="cuda") * 3)
    tgt = torch.randint(0, V, (rows,), device="cuda")
    tgt[::5] = pad
    lf = logits.float().detach().clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lf, tgt, ignore_index=pad)
    ref.backward()
    sc = torch.zeros(2, device="cuda")
    inv = (tgt != pad).sum().clamp(min=1).float().reciprocal().reshape(1)
    nat.ce_fwd_bwd(buf, tgt, pad, sc, gscale_ptr=inv, n_classes=V)
    loss = torch.empty(1, device="cuda")
    nat.ce_finalize(sc, loss)
    assert abs(loss.item() - ref.item()) < (2e-2 if dtype != torch.float32 else 1e-4)
    assert _rel_err(logits, lf.grad) < (3e-2 if dtype != torch.float32 else 1e-4)


def test_embedding(nat):
    torch.manual_seed(8)
    vocab, d = 50257, 256
    table = torch.randn(vocab, d, device="cuda", dtype=torch.bfloat16)
    idx = torch.randint(0, vocab, (32, 127), device="cuda")
    idx[:, 100:] = 50256
    out = nat.embedding_fwd(table, idx)
    assert torch.equal(out, table[idx])
    dout = torch.randn(32, 127, d, device="cuda", dtype=torch.bfloat16)
    dt = torch.zeros(vocab, d, device="cuda")
    nat.embedding_bwd(dout, idx, dt)
    ref = torch.zeros(vocab, d, device="cuda").index_add_(0, idx.reshape(-1), dout.reshape(-1, d).float())
    assert _rel_err(dt, ref) < 1e-4


def test_rope(nat):
    torch.manual_seed(9)
    B, S, H, D = 2, 128, 32, 128
    x = torch.randn(B, S, H, D, device="cuda", dtype=torch.bfloat16)
    inv_freq = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device="cuda").float() / D))
    ang = torch.arange(S, device="cuda").float()[:, None] * inv_freq[None]
    cos, sin = torch.cat([ang, ang], -1).cos()[None, :, None], torch.cat([ang, ang], -1).sin()[None, :, None]
    xf = x.float()
    rot = torch.cat([-xf[..., D // 2:], xf[..., : D // 2]], -1)
    ref = xf * cos + rot * sin
    y = nat.rope_(x.clone())
    assert _rel_err(y, ref) < 1e-2
    back = nat.rope_(y.clone(), inverse=True)
    assert _rel_err(back, xf) < 2e-2


ATTN_CASES = [
    # B, H, Sq, Sk, D, causal, padded
    (2, 4, 127, 127, 64, False, False),   # reference LM: no mask, ragged S
    (2, 8, 16, 16, 64, False, False),     # custom transformer
    (1, 12, 197, 197, 64, False, False),  # ViT
    (2, 16, 1024, 1024, 64, True, False), # GPT-2 medium
    (2, 4, 128, 128, 128, True, True),    # Llama: causal + key padding, d 128
    (1, 2, 300, 300, 128, True, False),
    (1, 3, 256, 384, 64, True, False),    # Sk > Sq (causal alignment at the end)
]


@pytest.mark.parametrize("B,H,Sq,Sk,D,causal,padded", ATTN_CASES)
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_flash_attention_forward(nat, B, H, Sq, Sk, D, causal, padded, dtype):
    from hyperion_b200.ops.attn import attention_reference
    from hyperion_b200.ops.flash import flash_fwd

    torch.manual_seed(10)
    # packed projection layout [B, S, 3, H, D] like the models produce (strided q/k/v views)
    if Sq == Sk:
        qkv = torch.randn(B, Sq, 3, H, D, device="cuda", dtype=dtype)
        q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    else:
        q = torch.randn(B, H, Sq, D, device="cuda", dtype=dtype)
        k = torch.randn(B, H, Sk, D, device="cuda", dtype=dtype)
        v = torch.randn(B, H, Sk, D, device="cuda", dtype=dtype)
    kpm = None
    if padded:
        lens = torch.randint(Sk // 2, Sk + 1, (B,), device="cuda")
        kpm = torch.arange(Sk, device="cuda")[None] < lens[:, None]
    o, lse, _ = flash_fwd(q, k, v, causal, kpm, 1.0 / D ** 0.5)
    ref = attention_reference(q, k, v, causal, kpm, 0.0, None)
    assert o.shape == ref.shape
    assert _rel_err(o, ref) < 2e-2, _rel_err(o, ref)
    s = torch.matmul(q.float(), k.float().transpose(-1, -2)) / D ** 0.5
    if causal:
        s = s.masked_fill(~torch.ones(Sq, Sk, dtype=torch.bool, device="cuda").tril(Sk - Sq), float("-inf"))
    if kpm is not None:
        s = s.masked_fill(~kpm[:, None, None, :], float("-inf"))
    assert torch.allclose(lse, torch.logsumexp(s, -1), atol=2e-2, rtol=1e-3)


@pytest.mark.parametrize("B,H,Sq,Sk,D,causal,padded", ATTN_CASES)
def test_flash_attention_backward(nat, B, H, Sq, Sk, D, causal, padded):
    from hyperion_b200.ops.attn import attention_reference
    from hyperion_b200.ops.flash import flash_attention

    torch.manual_seed(11)
    dtype = torch.bfloat16
    q = torch.randn(B, H, Sq, D, device="cuda", dtype=dtype, requires_grad=True)
    k = torch.randn(B, H, Sk, D, device="cuda", dtype=dtype, requires_grad=True)
    v = torch.randn(B, H, Sk, D, device="cuda", dtype=dtype, requires_grad=True)
    kpm = None
    if padded:
        lens = torch.randint(Sk // 2, Sk + 1, (B,), device="cuda")
        kpm = torch.arange(Sk, device="cuda")[None] < lens[:, None]
    do = torch.randn(B, H, Sq, D, device="cuda", dtype=dtype)
    o = flash_attention(q, k, v, causal=causal, key_padding_mask=kpm)
    dq, dk, dv = torch.autograd.grad(o, (q, k, v), do)
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    ref = attention_reference(qf, kf, vf, causal, kpm, 0.0, None)
    rq, rk, rv = torch.autograd.grad(ref, (qf, kf, vf), do.float())
    for got, want, name in ((dq, rq, "dq"), (dk, rk, "dk"), (dv, rv, "dv")):
        assert _rel_err(got, want) < 3e-2, (name, _rel_err(got, want))
