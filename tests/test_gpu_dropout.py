"""Native Philox dropout: the stand-alone kernel, the residual dropout fused into the norm forward, and dropout on the
attention probabilities inside the flash kernels (forward mask regenerated in backward).  Every check rebuilds the mask on
the host from the same (seed, step, salt) with ``ops.rng.reference_keep_mask`` and compares against a plain PyTorch fp32
formulation of the op that uses that mask.  Reference behaviour being covered: nn.TransformerEncoderLayer(dropout=0.1) in
train() mode (02_development/distributed_utils.py:79, 170-181) and lora_dropout=0.05 (:469-473)."""
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    from hyperion_b200 import ops
    from hyperion_b200.ops import _lib, native, rng

    assert _lib.available()
    rng.manual_seed(1234)
    return ops, native, rng


def _mask(rng, dev, salt, plane, rows, cols, p):
    seed, step = rng.state(dev).tolist()
    keep, scale = rng.reference_keep_mask(seed, step, salt, plane, rows, cols, p)
    return torch.from_numpy(keep).to(dev), scale


@pytest.mark.parametrize("rows,cols,dtype", [(257, 1024, torch.bfloat16), (64, 2048, torch.float16), (33, 100, torch.float32),
                                              (5, 8, torch.bfloat16)])
def test_dropout_kernel_matches_host_mask(env, rows, cols, dtype):
    ops, native, rng = env
    dev = torch.device("cuda")
    x = torch.randn(rows, cols, device=dev).to(dtype)
    p, salt = 0.1, 77
    y = native.dropout(x, rng.state(dev), salt, p)
    keep, scale = _mask(rng, dev, salt, 0x80000000, rows, cols, p)
    want = torch.where(keep, x.float() * scale, torch.zeros((), device=dev)).to(dtype)
    assert torch.equal(y, want)
    if rows * cols > 10000:
        assert abs(float(keep.float().mean()) - (1 - 26 / 256)) < 0.01


def test_dropout_autograd_and_step_advance(env):
    ops, native, rng = env
    dev = torch.device("cuda")
    x = torch.randn(64, 256, device=dev, dtype=torch.bfloat16, requires_grad=True)
    y = ops.dropout(x, 0.25, True)
    y.sum().backward()
    kept = y != 0
    assert torch.equal(x.grad != 0, kept)                      # backward used the forward's mask
    assert torch.allclose(x.grad[kept].float(), torch.full((), 256 / 192, device=dev), rtol=4e-3)   # bf16 rounding of 4/3
    y2 = ops.dropout(x.detach(), 0.25, True)                   # new salt -> different mask
    assert not torch.equal(y2 != 0, kept)
    s0 = rng.get_salt()
    rng.set_salt(s0 - 1)                                       # same salt as y2 ...
    rng.advance(dev)                                           # ... but the step word moved on: different mask again
    y3 = ops.dropout(x.detach(), 0.25, True)
    assert not torch.equal(y3 != 0, y2 != 0)
    assert ops.dropout(x, 0.25, False) is x


@pytest.mark.parametrize("d,rms", [(256, False), (1024, False), (4096, True)])
def test_norm_with_fused_residual_dropout(env, d, rms):
    ops, native, rng = env
    dev = torch.device("cuda")
    torch.manual_seed(0)
    rows, p = 300, 0.1
    x = torch.randn(rows, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    res = torch.randn(rows, d, device=dev, dtype=torch.bfloat16, requires_grad=True)
    g = (1 + 0.1 * torch.randn(d, device=dev)).to(torch.bfloat16).requires_grad_(True)
    b = (0.1 * torch.randn(d, device=dev)).to(torch.bfloat16).requires_grad_(True)
    dy = torch.randn(rows, d, device=dev, dtype=torch.bfloat16)
    if rms:
        from hyperion_b200.ops.functional import _NormFn

        y = _NormFn.apply(x, g, None, res, 1e-6, True, p)
    else:
        y = ops.layer_norm(x, g, b, 1e-5, residual=res, dropout_p=p)
    salt = rng.get_salt()
    y.backward(dy)
    keep, scale = _mask(rng, dev, salt, 0x80000000, rows, d, p)
    xf, rf, gf, bf = (t.detach().float().requires_grad_(True) for t in (x, res, g, b))
    z = torch.where(keep, xf * scale, torch.zeros((), device=dev)) + rf
    if rms:
        yo = z * torch.rsqrt(z.pow(2).mean(-1, keepdim=True) + 1e-6) * gf
    else:
        yo = torch.nn.functional.layer_norm(z, (d,), gf, bf, 1e-5)
    yo.backward(dy.float())

    def rel(a, w):
        return float((a.float() - w).norm() / (w.norm() + 1e-9))

    assert rel(y, yo) < 1e-2
    assert rel(x.grad, xf.grad) < 2e-2 and rel(res.grad, rf.grad) < 2e-2 and rel(g.grad, gf.grad) < 2e-2
    assert torch.equal(x.grad == 0, ~keep | (x.grad == 0))   # dropped positions carry no gradient
    if not rms:
        assert rel(b.grad, bf.grad) < 2e-2


def _attn_oracle(q, k, v, keep, scale_keep, causal, kpm):
    D = q.shape[-1]
    s = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(D)
    Sq, Sk = s.shape[-2:]
    if causal:
        s = s.masked_fill(~torch.ones(Sq, Sk, dtype=torch.bool, device=q.device).tril(Sk - Sq), float("-inf"))
    if kpm is not None:
        s = s.masked_fill(~kpm[:, None, None, :].bool(), float("-inf"))
    pr = torch.nan_to_num(torch.softmax(s, -1))
    pr = torch.where(keep, pr * scale_keep, torch.zeros((), device=q.device))
    return torch.matmul(pr, v)


@pytest.mark.parametrize("B,H,S,D,causal,use_kpm", [(2, 4, 127, 64, False, False), (2, 2, 300, 64, True, True),
                                                      (1, 2, 256, 128, True, False), (1, 2, 130, 128, False, True)])
def test_flash_attention_dropout_fwd_bwd(env, B, H, S, D, causal, use_kpm):
    ops, native, rng = env
    from hyperion_b200.ops.flash import flash_attention

    dev = torch.device("cuda")
    torch.manual_seed(1)
    q, k, v = (torch.randn(B, H, S, D, device=dev, dtype=torch.bfloat16, requires_grad=True) for _ in range(3))
    kpm = None
    if use_kpm:
        lens = torch.tensor([S - 7 * (i + 1) for i in range(B)], device=dev)
        kpm = (torch.arange(S, device=dev)[None] < lens[:, None])
    p = 0.1
    o = flash_attention(q, k, v, causal=causal, key_padding_mask=kpm, dropout_p=p)
    salt = rng.get_salt()
    do = torch.randn_like(o)
    o.backward(do)
    seed, step = rng.state(dev).tolist()
    keep = np.stack([rng.reference_keep_mask(seed, step, salt, bh, S, S, p)[0] for bh in range(B * H)]).reshape(B, H, S, S)
    keep = torch.from_numpy(keep).to(dev)
    scale_keep = 256.0 / (256 - 26)
    qf, kf, vf = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    oo = _attn_oracle(qf, kf, vf, keep, scale_keep, causal, kpm)
    oo.backward(do.float())

    def rel(a, w):
        return float((a.float() - w).norm() / (w.norm() + 1e-9))

    assert rel(o, oo) < 2e-2, rel(o, oo)
    assert rel(q.grad, qf.grad) < 3e-2, rel(q.grad, qf.grad)
    assert rel(k.grad, kf.grad) < 3e-2, rel(k.grad, kf.grad)
    assert rel(v.grad, vf.grad) < 3e-2, rel(v.grad, vf.grad)
    # and the dropout really happened: the un-dropped output differs
    o0 = flash_attention(q.detach(), k.detach(), v.detach(), causal=causal, key_padding_mask=kpm)
    assert rel(o0, oo.detach()) > 5e-2


def test_reference_lm_in_train_mode_stays_on_native_kernels(env, monkeypatch):
    """SimpleTransformerLM (dropout 0.1, train mode -- the reference's own job) must not touch F.dropout or the library
    SDPA: attention-probability dropout runs in the flash kernels, residual dropouts in the norm kernel, the FFN dropout in
    the Philox kernel."""
    import torch.nn.functional as F

    from hyperion_b200.models import SimpleTransformerLM
    from hyperion_b200.ops import _lib
    from hyperion_b200.train.engine import build_engine

    def boom(*a, **k):
        raise AssertionError("library dropout / SDPA reached from the native training path")

    dev = torch.device("cuda")
    torch.manual_seed(0)
    model = SimpleTransformerLM(50257)
    eng = build_engine(model, strategy="single", precision="fp16", lr=2e-4, device=dev)
    eng.model.train()
    ids = torch.randint(0, 50256, (8, 128), device=dev)
    monkeypatch.setattr(F, "dropout", boom)
    monkeypatch.setattr(F, "scaled_dot_product_attention", boom)
    n0 = _lib.launch_count()
    losses = []
    for _ in range(3):
        loss = eng.model.loss(ids[:, :-1].contiguous(), ids[:, 1:].contiguous(), ignore_index=50256)
        eng.backward_and_step(loss)
        losses.append(float(loss))
    assert all(math.isfinite(l) for l in losses) and losses[-1] < losses[0]
    assert _lib.launch_count() - n0 > 100
