"""``hyp_norm_bwd_noalloc``: the norm backward for callers that cannot allocate device memory where they call the
library (a plugin inside another runtime's callback, a fixed-budget real-time loop, a bare C wrapper).  Its partial
sums live in an array compiled into the library instead of caller scratch.

The test drives both entry points straight through the C ABI (ctypes, raw pointers, no Python wrapper in between) and
requires the no-alloc variant to reproduce ``hyp_norm_bwd`` -- twice in a row, the second time with a different,
smaller shape, so anything the first call left in the shared array would show up.  It also checks that a shape whose
partial sums exceed the built-in array is refused with a message instead of running.

This file sorts last on purpose: with ``-x`` a failure here cannot hide any other test.  It passes alone in a fresh
process (the first call does the one-time address look-up) and after the rest of the suite.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

_P, _I, _L = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong
DT = {torch.float32: 0, torch.float16: 1, torch.bfloat16: 2}


@pytest.fixture(scope="module")
def lib():
    from hyperion_b200.ops import _lib

    assert _lib.available(), "libhyperion_b200.so must load on a GPU box"
    L = _lib.require()
    L.hyp_last_error.restype = ctypes.c_char_p
    L.hyp_norm_bwd_scratch_bytes.restype = _L
    L.hyp_norm_bwd_scratch_bytes.argtypes = [_I, _I]
    L.hyp_norm_bwd_noalloc_max_scratch_bytes.restype = _L
    L.hyp_norm_bwd_noalloc_max_scratch_bytes.argtypes = []
    L.hyp_norm_fwd.restype = _I
    L.hyp_norm_fwd.argtypes = [_P] * 8 + [_I, _I, ctypes.c_float, _I, _I, _P]
    L.hyp_norm_bwd.restype = _I
    L.hyp_norm_bwd.argtypes = [_P] * 9 + [_L, _I, _I, _I, _I, _I, _P]
    L.hyp_norm_bwd_noalloc.restype = _I
    L.hyp_norm_bwd_noalloc.argtypes = [_P] * 8 + [_I, _I, _I, _I, _I, _P]
    return L


def _case(lib, rows, d, dtype, rms, seed):
    """Run forward once, then both backwards on the same inputs; return the two result triples."""
    torch.manual_seed(seed)
    dev = "cuda"
    x = torch.randn(rows, d, device=dev).to(dtype)
    g = (1 + 0.1 * torch.randn(d, device=dev)).to(dtype)
    b = (0.1 * torch.randn(d, device=dev)).to(dtype)
    dy = torch.randn(rows, d, device=dev).to(dtype)
    y = torch.empty_like(x)
    mean = torch.empty(rows, device=dev)
    rstd = torch.empty(rows, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.hyp_norm_fwd(x.data_ptr(), None, g.data_ptr(), None if rms else b.data_ptr(), y.data_ptr(), None,
                          mean.data_ptr(), rstd.data_ptr(), rows, d, 1e-5, int(rms), DT[dtype], st)
    assert rc == 0, lib.hyp_last_error()

    def outs():
        return torch.empty_like(x), torch.empty(d, device=dev), torch.empty(d, device=dev)

    # existing entry point: caller-owned scratch
    need = lib.hyp_norm_bwd_scratch_bytes(rows, d)
    scratch = torch.empty(need, device=dev, dtype=torch.uint8)
    dz0, dg0, db0 = outs()
    rc = lib.hyp_norm_bwd(dy.data_ptr(), x.data_ptr(), g.data_ptr(), mean.data_ptr(), rstd.data_ptr(), dz0.data_ptr(),
                          dg0.data_ptr(), None if rms else db0.data_ptr(), scratch.data_ptr(), need, rows, d, int(rms),
                          DT[dtype], 0, st)
    assert rc == 0, lib.hyp_last_error()
    # new entry point: no scratch argument at all
    dz1, dg1, db1 = outs()
    rc = lib.hyp_norm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), g.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                  dz1.data_ptr(), dg1.data_ptr(), None if rms else db1.data_ptr(), rows, d, int(rms),
                                  DT[dtype], 0, st)
    assert rc == 0, lib.hyp_last_error()
    torch.cuda.synchronize()
    # fp32 oracle of the same op (plain PyTorch autograd on upcast inputs)
    xf = x.float().requires_grad_(True)
    gf, bf = g.float().requires_grad_(True), b.float().requires_grad_(True)
    if rms:
        yo = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-5) * gf
    else:
        yo = torch.nn.functional.layer_norm(xf, (d,), gf, bf, 1e-5)
    yo.backward(dy.float())
    oracle = (xf.grad, gf.grad, None if rms else bf.grad)
    return (dz0, dg0, None if rms else db0), (dz1, dg1, None if rms else db1), oracle


def _rel(a, b):
    a, b = a.float(), b.float()
    return ((a - b).norm() / (b.norm() + 1e-12)).item()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_noalloc_matches_existing_entry_point_two_calls_in_a_row(lib, dtype):
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5  # the tolerances of tests/test_gpu_kernels.py
    assert lib.hyp_norm_bwd_noalloc_max_scratch_bytes() == 8 << 20

    def check(ref, got, oracle):
        # dz is computed row by row with no cross-block arithmetic: both entry points run the same code on the same
        # inputs, so it must agree bit for bit.  dgamma / dbeta are fp32 atomic accumulations over row blocks (order
        # varies from run to run), so they agree to rounding of an fp32 sum, far inside the suite's tolerance.
        assert torch.equal(got[0], ref[0])
        for r, o in zip(ref[1:], got[1:]):
            if r is not None:
                assert _rel(o, r) < 1e-5
        for want, o in zip(oracle, got):
            if want is not None:
                assert _rel(o, want) < tol

    # call 1: a big LayerNorm problem fills a large part of the built-in array
    check(*_case(lib, 8192, 1024, dtype, rms=False, seed=11))
    # call 2, straight after: smaller RMSNorm problem; stale partial sums from call 1 must not leak in
    check(*_case(lib, 333, 256, dtype, rms=True, seed=12))
    # and a wide row (one block per row path)
    check(*_case(lib, 64, 4096, dtype, rms=False, seed=13))


def test_noalloc_rejects_shapes_beyond_builtin_scratch(lib):
    rows, d = 4096, 8192
    assert lib.hyp_norm_bwd_scratch_bytes(rows, d) > lib.hyp_norm_bwd_noalloc_max_scratch_bytes()
    x = torch.zeros(8, device="cuda")  # never touched: the call must fail before launching anything
    rc = lib.hyp_norm_bwd_noalloc(x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(), x.data_ptr(),
                                  x.data_ptr(), None, rows, d, 0, 2, 0, None)
    assert rc != 0
    msg = lib.hyp_last_error().decode()
    assert "built-in scratch" in msg and "hyp_norm_bwd" in msg
