"""Which part of the fused epilogues costs time: fc1-shaped GEMM (8192 x 4096 x 1024) with epilogue variants."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat
from gemm_layer_bench import timeit

dt = torch.bfloat16
M, E, F = 8192, 1024, 4096
r = lambda *s: torch.randn(*s, device="cuda", dtype=dt)
x, w1, b1, dy, w2, h4 = r(M, E), r(F, E), r(F), r(M, E), r(E, F), r(M, F)
out, pre = torch.empty(M, F, device="cuda", dtype=dt), torch.empty(M, F, device="cuda", dtype=dt)
fl = 2.0 * M * F * E / 1e9
cases = [
    ("fwd plain", lambda: nat.gemm(x, w1, out=out)),
    ("fwd bias", lambda: nat.gemm(x, w1, out=out, bias=b1)),
    ("fwd bias+relu", lambda: nat.gemm(x, w1, out=out, bias=b1, act="relu")),
    ("fwd bias+gelu_tanh", lambda: nat.gemm(x, w1, out=out, bias=b1, act="gelu_tanh")),
    ("fwd bias+gelu_tanh+preact", lambda: nat.gemm(x, w1, out=out, bias=b1, act="gelu_tanh", preact_out=pre)),
    ("fwd bias+silu+preact", lambda: nat.gemm(x, w1, out=out, bias=b1, act="silu", preact_out=pre)),
    ("dgrad plain", lambda: nat.gemm(dy, w2, b_mn_major=True, out=out)),
    ("dgrad * relu'(aux)", lambda: nat.gemm(dy, w2, b_mn_major=True, out=out, aux=h4, act="relu", mode=1)),
    ("dgrad * gelu_tanh'(aux)", lambda: nat.gemm(dy, w2, b_mn_major=True, out=out, aux=h4, act="gelu_tanh", mode=1)),
    ("dgrad + aux (residual)", lambda: nat.gemm(dy, w2, b_mn_major=True, out=out, aux=h4)),
]
for name, fn in cases:
    t = timeit(fn)
    print(f"{name:28s} {t * 1e3:7.1f} us {fl / t:6.0f} TF", flush=True)
