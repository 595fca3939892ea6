"""Times every GEMM variant of one GPT-2-medium layer (fwd + bwd) at M = 8192 tokens, vs cuBLAS for the bare matmul."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat

dt = torch.bfloat16
M, E, F = 8192, 1024, 4096
dev = "cuda"
r = lambda *s: torch.randn(*s, device=dev, dtype=dt)
x, h4, dy, dy4 = r(M, E), r(M, F), r(M, E), r(M, F)
wqkv, wo, w1, w2 = r(3 * E, E), r(E, E), r(F, E), r(E, F)
bqkv, bo, b1, b2 = r(3 * E), r(E), r(F), r(E)
res = r(M, E)
flush = torch.empty(256 << 20, device=dev, dtype=torch.uint8)


def t(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


cases = [
    ("qkv  fwd  bias            8192x3072x1024", lambda: nat.gemm(x, wqkv, bias=bqkv), lambda: x @ wqkv.t(), 2 * M * 3 * E * E),
    ("proj fwd  bias+resid      8192x1024x1024", lambda: nat.gemm(x, wo, bias=bo, aux=res), lambda: x @ wo.t(), 2 * M * E * E),
    ("fc   fwd  bias+gelu+pre   8192x4096x1024", lambda: nat.gemm(x, w1, bias=b1, act="gelu_tanh", preact_out=h4.clone()), lambda: x @ w1.t(), 2 * M * F * E),
    ("fc2  fwd  bias+resid      8192x1024x4096", lambda: nat.gemm(h4, w2, bias=b2, aux=res), lambda: h4 @ w2.t(), 2 * M * F * E),
    ("fc2  dgrad*dact           8192x4096x1024", lambda: nat.gemm(dy, w2, b_mn_major=True, aux=h4, act="gelu_tanh", mode=1), lambda: dy @ w2, 2 * M * F * E),
    ("fc2  wgrad                1024x4096x8192", lambda: nat.gemm(dy, h4, a_mn_major=True, b_mn_major=True), lambda: dy.t() @ h4, 2 * M * F * E),
    ("fc   wgrad                4096x1024x8192", lambda: nat.gemm(dy4, x, a_mn_major=True, b_mn_major=True), lambda: dy4.t() @ x, 2 * M * F * E),
    ("fc   dgrad                8192x1024x4096", lambda: nat.gemm(dy4, w1, b_mn_major=True), lambda: dy4 @ w1, 2 * M * F * E),
    ("proj wgrad                1024x1024x8192", lambda: nat.gemm(dy, x, a_mn_major=True, b_mn_major=True), lambda: dy.t() @ x, 2 * M * E * E),
    ("qkv  wgrad                3072x1024x8192", lambda: nat.gemm(r(M, 3 * E), x, a_mn_major=True, b_mn_major=True), lambda: dy.t() @ x, 2 * M * 3 * E * E),
]
tot_h = tot_c = 0.0
for name, fh, fc, fl in cases:
    th, tc = t(fh), t(fc)
    tot_h += th; tot_c += tc
    print(f"{name}: hyperion {th:7.1f} us {fl / th / 1e6:6.0f} TF | cublas(bare) {tc:7.1f} us {fl / tc / 1e6:6.0f} TF", flush=True)
print(f"sum hyperion {tot_h:.0f} us, cublas bare {tot_c:.0f} us")
