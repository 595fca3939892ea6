"""Runs one gemm_tcgen05 shape a few times (target of `ncu --set full -k regex:gemm_kernel`)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat

M, N, K = (int(x) for x in sys.argv[1:4])
layout = sys.argv[4] if len(sys.argv) > 4 else "tn"   # tn: fwd, nn: dgrad (B mn-major), wg: wgrad (both mn-major)
iters = int(sys.argv[5]) if len(sys.argv) > 5 else 5
epi = sys.argv[6] if len(sys.argv) > 6 else "plain"   # plain | dact (gelu' * aux) | act (bias + gelu + preact)
dt = torch.bfloat16
a = torch.randn(M, K, device="cuda", dtype=dt)
b = torch.randn(N, K, device="cuda", dtype=dt)
if layout == "nn":
    b = b.t().contiguous()
    kw = dict(b_mn_major=True)
elif layout == "wg":
    a, b = a.t().contiguous(), b.t().contiguous()
    kw = dict(a_mn_major=True, b_mn_major=True)
else:
    kw = {}
d = torch.empty(M, N, device="cuda", dtype=dt)
if epi == "dact":
    kw.update(aux=torch.randn(M, N, device="cuda", dtype=dt), act="gelu_tanh", mode=1)
elif epi == "act":
    kw.update(bias=torch.randn(N, device="cuda", dtype=dt), act="gelu_tanh", preact_out=torch.empty(M, N, device="cuda", dtype=dt))
for _ in range(iters):
    nat.gemm(a, b, out=d, **kw)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    nat.gemm(a, b, out=d, **kw)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"{layout} {M}x{N}x{K}: {ms*1e3:.1f} us  {2*M*N*K/ms/1e9:.0f} TFLOPS")
