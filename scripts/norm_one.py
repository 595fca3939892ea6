"""One LayerNorm forward + backward on the GPT-2-medium activation shape (for ncu captures and quick timing)."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200 import ops

rows, d = 8192, 1024
x = torch.randn(8, 1024, d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
g = torch.randn(d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
b = torch.randn(d, device="cuda", dtype=torch.bfloat16, requires_grad=True)
dy = torch.randn_like(x)
ds = torch.randn_like(x)
flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)


def step():
    y, skip = ops.layer_norm_fork(x, g, b, 1e-5)
    torch.autograd.grad([y, skip], [x, g, b], [dy, ds])


for _ in range(3):
    step()
torch.cuda.synchronize()
if len(sys.argv) > 1 and sys.argv[1] == "time":
    ts = []
    for _ in range(20):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); step(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    print(f"fwd+bwd median {ts[len(ts) // 2]:.1f} us (includes launch gaps)")
print("ok")
