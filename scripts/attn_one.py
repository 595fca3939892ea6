import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from hyperion_b200.ops.flash import flash_attention
B, H, S, D = 8, 16, 1024, 64
mode = sys.argv[1] if len(sys.argv) > 1 else "fwd"
qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
do = torch.randn(B, H, S, D, device="cuda", dtype=torch.bfloat16)
for _ in range(4):
    o = flash_attention(q, k, v, causal=True)
    if mode == "bwd":
        torch.autograd.grad(o, qkv, do)
torch.cuda.synchronize()
print("ok")
