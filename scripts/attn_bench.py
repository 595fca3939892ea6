"""Times flash_attn_sm100 fwd / fwd+bwd against torch SDPA (cuDNN / flash) for the GPT-2 and Llama shapes."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.nn.functional as F

from hyperion_b200.ops.flash import flash_attention


def t(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


for (B, H, S, D, causal) in [(8, 16, 1024, 64, True), (32, 4, 127, 64, False), (4, 32, 128, 128, True), (2, 32, 4096, 128, True)]:
    qkv = torch.randn(B, S, 3, H, D, device="cuda", dtype=torch.bfloat16, requires_grad=True)
    q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    do = torch.randn(B, H, S, D, device="cuda", dtype=torch.bfloat16)
    fl = 4.0 * B * H * S * S * D * (0.5 if causal else 1.0)
    f_h = t(lambda: flash_attention(q.detach(), k.detach(), v.detach(), causal=causal))
    f_t = t(lambda: F.scaled_dot_product_attention(q.detach(), k.detach(), v.detach(), is_causal=causal))

    def fb_h():
        o = flash_attention(q, k, v, causal=causal)
        torch.autograd.grad(o, qkv, do)

    def fb_t():
        o = F.scaled_dot_product_attention(q, k, v, is_causal=causal)
        torch.autograd.grad(o, qkv, do)
    b_h, b_t = t(fb_h), t(fb_t)
    print(f"B{B} H{H} S{S} D{D} causal={causal}: fwd hyperion {f_h:.0f} us ({fl / f_h / 1e6:.0f} TF) sdpa {f_t:.0f} us ({fl / f_t / 1e6:.0f} TF) | "
          f"fwd+bwd hyperion {b_h:.0f} us ({3.5 * fl / b_h / 1e6:.0f} TF) sdpa {b_t:.0f} us ({3.5 * fl / b_t / 1e6:.0f} TF)", flush=True)
