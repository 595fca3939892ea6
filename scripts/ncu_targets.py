"""One launch of every shipping kernel at the shape the BASELINE configs use, each inside an NVTX range "profile" (after two
warm-up launches outside it), so that

    ncu --set full --clock-control none --import-source on --nvtx --nvtx-include "profile/" -o gpurun_out/prof python scripts/ncu_targets.py

captures exactly one profile per kernel.  Run plain first (exit code 0) as the profiling recipe asks."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200 import ops
from hyperion_b200.ops import functional as HF
from hyperion_b200.ops import native as nat
from hyperion_b200.ops import rng

dev = torch.device("cuda")
bf = torch.bfloat16
torch.manual_seed(0)


def prof(name, fn, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_push("profile")
    fn()
    torch.cuda.synchronize()
    torch.cuda.nvtx.range_pop()
    print("profiled", name, flush=True)


T = 8192
x = torch.randn(T, 1024, device=dev, dtype=bf)
w1 = torch.randn(4096, 1024, device=dev, dtype=bf) * 0.02
b1 = torch.zeros(4096, device=dev, dtype=bf)
dy = torch.randn(T, 4096, device=dev, dtype=bf)
y = torch.empty(T, 4096, device=dev, dtype=bf)
h = torch.empty_like(y)
prof("gemm2 plain fc1 fwd 8192x4096x1024", lambda: nat.gemm(x, w1, out=y))
prof("gemm2 bias+gelu fc1 fwd", lambda: nat.gemm(x, w1, out=y, bias=b1, act="gelu_tanh", preact_out=h))
prof("gemm2 dgrad 8192x1024x4096 (B MN-major)", lambda: nat.gemm(dy, w1, b_mn_major=True))
prof("gemm2 wgrad 4096x1024x8192 (both MN-major)", lambda: HF._wgrad(dy, x, bf))
wq = torch.randn(1024, 1024, device=dev, dtype=bf) * 0.02
dq = torch.randn(T, 1024, device=dev, dtype=bf)
prof("gemm2 split-K wgrad 1024x1024x8192", lambda: HF._wgrad(dq, x, bf))

qkv = torch.randn(8, 1024, 3, 16, 64, device=dev, dtype=bf, requires_grad=True)
from hyperion_b200.ops.flash import flash_attention_qkvpacked

def attn_fb(p=0.0):
    o = flash_attention_qkvpacked(qkv, causal=True, dropout_p=p)
    o.backward(torch.ones_like(o))
    qkv.grad = None
prof("flash fwd+bwd B8 H16 S1024 D64 causal", attn_fb)
prof("flash fwd+bwd with dropout 0.1", lambda: attn_fb(0.1))
q128 = torch.randn(4, 128, 3, 32, 128, device=dev, dtype=bf, requires_grad=True)

def attn128():
    o = flash_attention_qkvpacked(q128, causal=True)
    o.backward(torch.ones_like(o))
    q128.grad = None
prof("flash fwd+bwd B4 H32 S128 D128 causal (Llama)", attn128)

g = torch.ones(1024, device=dev, dtype=bf, requires_grad=True)
bb = torch.zeros(1024, device=dev, dtype=bf, requires_grad=True)
xn = torch.randn(T, 1024, device=dev, dtype=bf, requires_grad=True)
rn = torch.randn(T, 1024, device=dev, dtype=bf)

def norm_fb(p=0.0):
    yy = ops.layer_norm(xn, g, bb, 1e-5, residual=rn, dropout_p=p)
    yy.backward(torch.ones_like(yy))
    xn.grad = g.grad = bb.grad = None
prof("layer_norm fwd+bwd 8192x1024 (+residual)", norm_fb)
prof("layer_norm fwd+bwd with residual dropout", lambda: norm_fb(0.1))

logits = torch.randn(8184, 50304, device=dev, dtype=bf)[:, :50257]
tgt = torch.randint(0, 50257, (8184,), device=dev)
sc = torch.zeros(2, device=dev)
inv = torch.full((1,), 1.0 / 8184, device=dev)
prof("ce_fwd_bwd 8184x50257", lambda: nat.ce_fwd_bwd(logits, tgt, -100, sc, gscale_ptr=inv, n_classes=50257))
prof("colsum 8184x50257 (ragged, vector path)", lambda: nat.colsum(logits))

n = 354_823_168
master = torch.zeros(n, device=dev)
m = torch.zeros(n, device=dev)
v = torch.zeros(n, device=dev)
gr = torch.zeros(n, device=dev, dtype=bf)
p16 = torch.zeros(n, device=dev, dtype=bf)
step_t = torch.ones(1, device=dev)
prof("adamw 354.8M parameters", lambda: nat.adamw_step(master, gr, m, v, p16, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01,
                                                       step=1, step_t=step_t))
xd = torch.randn(T, 4096, device=dev, dtype=bf)
prof("dropout 8192x4096", lambda: nat.dropout(xd, rng.state(dev), 5, 0.1))

import torch.nn as nn
from hyperion_b200.ops import conv as C

conv = nn.Conv2d(256, 256, 3, 1, 1, bias=False).to(dev).to(bf)
conv.weight.data = conv.weight.data.contiguous(memory_format=torch.channels_last)
bn = nn.BatchNorm2d(256).to(dev)
bn.weight.data = bn.weight.data.to(bf); bn.bias.data = bn.bias.data.to(bf)
xi = torch.randn(64, 256, 14, 14, device=dev, dtype=bf).contiguous(memory_format=torch.channels_last).requires_grad_(True)

def conv_fb():
    yy = C.conv_bn_act(xi, conv, bn, relu=True)
    yy.backward(torch.ones_like(yy))
    xi.grad = conv.weight.grad = bn.weight.grad = bn.bias.grad = None
prof("conv3x3 256->256 14x14 b64 + BN + ReLU fwd+bwd (ResNet-50 layer3)", conv_fb)
print("done")
