"""Sweeps split-K for the wgrad GEMMs (both operands MN-major); run once per HYPERION_GEMM_2CTA setting."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat
from gemm_layer_bench import timeit

T = 8192
dt = torch.bfloat16
print("2CTA =", os.environ.get("HYPERION_GEMM_2CTA", "1"))
for name, n_out, k_in in [("qkv", 3072, 1024), ("proj", 1024, 1024), ("fc1", 4096, 1024), ("fc2", 1024, 4096)]:
    x = torch.randn(T, k_in, device="cuda", dtype=dt)
    dy = torch.randn(T, n_out, device="cuda", dtype=dt)
    fl = 2.0 * T * n_out * k_in / 1e9
    row = []
    for splits in (1, 2, 3, 4, 6, 8):
        if splits == 1:
            out = torch.empty(n_out, k_in, device="cuda", dtype=dt)
            t = timeit(lambda: nat.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=out))
        else:
            acc = torch.zeros(n_out, k_in, device="cuda", dtype=torch.float32)
            t = timeit(lambda: nat.gemm(dy, x, a_mn_major=True, b_mn_major=True, out=acc, splits=splits))
        row.append(f"s{splits}: {t * 1e3:6.1f}us {fl / t:6.0f}TF")
    print(f"{name:5s} " + " | ".join(row), flush=True)
