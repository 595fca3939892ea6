"""Operand-majorness sweep: the same M x N x K product with A and/or B stored MN-major."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat
from gemm_layer_bench import timeit

dt = torch.bfloat16
print("2CTA =", os.environ.get("HYPERION_GEMM_2CTA", "1"))
for (M, N, K) in [(4096, 4096, 8192), (4096, 1024, 8192), (8192, 8192, 8192)]:
    fl = 2.0 * M * N * K / 1e9
    out = torch.empty(M, N, device="cuda", dtype=dt)
    row = []
    for amn in (False, True):
        for bmn in (False, True):
            a = torch.randn((K, M) if amn else (M, K), device="cuda", dtype=dt)
            b = torch.randn((K, N) if bmn else (N, K), device="cuda", dtype=dt)
            t = timeit(lambda: nat.gemm(a, b, a_mn_major=amn, b_mn_major=bmn, out=out))
            row.append(f"A{'mn' if amn else 'k '} B{'mn' if bmn else 'k '}: {t * 1e3:6.1f}us {fl / t:6.0f}TF")
    print(f"{M}x{N}x{K} " + " | ".join(row), flush=True)
