"""Small-M GEMMs (Llama-7B at 512 tokens) under each tile shape: run once per HYPERION_GEMM_TILE value."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat
from gemm_layer_bench import timeit

dt = torch.bfloat16
print("TILE =", os.environ.get("HYPERION_GEMM_TILE", "0 (auto)"))
for (M, N, K) in [(512, 4096, 4096), (512, 11008, 4096), (512, 4096, 11008), (512, 32000, 4096), (2048, 4096, 4096), (4064, 2048, 256), (4064, 256, 2048)]:
    a = torch.randn(M, K, device="cuda", dtype=dt)
    w = torch.randn(N, K, device="cuda", dtype=dt)
    wt = torch.randn(K, N, device="cuda", dtype=dt)
    out = torch.empty(M, N, device="cuda", dtype=dt)
    fl = 2.0 * M * N * K / 1e9
    t1 = timeit(lambda: nat.gemm(a, w, out=out))
    t2 = timeit(lambda: nat.gemm(a, wt, b_mn_major=True, out=out))
    t3 = timeit(lambda: torch.matmul(a, w.t(), out=out))
    print(f"{M}x{N}x{K}: fwd {t1 * 1e3:6.1f} us {fl / t1:5.0f} TF | dgrad-layout {t2 * 1e3:6.1f} us {fl / t2:5.0f} TF | cublas {t3 * 1e3:6.1f} us {fl / t3:5.0f} TF", flush=True)
