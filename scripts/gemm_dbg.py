import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from hyperion_b200.ops import native as nat
from gemm_layer_bench import timeit
M = N = K = 8192
dt = torch.bfloat16
a = torch.randn(K, M, device="cuda", dtype=dt); b = torch.randn(K, N, device="cuda", dtype=dt)
out = torch.empty(M, N, device="cuda", dtype=dt)
fl = 2.0 * M * N * K / 1e9
for amn, bmn in ((False, False), (True, False), (False, True), (True, True)):
    t = timeit(lambda: nat.gemm(a, b, a_mn_major=amn, b_mn_major=bmn, out=out))
    print("DBG", os.environ.get("HYPERION_GEMM_DBG", "0"), "amn", amn, "bmn", bmn, f"{t*1e3:.1f}us {fl/t:.0f}TF", flush=True)
