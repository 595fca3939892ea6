"""Times gemm_tcgen05 against torch.matmul (cuBLAS) with CUDA events; prints TFLOPS per shape."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import native as nat


def timeit(fn, iters=20, warmup=5):
    flush = torch.empty(256 << 20, device="cuda", dtype=torch.uint8)
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    out = []
    for (M, N, K) in [(1024,) * 3, (2048,) * 3, (4096,) * 3, (8192,) * 3, (16384, 4096, 1024), (16384, 1024, 4096),
                      (16384, 3072, 1024), (4064, 2048, 256), (4064, 50264, 256)]:
        for dt in (torch.bfloat16, torch.float16, torch.float32):
            a = torch.randn(M, K, device="cuda", dtype=dt)
            b = torch.randn(N, K, device="cuda", dtype=dt)
            d = torch.empty(M, N, device="cuda", dtype=dt)
            if dt == torch.float32:
                torch.backends.cuda.matmul.allow_tf32 = True
            t_h, t_hb = timeit(lambda: nat.gemm(a, b, out=d))
            t_t, t_tb = timeit(lambda: torch.matmul(a, b.t(), out=d))
            fl = 2.0 * M * N * K / 1e9
            rec = dict(M=M, N=N, K=K, dtype=str(dt).split(".")[-1], hyperion_ms=t_h, hyperion_tflops=fl / t_h,
                       hyperion_best_tflops=fl / t_hb, cublas_ms=t_t, cublas_tflops=fl / t_t, cublas_best_tflops=fl / t_tb)
            print(json.dumps(rec), flush=True)
            out.append(rec)
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/gemm_bench.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
