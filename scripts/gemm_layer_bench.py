"""GPT-2-medium layer GEMMs (tokens 8192, d 1024, ff 4096) in the three layouts a training step uses -- forward
(NT), dgrad (NN: B MN-major) and wgrad (TN: both MN-major, split-K as ops.functional picks it) -- against cuBLAS."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

from hyperion_b200.ops import functional as HF
from hyperion_b200.ops import native as nat


def timeit(fn, iters=15, warmup=4):
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
    return ts[len(ts) // 2]


def main():
    T = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    dt = torch.bfloat16
    out = []
    for name, n_out, k_in in [("qkv", 3072, 1024), ("proj", 1024, 1024), ("fc1", 4096, 1024), ("fc2", 1024, 4096),
                              ("lm_head", 50264, 1024)]:
        x = torch.randn(T, k_in, device="cuda", dtype=dt)
        w = torch.randn(n_out, k_in, device="cuda", dtype=dt)
        dy = torch.randn(T, n_out, device="cuda", dtype=dt)
        fl = 2.0 * T * n_out * k_in / 1e9
        y = torch.empty(T, n_out, device="cuda", dtype=dt)
        dx = torch.empty(T, k_in, device="cuda", dtype=dt)
        dw = torch.empty(n_out, k_in, device="cuda", dtype=dt)
        rec = dict(name=name, T=T, n_out=n_out, k_in=k_in)
        rec["fwd_h"] = fl / timeit(lambda: nat.gemm(x, w, out=y))
        rec["fwd_c"] = fl / timeit(lambda: torch.matmul(x, w.t(), out=y))
        rec["dgrad_h"] = fl / timeit(lambda: nat.gemm(dy, w, b_mn_major=True, out=dx))
        rec["dgrad_c"] = fl / timeit(lambda: torch.matmul(dy, w, out=dx))
        rec["wgrad_h"] = fl / timeit(lambda: HF._wgrad(dy, x, dt))
        rec["wgrad_c"] = fl / timeit(lambda: torch.matmul(dy.t(), x, out=dw))
        print(json.dumps({k: (round(v, 1) if isinstance(v, float) else v) for k, v in rec.items()}), flush=True)
        out.append(rec)
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/gemm_layer_bench.json").write_text(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
