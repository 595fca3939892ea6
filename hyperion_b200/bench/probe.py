"""Hardware probe: the reference's ``Phase 1/01_hardware_exploration.ipynb`` (c1) as a CLI.

    python -m hyperion_b200 probe [--out results/benchmarks/hardware] [--sizes 1024,2048,4096,8192]

* environment banner, device list, allocated / reserved memory (c1 L16-32)
* basic-op timings: tensor creation, one 5000^2 fp32 matmul, ``sin + cos`` (L35-67)
* matmul TFLOPS sweep n in {1024..8192} x {fp32, fp16, bf16}, ``TFLOPS = 2 n^3 / t / 1e12`` (L72-106) -- reported for
  torch.matmul (cuBLAS; the baseline) AND for the framework's own tcgen05 GEMM
* HBM bandwidth sweep ``z = x + y`` fp32 for {10..500} M elements, ``bytes = 12 n`` (L127-165) -- torch and the
  framework's stream kernel
* CSVs with the reference's schemas: ``precision_results.csv`` (Size,Precision,Time (s),TFLOPS) and
  ``bandwidth_results.csv`` (Size (M elements),Bandwidth (GB/s)); extra columns carry the own-kernel numbers.

Fixes of the reference's method: warm-up, CUDA events, best of N, RNG outside the timed window, L2
flushed between iterations.
"""
from __future__ import annotations

import argparse
import os
import platform
import sys

import pandas as pd
import torch

from ..memory import get_gpu_memory
from ..utils.timing import time_fn


def environment() -> dict:
    info = {"python": platform.python_version(), "torch": torch.__version__, "cuda_available": torch.cuda.is_available(),
            "cuda": torch.version.cuda, "devices": []}
    if torch.cuda.is_available():
        for i in range(torch.cuda.device_count()):
            p = torch.cuda.get_device_properties(i)
            info["devices"].append({"index": i, "name": p.name, "sms": p.multi_processor_count,
                                    "memory_gb": p.total_memory / 2 ** 30, "cc": f"{p.major}.{p.minor}"})
    return info


def test_gpu_operations(device) -> dict:
    """Tensor creation, 5000^2 fp32 matmul, sin+cos (median ms)."""
    out = {}
    out["tensor_creation_ms"], _ = time_fn(lambda: torch.rand(5000, 5000, device=device), 10, 3, False)
    a, b = torch.rand(5000, 5000, device=device), torch.rand(5000, 5000, device=device)
    out["matmul_5000_fp32_ms"], _ = time_fn(lambda: torch.matmul(a, b), 10, 3, False)
    out["sin_cos_ms"], _ = time_fn(lambda: torch.sin(a) + torch.cos(b), 10, 3, False)
    return out


def test_precision_formats(device, sizes=(1024, 2048, 4096, 8192), iters: int = 10) -> pd.DataFrame:
    from ..ops import _lib, native

    rows = []
    # FP32 = true IEEE fp32 like the reference's sweep (library matmul, TF32 off; the tensor-core GEMM has no such mode, so
    # its column stays empty); TF32 = fp32 storage on the tensor cores, library with allow_tf32 vs tcgen05 kind::tf32
    dts = [(torch.float32, "FP32"), (torch.float32, "TF32"), (torch.float16, "FP16"), (torch.bfloat16, "BF16")]
    for n in sizes:
        for dt, name in dts:
            x = torch.rand(n, n, device=device).to(dt)
            y = torch.rand(n, n, device=device).to(dt)
            out = torch.empty(n, n, device=device, dtype=dt)
            torch.backends.cuda.matmul.allow_tf32 = (name == "TF32")
            _, t_torch = time_fn(lambda: torch.matmul(x, y, out=out), iters, 3)
            torch.backends.cuda.matmul.allow_tf32 = False
            row = {"Size": n, "Precision": name, "Time (s)": t_torch / 1e3, "TFLOPS": 2 * n ** 3 / (t_torch / 1e3) / 1e12}
            if torch.device(device).type == "cuda" and _lib.available() and name != "FP32":
                yt = y.t().contiguous()  # the own GEMM takes B as [N, K]
                _, t_own = time_fn(lambda: native.gemm(x, yt, out=out), iters, 3)
                row["Time tcgen05 (s)"] = t_own / 1e3
                row["TFLOPS tcgen05"] = 2 * n ** 3 / (t_own / 1e3) / 1e12
            rows.append(row)
    return pd.DataFrame(rows)


def test_memory_bandwidth(device, sizes_m=(10, 20, 50, 100, 200, 500), iters: int = 10) -> pd.DataFrame:
    from ..ops import _lib, native

    rows = []
    for m in sizes_m:
        n = m * 1_000_000
        x, y = torch.rand(n, device=device), torch.rand(n, device=device)
        z = torch.empty_like(x)
        _, t = time_fn(lambda: torch.add(x, y, out=z), iters, 3)
        row = {"Size (M elements)": m, "Bandwidth (GB/s)": 12 * n / (t / 1e3) / 1e9}
        if torch.device(device).type == "cuda" and _lib.available():
            _, t2 = time_fn(lambda: native.stream_add(x, y, out=z), iters, 3)
            row["Bandwidth stream_add (GB/s)"] = 12 * n / (t2 / 1e3) / 1e9
        rows.append(row)
        del x, y, z
    return pd.DataFrame(rows)


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="hyperion probe")
    ap.add_argument("--out", default="results/benchmarks/hardware")
    ap.add_argument("--sizes", default="1024,2048,4096,8192")
    ap.add_argument("--bw-sizes", default="10,20,50,100,200,500")
    a = ap.parse_args(argv)
    env = environment()
    print(f"Python {env['python']} | torch {env['torch']} | CUDA {env['cuda']} | available: {env['cuda_available']}")
    for d in env["devices"]:
        print(f"  GPU {d['index']}: {d['name']} ({d['sms']} SMs, {d['memory_gb']:.1f} GB, cc {d['cc']})")
    device = "cuda" if torch.cuda.is_available() else "cpu"
    print("Memory allocated / reserved (MB):", get_gpu_memory())
    print("Basic ops:", test_gpu_operations(device))
    sizes = [int(s) for s in a.sizes.split(",")]
    if device == "cpu":
        sizes = [s for s in sizes if s <= 1024]
    prec = test_precision_formats(device, sizes)
    print(prec.to_string(index=False))
    bw_sizes = [int(s) for s in a.bw_sizes.split(",")] if device == "cuda" else [10]
    bw = test_memory_bandwidth(device, bw_sizes)
    print(bw.to_string(index=False))
    os.makedirs(a.out, exist_ok=True)
    prec.to_csv(os.path.join(a.out, "precision_results.csv"), index=False)
    bw.to_csv(os.path.join(a.out, "bandwidth_results.csv"), index=False)
    print(f"CSVs written to {a.out}")
    plot_results(prec, bw, a.out)
    return 0


def plot_results(prec: pd.DataFrame, bw: pd.DataFrame, out: str) -> list[str]:
    """The notebook's two figures (01_hardware_exploration.ipynb c1 L110-123, L165-176): TFLOPS per precision and matrix size,
    bandwidth per vector size.  PNG with matplotlib when it exists, otherwise SVG from ``utils.svgplot``."""
    sizes = sorted(prec["Size"].unique())
    curves = {}
    for name in prec["Precision"].unique():
        sub = prec[prec.Precision == name].set_index("Size")
        curves[name] = (sizes, [sub["TFLOPS"].get(s) for s in sizes])
        if "TFLOPS tcgen05" in sub.columns and sub["TFLOPS tcgen05"].notna().any():
            curves[f"{name} tcgen05"] = (sizes, [sub["TFLOPS tcgen05"].get(s) for s in sizes])
    bws = {c.replace(" (GB/s)", ""): (list(bw["Size (M elements)"]), list(bw[c])) for c in bw.columns if c.endswith("(GB/s)")}
    from ..utils.svgplot import line_chart, pyplot

    plt = pyplot()
    if plt is not None:
        paths = [os.path.join(out, "precision_performance.png"), os.path.join(out, "memory_bandwidth.png")]
        for path, data, xl, yl in ((paths[0], curves, "Matrix size", "TFLOPS"), (paths[1], bws, "Size (M elements)", "GB/s")):
            fig, ax = plt.subplots(figsize=(10, 6))
            for k, (xs, ys) in data.items():
                ax.plot(xs, ys, "o-", label=k)
            ax.set_xlabel(xl); ax.set_ylabel(yl); ax.grid(True); ax.legend()
            fig.savefig(path, dpi=120)
            plt.close(fig)
    else:
        paths = [str(line_chart(os.path.join(out, "precision_performance.svg"), curves,
                                title="Performance across precision formats", xlabel="Matrix size", ylabel="TFLOPS")),
                 str(line_chart(os.path.join(out, "memory_bandwidth.svg"), bws,
                                title="Memory bandwidth vs data size", xlabel="Size (M elements)", ylabel="GB/s"))]
    print("Figures: " + ", ".join(paths))
    return paths


if __name__ == "__main__":
    sys.exit(main())
