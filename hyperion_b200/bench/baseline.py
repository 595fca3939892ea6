"""Baseline model benchmarks: ``Phase 1/baseline_performance.ipynb`` (c0) as a module + CLI.

    python -m hyperion_b200 baseline [--out results/benchmarks] [--precision fp32|bf16] [--batch 32]

``benchmark_model(model_fn, input_shape, target_shape)`` times three loops like the reference (c0 L70-176): forward only,
forward+backward, full step with Adam(1e-3) and ``MSELoss`` against random targets; derives
``backward = (fwd+bwd) - fwd`` and ``optimizer = full - fwd - bwd``; records peak memory and ``throughput = batch/total``.
``run_baseline_benchmarks`` covers ResNet-50 ``(32,3,224,224)->(32,1000)``, ViT ``(32,3,224,224)->(32,1000)`` and the
custom Transformer ``(32,16,512)->(32,16,512)`` (L179-224); ``test_batch_size_scaling`` sweeps ResNet-50 over 1..64
(L295-340).  CSV columns equal the reference's: ``Model, Forward Time (ms), Backward Time (ms), Optimizer Time (ms),
Total Time (ms), Memory Usage (MB), Throughput (samples/s)`` (+ ``Batch Size``).

The "ViT" row is the real ViT-B/16 here; the reference's published row was its silent CNN fallback, which
is benchmarked too and labelled as what it is.
"""
from __future__ import annotations

import argparse
import os
import sys

import pandas as pd
import torch
import torch.nn as nn

from ..memory import MemoryStats
from ..models import FallbackCNN, create_custom_transformer, create_resnet50, create_vit_model
from ..train.engine import build_engine

COLUMNS = ["Model", "Forward Time (ms)", "Backward Time (ms)", "Optimizer Time (ms)", "Total Time (ms)",
           "Memory Usage (MB)", "Throughput (samples/s)", "Math"]   # the reference's seven columns + the arithmetic used


def _loop_ms(fn, iters: int, device) -> float:
    if device.type == "cuda":
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(device)
        return e0.elapsed_time(e1) / iters
    import time

    t0 = time.time()
    for _ in range(iters):
        fn()
    return (time.time() - t0) * 1e3 / iters


def benchmark_model(model_fn, input_shape, target_shape, num_iterations: int = 50, warmup_iterations: int = 10,
                    precision: str = "fp32", device=None, name: str | None = None) -> dict:
    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    model = model_fn()
    # "fp32" is true IEEE fp32 as in the reference's run (fp32 tensors stay on the library path with TF32 off); "tf32" is
    # fp32 storage on the tensor cores (tcgen05 kind::tf32)
    from ..ops.functional import fp32_math, set_fp32_math

    prev_math = fp32_math()
    if precision in ("fp32", "tf32"):
        set_fp32_math("ieee" if precision == "fp32" else "tf32")
    math_label = {"fp32": "fp32 (IEEE)", "tf32": "tf32 (fp32 storage, tensor cores)"}.get(precision, precision)
    precision = "fp32" if precision == "tf32" else precision
    eng = build_engine(model, strategy="single", precision=precision if device.type == "cuda" else "fp32", lr=1e-3,
                       weight_decay=0.0, device=device)
    dt = eng.precision.dtype
    inputs = torch.rand(*input_shape, device=device).to(dt)
    targets = torch.rand(*target_shape, device=device)
    crit = nn.MSELoss()
    model.train()

    def fwd():
        return eng.model(inputs)

    def fwd_bwd():
        eng.flat.zero_grad()
        crit(eng.model(inputs).float(), targets).backward()

    def full():
        eng.backward_and_step(crit(eng.model(inputs).float(), targets))

    for _ in range(warmup_iterations):
        full()
    f = _loop_ms(fwd, num_iterations, device)
    fb = _loop_ms(fwd_bwd, num_iterations, device)
    t = _loop_ms(full, num_iterations, device)
    mem = MemoryStats(device)
    mem.reset_peak()
    fwd_bwd()
    if device.type == "cuda":
        torch.cuda.synchronize(device)
    bwd = max(fb - f, 0.0)
    # the optimizer is timed on its own (the gradients of the last backward are still in the flat buffer): the difference
    # of two loop timings used before was noise for models whose step is a single 1 ms kernel
    opt = _loop_ms(lambda: eng.optimizer.step(), num_iterations, device) if eng.optimizer is not None else max(t - fb, 0.0)
    total = f + bwd + opt
    set_fp32_math(prev_math)
    return {"Model": name or getattr(model_fn, "__name__", "model"), "Forward Time (ms)": f, "Backward Time (ms)": bwd,
            "Optimizer Time (ms)": opt, "Total Time (ms)": total, "Memory Usage (MB)": mem.sample().peak_mb,
            "Throughput (samples/s)": input_shape[0] / (total / 1e3), "Math": math_label}


def run_baseline_benchmarks(batch: int = 32, precision: str = "fp32", out_dir: str | None = None, iters: int = 50,
                            warmup: int = 10, device=None) -> pd.DataFrame:
    cases = [
        ("ResNet-50", create_resnet50, (batch, 3, 224, 224), (batch, 1000)),
        ("Vision Transformer (ViT-B/16)", create_vit_model, (batch, 3, 224, 224), (batch, 1000)),
        ("Fallback CNN (what the reference published as 'ViT')", FallbackCNN, (batch, 3, 224, 224), (batch, 1000)),
        ("Custom Transformer", create_custom_transformer, (batch, 16, 512), (batch, 16, 512)),
    ]
    rows = [benchmark_model(fn, i, t, iters, warmup, precision, device, name) for name, fn, i, t in cases]
    df = pd.DataFrame(rows, columns=COLUMNS)
    if out_dir:
        d = os.path.join(out_dir, "baseline")
        os.makedirs(d, exist_ok=True)
        df.to_csv(os.path.join(d, "model_benchmarks.csv"), index=False)
    return df


def test_batch_size_scaling(model_fn=create_resnet50, input_hw=(3, 224, 224), out_features: int = 1000,
                            batch_sizes=(1, 2, 4, 8, 16, 32, 64), precision: str = "fp32", out_dir: str | None = None,
                            iters: int = 20, warmup: int = 5, device=None) -> pd.DataFrame:
    rows = []
    for bs in batch_sizes:
        r = benchmark_model(model_fn, (bs, *input_hw), (bs, out_features), iters, warmup, precision, device,
                            getattr(model_fn, "__name__", "model"))
        r["Batch Size"] = bs
        rows.append(r)
    df = pd.DataFrame(rows, columns=COLUMNS + ["Batch Size"])
    if out_dir:
        d = os.path.join(out_dir, "scaling")
        os.makedirs(d, exist_ok=True)
        df.to_csv(os.path.join(d, f"{getattr(model_fn, '__name__', 'model')}_batch_scaling.csv"), index=False)
    return df


def visualize_baseline_results(results: pd.DataFrame, out_dir: str = "results") -> str | None:
    """The notebook's 2x2 page (baseline_performance.ipynb c0 L236-288): time breakdown, memory, throughput, total time per
    model -> ``<out_dir>/visualizations/baseline_benchmarks.svg`` (no matplotlib in this image, so SVG)."""
    if results is None or len(results) == 0:
        print("No results to visualize.")
        return None
    from ..utils.svgplot import bar_chart, compose

    d = os.path.join(out_dir, "visualizations")
    models = [str(m).split(" (")[0] for m in results["Model"]]
    panels = [bar_chart(os.path.join(d, "_p0.svg"), models,
                        {"Forward": list(results["Forward Time (ms)"]), "Backward": list(results["Backward Time (ms)"]),
                         "Optimizer": list(results["Optimizer Time (ms)"])}, title="Time Breakdown by Model", ylabel="Time (ms)")]
    for i, col in enumerate(("Memory Usage (MB)", "Throughput (samples/s)", "Total Time (ms)"), start=1):
        panels.append(bar_chart(os.path.join(d, f"_p{i}.svg"), models, {col: list(results[col])},
                                title=col.split(" (")[0] + " by Model", ylabel=col))
    return str(compose(os.path.join(d, "baseline_benchmarks.svg"), panels))


def visualize_batch_scaling(results: pd.DataFrame, model_name: str, out_dir: str = "results") -> str | None:
    """The notebook's batch-scaling page (c0 L350-400): throughput, memory, total time and scaling efficiency (throughput over
    linear scaling from the first batch size, in %) against batch size."""
    if results is None or len(results) == 0:
        print(f"No scaling results available for {model_name}.")
        return None
    from ..utils.svgplot import compose, line_chart

    d = os.path.join(out_dir, "visualizations")
    bs, thr = list(results["Batch Size"]), list(results["Throughput (samples/s)"])
    eff = [t / (thr[0] * b / bs[0]) * 100 for t, b in zip(thr, bs)]
    panels = []
    for i, (ys, yl) in enumerate(((thr, "Throughput (samples/s)"), (list(results["Memory Usage (MB)"]), "Memory Usage (MB)"),
                                  (list(results["Total Time (ms)"]), "Total Time (ms)"), (eff, "Scaling Efficiency (%)"))):
        panels.append(line_chart(os.path.join(d, f"_p{i}.svg"), {model_name: (bs, ys)},
                                 title=f"{model_name} {yl.split(' (')[0]} vs Batch Size", xlabel="Batch Size", ylabel=yl))
    return str(compose(os.path.join(d, f"{model_name.lower().replace(' ', '_')}_batch_scaling.svg"), panels))


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="hyperion baseline")
    ap.add_argument("--out", default="results/benchmarks")
    ap.add_argument("--precision", default="fp32", choices=["fp32", "tf32", "fp16", "bf16"])
    ap.add_argument("--batch", type=int, default=32)
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--no-scaling", action="store_true")
    a = ap.parse_args(argv)
    df = run_baseline_benchmarks(a.batch, a.precision, a.out, a.iters, a.warmup)
    print(df.to_string(index=False))
    viz = os.path.dirname(os.path.normpath(a.out)) or "."          # results/benchmarks -> results/visualizations
    print("Figure:", visualize_baseline_results(df, viz))
    if not a.no_scaling:
        sc = test_batch_size_scaling(precision=a.precision, out_dir=a.out)
        print(sc.to_string(index=False))
        print("Figure:", visualize_batch_scaling(sc, "ResNet-50", viz))
    return 0


if __name__ == "__main__":
    sys.exit(main())
