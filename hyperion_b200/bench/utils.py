"""Benchmark utilities with the reference's four function names, signatures and return keys
(``Phase 1/benchmarking.py``: ``benchmark_forward_pass`` :12-55, ``benchmark_training_step`` :57-151,
``compare_precision_formats`` :153-221, ``save_benchmark_results`` :223-239).

Return dict keys: ``total_time, avg_time, throughput, memory_usage`` (seconds, samples/s, MB).  Differences: timing uses
CUDA events after the warm-up (the reference brackets ``time.time()`` with synchronize), AMP means "run the module in
16-bit with fp32 master weights" through :func:`hyperion_b200.train.build_engine` rather than autocast + GradScaler
objects, and peak memory is reset per measurement.
"""
from __future__ import annotations

import os
import time

import pandas as pd
import torch
import torch.nn as nn

from ..memory import MemoryStats


def _sync(device):
    if torch.device(device).type == "cuda":
        torch.cuda.synchronize(device)


def _timed(fn, iterations: int, device) -> float:
    """Total seconds for ``iterations`` calls of ``fn`` (device time on CUDA)."""
    if torch.device(device).type == "cuda":
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iterations):
            fn()
        e1.record()
        torch.cuda.synchronize(device)
        return e0.elapsed_time(e1) / 1e3
    t0 = time.time()
    for _ in range(iterations):
        fn()
    return time.time() - t0


def benchmark_forward_pass(model, input_shape, batch_size: int = 32, iterations: int = 100, warmup: int = 10,
                           device=None, dtype: torch.dtype | None = None) -> dict:
    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    model = model.to(device).eval()
    dtype = dtype or next(model.parameters()).dtype
    x = torch.rand(batch_size, *input_shape, device=device).to(dtype)
    mem = MemoryStats(device)
    with torch.no_grad():
        for _ in range(warmup):
            model(x)
        _sync(device)
        mem.reset_peak()
        total = _timed(lambda: model(x), iterations, device)
    avg = total / iterations
    return {"total_time": total, "avg_time": avg, "throughput": batch_size / avg, "memory_usage": mem.sample().peak_mb}


def benchmark_training_step(model, input_shape, target_shape, batch_size: int = 32, iterations: int = 50, warmup: int = 10,
                            use_amp: bool = False, optimizer_class=None, criterion=None, device=None,
                            amp_dtype: str = "bf16", lr: float = 1e-3) -> dict:
    """Full step (forward, loss, backward, optimizer).  ``optimizer_class`` is accepted for signature parity; the step
    always runs the fused AdamW (``torch.optim.Adam``/``AdamW`` semantics with weight_decay=0 for Adam)."""
    from ..train.engine import build_engine

    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    precision = amp_dtype if (use_amp and device.type == "cuda") else "fp32"
    wd = 0.0 if (optimizer_class is torch.optim.Adam or optimizer_class is None) else 0.01
    eng = build_engine(model, strategy="single", precision=precision, lr=lr, weight_decay=wd, device=device)
    criterion = criterion or nn.MSELoss()
    dt = eng.precision.dtype
    x = torch.rand(batch_size, *input_shape, device=device).to(dt)
    if isinstance(criterion, nn.CrossEntropyLoss):
        y = torch.randint(0, target_shape[-1] if len(target_shape) else 10, (batch_size,), device=device)
    else:
        y = torch.rand(batch_size, *target_shape, device=device).to(dt)
    model.train()

    def step():
        out = eng.model(x)
        loss = criterion(out.float(), y.float()) if not isinstance(criterion, nn.CrossEntropyLoss) else criterion(out.float(), y)
        eng.backward_and_step(loss)

    for _ in range(warmup):
        step()
    _sync(device)
    mem = MemoryStats(device)
    mem.reset_peak()
    total = _timed(step, iterations, device)
    avg = total / iterations
    return {"total_time": total, "avg_time": avg, "throughput": batch_size / avg, "memory_usage": mem.sample().peak_mb}


def compare_precision_formats(model_fn, input_shape, target_shape, batch_sizes=(1, 2, 4, 8, 16, 32, 64, 128),
                              precisions=("fp32", "fp16", "bf16"), iterations: int = 50, warmup: int = 10, device=None) -> pd.DataFrame:
    """Batch x precision sweep of the training step; one row per (batch, precision) like the reference's DataFrame."""
    rows = []
    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    for bs in batch_sizes:
        for prec in precisions:
            if prec != "fp32" and device.type != "cuda":
                continue
            try:
                r = benchmark_training_step(model_fn(), input_shape, target_shape, bs, iterations, warmup,
                                            use_amp=prec != "fp32", amp_dtype=prec if prec != "fp32" else "bf16", device=device)
                rows.append({"batch_size": bs, "precision": prec, "avg_time_ms": r["avg_time"] * 1e3,
                             "throughput": r["throughput"], "memory_mb": r["memory_usage"]})
            except RuntimeError as e:  # out of memory at large batch: record and continue like the reference's try/except
                rows.append({"batch_size": bs, "precision": prec, "avg_time_ms": None, "throughput": None,
                             "memory_mb": None, "error": str(e)[:80]})
                if device.type == "cuda":
                    torch.cuda.empty_cache()
    return pd.DataFrame(rows)


def save_benchmark_results(results, name: str, base_dir: str = "results/benchmarks") -> str:
    os.makedirs(base_dir, exist_ok=True)
    path = os.path.join(base_dir, f"{name}.csv")
    df = results if isinstance(results, pd.DataFrame) else pd.DataFrame(results if isinstance(results, list) else [results])
    df.to_csv(path, index=False)
    print(f"Results saved to {path}")
    return path
