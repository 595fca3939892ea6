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
                           device=None, dtype: torch.dtype | None = None, num_iterations: int | None = None) -> dict:
    """Two call forms: the reference's ``(model, inputs, num_iterations=50, warmup=10)`` with a ready input tensor
    (Phase 1/benchmarking.py:12-17), or ``(model, input_shape, batch_size, iterations, warmup)`` with a synthetic one."""
    if torch.is_tensor(input_shape):
        x = input_shape
        if num_iterations is None and isinstance(batch_size, int) and batch_size != 32:
            num_iterations, batch_size = batch_size, x.shape[0]      # positional (model, inputs, num_iterations, warmup)
            if isinstance(iterations, int) and iterations != 100:
                warmup = iterations
        iterations = num_iterations or 50
        batch_size = x.shape[0]
        device = x.device if device is None else torch.device(device)
        model = model.to(device).eval()
        x = x.to(device)
    else:
        iterations = num_iterations or iterations
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


def benchmark_training_step(model, input_shape, target_shape, batch_size=32, iterations=50, warmup: int = 10,
                            use_amp: bool = False, optimizer_class=None, criterion=None, device=None,
                            amp_dtype: str = "bf16", lr: float = 1e-3, num_iterations: int | None = None, **kw) -> dict:
    """Full step (forward, loss, backward, optimizer).  Two call forms:

    * the reference's ``(model, inputs, targets, optimizer, criterion, num_iterations=50, warmup=10, use_amp=False)``
      (Phase 1/benchmarking.py:57-66): ready tensors; the hyper-parameters (lr, betas, eps, weight_decay) are read from the
      passed ``torch.optim`` optimizer and the step runs the fused AdamW with them (Adam = weight_decay 0);
    * ``(model, input_shape, target_shape, batch_size, iterations, warmup, ...)`` with synthetic tensors."""
    from ..train.engine import build_engine

    ref_form = torch.is_tensor(input_shape)
    betas, eps = (0.9, 0.999), 1e-8
    if ref_form:
        x_in, y_in, opt = input_shape, target_shape, batch_size
        if callable(iterations) and not isinstance(iterations, int):
            criterion = iterations
        if isinstance(opt, torch.optim.Optimizer):
            g = opt.param_groups[0]
            lr, betas, eps = g.get("lr", lr), tuple(g.get("betas", betas)), g.get("eps", eps)
            wd = g.get("weight_decay", 0.0)
        else:
            wd = 0.0
        iterations = num_iterations or 50
        batch_size = x_in.shape[0]
        device = x_in.device if device is None else torch.device(device)
    else:
        iterations = num_iterations or iterations
        device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
        wd = 0.0 if (optimizer_class is torch.optim.Adam or optimizer_class is None) else 0.01
    precision = amp_dtype if (use_amp and device.type == "cuda") else "fp32"
    eng = build_engine(model, strategy="single", precision=precision, lr=lr, betas=betas, eps=eps, weight_decay=wd, device=device)
    criterion = criterion or nn.MSELoss()
    dt = eng.precision.dtype
    if ref_form:
        x = x_in.to(device).to(dt) if x_in.is_floating_point() else x_in.to(device)
        y = y_in.to(device)
    else:
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


def compare_precision_formats(model_fn, input_shape, target_shape=None, batch_sizes=(1, 2, 4, 8, 16, 32, 64),
                              precisions=("fp32", "fp16", "bf16"), iterations: int = 50, warmup: int = 10, device=None,
                              num_iterations: int | None = None) -> pd.DataFrame:
    """Batch x precision sweep of the training step; one row per (batch, precision) like the reference's DataFrame.
    The reference's form ``(model_fn, input_shape, batch_sizes=[...], num_iterations=20, warmup=5)``
    (Phase 1/benchmarking.py:153-158) has no target shape: pass a list of batch sizes third (or ``target_shape=None``) and
    the target takes the shape of the model's output."""
    rows = []
    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    if target_shape is not None and len(target_shape) and isinstance(target_shape, list):
        batch_sizes, target_shape = target_shape, None      # positional reference form: third argument = batch sizes
    iterations = num_iterations or iterations
    if target_shape is None:
        with torch.no_grad():
            target_shape = tuple(model_fn()(torch.rand(1, *input_shape)).shape[1:])
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


def save_benchmark_results(results, name: str | None = None, base_dir: str = "results/benchmarks", *, filename: str | None = None,
                           results_dir: str | None = None) -> str:
    """``(results, filename, results_dir)`` as in the reference (Phase 1/benchmarking.py:223-227; ``filename`` with or without
    ``.csv``) or ``(results, name, base_dir)``."""
    name = filename or name
    base_dir = results_dir or base_dir
    os.makedirs(base_dir, exist_ok=True)
    path = os.path.join(base_dir, name if name.endswith(".csv") else f"{name}.csv")
    df = results if isinstance(results, pd.DataFrame) else pd.DataFrame(results if isinstance(results, list) else [results])
    df.to_csv(path, index=False)
    print(f"Results saved to {path}")
    return path
