"""Fusion benchmark: the reference's ``02_development/compilation_optimization.py`` re-targeted.

The reference compares eager vs ``torch.compile`` (default / max-autotune) on two random-init inference models -- the
768-d 4-layer GELU LM on ``idx[128, 8]`` and ``resnet18(num_classes=10)`` on channels-last ``[32,3,32,32]`` -- with 3
warm-up + ``--repeat`` timed ``no_grad`` calls, and writes ``compilation_ckpt_benchmark.{csv,json}``, an analysis text
and a plot (compilation_optimization.py:57-93, 105-156, 232-249).

Here the variants are ``eager`` (stock torch modules, cuBLAS/cuDNN), ``inductor`` (``torch.compile`` default -- a
*baseline*, Triton is never on the product path) and ``hyperion`` (this framework's modules on the hand-written sm_100a
kernels).  Same CSV/JSON keys (``model, variant, time_ms, mem_gb``).  Fixed vs the reference (A-13): the dtype context
is applied inside the timed region and ``mem_gb`` is the per-variant *peak* (reset before each variant).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import pandas as pd
import torch
import torch.nn as nn


def _ref_lm(vocab=50257, d_model=768, nhead=12, dim_ff=3072, n_layers=4):
    class SimpleTransformerLM(nn.Module):  # compilation_optimization.py:57-71
        def __init__(self):
            super().__init__()
            self.embed = nn.Embedding(vocab, d_model)
            layer = nn.TransformerEncoderLayer(d_model, nhead, dim_ff, activation="gelu")
            self.tr = nn.TransformerEncoder(layer, n_layers, enable_nested_tensor=False)
            self.fc = nn.Linear(d_model, vocab)

        def forward(self, idx):
            return self.fc(self.tr(self.embed(idx)))

    return SimpleTransformerLM()


def builders(device):
    import torchvision

    from ..models import CompileBenchLM, resnet18

    lm_in = torch.randint(0, 50257, (128, 8), device=device)
    im_in = torch.randn(32, 3, 32, 32, device=device).contiguous(memory_format=torch.channels_last)
    return {
        "simple_transformer_lm": {"input": lm_in, "eager": _ref_lm, "hyperion": CompileBenchLM},
        "resnet18_cifar10": {"input": im_in, "eager": lambda: torchvision.models.resnet18(num_classes=10),
                             "hyperion": lambda: resnet18(num_classes=10)},
    }


def _graphed(model, x):
    """Capture the no-grad forward in a CUDA graph; returns a callable with the module's call signature."""
    with torch.no_grad():
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                model(x)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            out = model(x)

    def run(_x):
        g.replay()
        return out
    return run


def bench(model, x, repeat: int, device) -> float:
    with torch.no_grad():
        for _ in range(3):
            model(x)
        if device.type == "cuda":
            torch.cuda.synchronize(device)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(repeat):
                model(x)
            e1.record()
            torch.cuda.synchronize(device)
            return e0.elapsed_time(e1) / repeat
        import time

        t0 = time.time()
        for _ in range(repeat):
            model(x)
        return (time.time() - t0) * 1e3 / repeat


def run(base_dir: str = ".", dtype: str = "fp32", repeat: int = 50, variants=("eager", "inductor", "hyperion"), device=None):
    device = torch.device(device or ("cuda" if torch.cuda.is_available() else "cpu"))
    tdt = {"fp32": torch.float32, "bf16": torch.bfloat16, "fp16": torch.float16}[dtype]
    if device.type != "cuda":
        tdt = torch.float32
    rows = []
    for name, spec in builders(device).items():
        x = spec["input"]
        if x.is_floating_point():
            x = x.to(tdt)
        for variant in variants:
            torch.manual_seed(0)
            try:
                if variant == "hyperion":
                    model = spec["hyperion"]()
                else:
                    model = spec["eager"]()
                model = model.to(device).to(tdt).eval()
                if name.startswith("resnet") and device.type == "cuda":
                    model = model.to(memory_format=torch.channels_last)
                if variant == "inductor":
                    model = torch.compile(model, mode="default", fullgraph=False)
                if variant == "hyperion" and device.type == "cuda":
                    model = _graphed(model, x)  # CUDA graph instead of a tracing compiler: one launch per forward
                if device.type == "cuda":
                    torch.cuda.reset_peak_memory_stats(device)
                t = bench(model, x, repeat, device)
                mem = torch.cuda.max_memory_allocated(device) / 1e9 if device.type == "cuda" else 0.0
                rows.append({"model": name, "variant": variant, "time_ms": t, "mem_gb": mem})
                print(f"{name:24s} {variant:10s} {t:8.3f} ms  peak {mem:.3f} GB", flush=True)
            except Exception as e:  # noqa: BLE001  (the reference also continues past a failing variant)
                print(f"{name} {variant} failed: {e!r}", file=sys.stderr)
            finally:
                model = None
                if device.type == "cuda":
                    torch.cuda.empty_cache()
    df = pd.DataFrame(rows)
    out = os.path.join(base_dir, "results/benchmarks/compilation")
    os.makedirs(out, exist_ok=True)
    df.to_csv(os.path.join(out, "compilation_ckpt_benchmark.csv"), index=False)
    with open(os.path.join(out, "compilation_ckpt_benchmark.json"), "w") as f:
        json.dump(rows, f, indent=2)
    lines = []
    for name in df["model"].unique() if len(df) else []:
        sub = df[df.model == name].set_index("variant")
        if "eager" in sub.index:
            for v in sub.index:
                if v != "eager":
                    lines.append(f"{name}: {v} speed-up vs eager = {sub.loc['eager', 'time_ms'] / sub.loc[v, 'time_ms']:.2f}x, "
                                 f"peak memory {sub.loc[v, 'mem_gb']:.3f} GB vs {sub.loc['eager', 'mem_gb']:.3f} GB")
    with open(os.path.join(out, "compilation_ckpt_analysis.txt"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines))
    from ..utils.svgplot import bar_chart, pyplot

    plt = pyplot()
    if plt is not None and len(df):
        fig, ax = plt.subplots(1, 2, figsize=(10, 4))
        df.pivot(index="model", columns="variant", values="time_ms").plot.bar(ax=ax[0], title="time (ms)")
        df.pivot(index="model", columns="variant", values="mem_gb").plot.bar(ax=ax[1], title="peak memory (GB)")
        fig.savefig(os.path.join(out, "compilation_ckpt_speed_mem.png"), dpi=120)
    elif len(df):  # no matplotlib in this image: the same two panels as SVG files
        models = list(df["model"].unique())
        for col, stem, label in (("time_ms", "speed", "time (ms)"), ("mem_gb", "mem", "peak memory (GB)")):
            piv = df.pivot(index="model", columns="variant", values=col)
            bar_chart(os.path.join(out, f"compilation_ckpt_{stem}.svg"), models,
                      {v: [piv.loc[m, v] if v in piv.columns and m in piv.index else None for m in models] for v in piv.columns},
                      title=label, ylabel=label)
    return df


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description="eager vs inductor vs hand-written sm_100a kernels")
    ap.add_argument("--base_dir", default=".")
    ap.add_argument("--dtype", choices=["fp32", "bf16", "fp16"], default="fp32")
    ap.add_argument("--repeat", type=int, default=50)
    ap.add_argument("--variants", default="eager,inductor,hyperion")
    a = ap.parse_args(argv)
    run(a.base_dir, a.dtype, a.repeat, tuple(a.variants.split(",")))
    return 0


if __name__ == "__main__":
    sys.exit(main())
