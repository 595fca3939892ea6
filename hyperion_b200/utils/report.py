"""Scaling report: the reference's ``create_scaling_report`` (02_development/distributed_utils.py:563-773).

Same inputs (``*_metrics.csv`` in a directory, classified by file-name substring), same rule (drop the first
``max(1, n//3)`` epochs as warm-up, mean epoch duration, speedup vs the 1-GPU run, efficiency = speedup/gpus),
same outputs (``scaling_analysis.csv`` with columns ``gpus, ideal, {model}_speedup, {model}_efficiency`` and an
optional ``scaling_analysis.png`` when matplotlib exists).  Deviations, all listed in SURVEY.md appendix A:
accepts ``duration_s`` (A-11), never fabricates sample curves when no CSVs exist (A-10), and files of other jobs
(``gpt2_fsdp_*``, ``resnet50_ddp_*``, ``vit_ddp_*``) become columns of their own instead of being skipped.
"""
from __future__ import annotations

import glob
import os
import re
from pathlib import Path

import pandas as pd

MODEL_KEYS = ["language_ddp", "cifar", "language_fsdp", "llama"]


def classify(filename: str) -> str | None:
    base = os.path.basename(filename)
    if "language_ddp" in base:
        return "language_ddp"
    if "cifar" in base:
        return "cifar"
    if "language_fsdp" in base:
        return "language_fsdp"
    if "llama" in base:
        return "llama"
    m = re.match(r"(.+?)_\d+gpus_", base)          # jobs beyond the reference's four: <kind>_<N>gpus_<stamp>_metrics.csv
    return m.group(1) if m else None


def steady_mean_duration(df: pd.DataFrame) -> float | None:
    col = "duration" if "duration" in df.columns else ("duration_s" if "duration_s" in df.columns else None)
    if col is None or "gpus" not in df.columns or len(df) == 0:
        return None
    n = len(df)
    skip = max(1, n // 3) if n > 1 else 0
    d = df[col].iloc[skip:]
    return float(d.mean()) if len(d) else float(df[col].mean())


def collect(data_dir) -> dict:
    """{model: {gpus: mean steady-state epoch seconds}} (latest file wins for a (model, gpus) pair)."""
    out: dict = {}
    for f in sorted(glob.glob(os.path.join(str(data_dir), "*_metrics.csv"))):
        if f.endswith("_device_metrics.csv"):
            continue
        kind = classify(f)
        if kind is None:
            continue
        try:
            df = pd.read_csv(f)
        except Exception:
            continue
        m = steady_mean_duration(df)
        if m is None:
            continue
        out.setdefault(kind, {})[int(df["gpus"].iloc[0])] = m
    return out


def scaling_table(durations: dict) -> pd.DataFrame:
    all_g = sorted({g for d in durations.values() for g in d})
    rows = []
    for g in all_g:
        row = {"gpus": g, "ideal": float(g)}
        for model, d in durations.items():
            base = d.get(1)
            if base is None or g not in d or d[g] <= 0 or base <= 0:  # Llama logs integer seconds: 0 is possible
                row[f"{model}_speedup"] = None
                row[f"{model}_efficiency"] = None
            else:
                sp = base / d[g]
                row[f"{model}_speedup"] = sp
                row[f"{model}_efficiency"] = sp / g
        rows.append(row)
    return pd.DataFrame(rows)


def create_scaling_report(data_dir, plot: bool = True) -> pd.DataFrame | None:
    data_dir = Path(data_dir)
    durations = collect(data_dir)
    if not durations:
        print(f"No *_metrics.csv with 'gpus' and 'duration' columns under {data_dir}; nothing to report.")
        return None
    table = scaling_table(durations)
    data_dir.mkdir(parents=True, exist_ok=True)
    table.to_csv(data_dir / "scaling_analysis.csv", index=False)
    print("\n=== Scaling analysis ===")
    print(table.to_string(index=False))
    if plot:
        from .svgplot import line_chart, pyplot

        plt = pyplot()
        if plt is not None:
            fig, ax = plt.subplots(1, 2, figsize=(12, 5))
            ax[0].plot(table["gpus"], table["ideal"], "k--", label="ideal")
            for model in durations:
                ax[0].plot(table["gpus"], table[f"{model}_speedup"], "o-", label=model)
                ax[1].plot(table["gpus"], table[f"{model}_efficiency"], "o-", label=model)
            ax[0].set_xlabel("GPUs"); ax[0].set_ylabel("speedup"); ax[0].legend()
            ax[1].set_xlabel("GPUs"); ax[1].set_ylabel("efficiency")
            fig.savefig(data_dir / "scaling_analysis.png", dpi=120)
        else:  # no matplotlib in this image: the same two panels as SVG files
            g = list(table["gpus"])
            line_chart(data_dir / "scaling_analysis.svg",
                       {"ideal": (g, list(table["ideal"])), **{m: (g, list(table[f"{m}_speedup"])) for m in durations}},
                       title="Speedup vs 1 GPU", xlabel="GPUs", ylabel="speedup")
            line_chart(data_dir / "scaling_efficiency.svg", {m: (g, list(table[f"{m}_efficiency"])) for m in durations},
                       title="Scaling efficiency", xlabel="GPUs", ylabel="speedup / GPUs")
    return table
