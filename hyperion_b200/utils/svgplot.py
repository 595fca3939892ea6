"""Dependency-free SVG charts: the fallback for the four plots the reference draws with matplotlib
(``precision_performance.png`` / ``memory_bandwidth.png``: Phase 1/01_hardware_exploration.ipynb c1 L110-123, L165-176;
``scaling_analysis.png``: 02_development/distributed_utils.py:700-760; ``compilation_ckpt_speed_mem.png``:
02_development/compilation_optimization.py:232-249).  matplotlib is not part of this image, so when it cannot be imported the
same data are written as ``.svg`` next to the CSVs (the CSVs stay the artefacts of record).

    line_chart(path, {"ideal": ([1, 2, 4, 8], [1, 2, 4, 8]), ...}, title=..., xlabel=..., ylabel=...)
    bar_chart(path, ["FP16", "BF16"], {"1024": [..], "8192": [..]}, title=..., ylabel=...)
"""
from __future__ import annotations

import math
from pathlib import Path

_COLORS = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd", "#8c564b", "#e377c2", "#7f7f7f"]
_W, _H, _ML, _MR, _MT, _MB = 700, 400, 70, 190, 40, 50


def pyplot():
    """``matplotlib.pyplot`` on the Agg backend, or None when matplotlib is absent (or is a stand-in module without a backend, as
    the reference arm installs for the reference's module-scope import)."""
    try:
        import matplotlib

        matplotlib.use("Agg")
        import matplotlib.pyplot as plt

        return plt if hasattr(plt, "subplots") else None
    except (ImportError, AttributeError):
        return None


def _esc(s) -> str:
    return str(s).replace("&", "&amp;").replace("<", "&lt;").replace(">", "&gt;")


def _ticks(lo: float, hi: float, n: int = 5) -> list[float]:
    if hi <= lo:
        hi = lo + 1.0
    raw = (hi - lo) / n
    mag = 10.0 ** math.floor(math.log10(raw))
    step = next(m * mag for m in (1, 2, 5, 10) if m * mag >= raw * (1 - 1e-9))
    t, out = math.floor(lo / step) * step, []
    while t <= hi + 1e-9:
        if t >= lo - 1e-9:
            out.append(round(t, 10))
        t += step
    return out or [lo, hi]


def _frame(title, xlabel, ylabel, extra_bottom: int = 0) -> list[str]:
    return [f'<svg xmlns="http://www.w3.org/2000/svg" width="{_W}" height="{_H + extra_bottom}" font-family="sans-serif" font-size="12">',
            f'<rect width="{_W}" height="{_H + extra_bottom}" fill="white"/>',
            f'<text x="{(_ML + _W - _MR) / 2}" y="20" text-anchor="middle" font-size="14">{_esc(title)}</text>',
            f'<text x="{(_ML + _W - _MR) / 2}" y="{_H + extra_bottom - 10}" text-anchor="middle">{_esc(xlabel)}</text>',
            f'<text x="15" y="{(_MT + _H - _MB) / 2}" text-anchor="middle" transform="rotate(-90 15 {(_MT + _H - _MB) / 2})">{_esc(ylabel)}</text>',
            f'<rect x="{_ML}" y="{_MT}" width="{_W - _ML - _MR}" height="{_H - _MT - _MB}" fill="none" stroke="black"/>']


def _yaxis(out: list[str], lo: float, hi: float):
    def ypix(v):
        return _H - _MB - (v - lo) / (hi - lo) * (_H - _MT - _MB)
    for t in _ticks(lo, hi):
        y = ypix(t)
        out.append(f'<line x1="{_ML}" y1="{y:.1f}" x2="{_W - _MR}" y2="{y:.1f}" stroke="#dddddd"/>')
        out.append(f'<text x="{_ML - 6}" y="{y + 4:.1f}" text-anchor="end">{t:g}</text>')
    return ypix


def _legend(out: list[str], names) -> None:
    for i, name in enumerate(names):
        y = _MT + 14 + 18 * i
        out.append(f'<rect x="{_W - _MR + 10}" y="{y - 9}" width="12" height="12" fill="{_COLORS[i % len(_COLORS)]}"/>')
        out.append(f'<text x="{_W - _MR + 28}" y="{y + 1}">{_esc(name)}</text>')


def line_chart(path, series: dict, title: str = "", xlabel: str = "", ylabel: str = "") -> Path:
    """``series``: name -> (xs, ys).  Points with ``None`` / NaN values are skipped."""
    pts = {k: [(float(x), float(y)) for x, y in zip(*v) if y is not None and y == y] for k, v in series.items()}
    pts = {k: v for k, v in pts.items() if v}
    xs = [x for v in pts.values() for x, _ in v] or [0.0, 1.0]
    ys = [y for v in pts.values() for _, y in v] or [0.0, 1.0]
    x0, x1, y0, y1 = min(xs), max(xs), min(0.0, min(ys)), max(ys) * 1.05 or 1.0
    if x1 == x0:
        x1 = x0 + 1.0
    out = _frame(title, xlabel, ylabel)
    ypix = _yaxis(out, y0, y1)

    def xpix(v):
        return _ML + (v - x0) / (x1 - x0) * (_W - _ML - _MR)
    for x in sorted(set(xs)):
        out.append(f'<text x="{xpix(x):.1f}" y="{_H - _MB + 16}" text-anchor="middle">{x:g}</text>')
    for i, (name, v) in enumerate(pts.items()):
        c = _COLORS[i % len(_COLORS)]
        dash = ' stroke-dasharray="6 4"' if name.lower() == "ideal" else ""
        out.append(f'<polyline fill="none" stroke="{c}" stroke-width="2"{dash} points="' +
                   " ".join(f"{xpix(x):.1f},{ypix(y):.1f}" for x, y in v) + '"/>')
        out += [f'<circle cx="{xpix(x):.1f}" cy="{ypix(y):.1f}" r="3" fill="{c}"/>' for x, y in v]
    _legend(out, list(pts))
    out.append("</svg>")
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_text("\n".join(out))
    return path


def bar_chart(path, categories: list, groups: dict, title: str = "", xlabel: str = "", ylabel: str = "") -> Path:
    """Grouped bars: ``groups``: series name -> one value per category."""
    vals = [float(v) for g in groups.values() for v in g if v is not None and v == v] or [1.0]
    y1 = max(vals) * 1.05 or 1.0
    tilt = len(categories) > 6                      # many categories: slanted labels below the axis instead of overlapping ones
    out = _frame(title, xlabel, ylabel, 70 if tilt else 0)
    ypix = _yaxis(out, 0.0, y1)
    n, k = max(1, len(categories)), max(1, len(groups))
    slot = (_W - _ML - _MR) / n
    bw = slot * 0.8 / k
    for ci, cat in enumerate(categories):
        x, y = _ML + slot * (ci + 0.5), _H - _MB + 16
        if tilt:
            out.append(f'<text x="{x:.1f}" y="{y - 4}" text-anchor="end" font-size="10" transform="rotate(-40 {x:.1f} {y - 4})">{_esc(cat)}</text>')
        else:
            out.append(f'<text x="{x:.1f}" y="{y}" text-anchor="middle">{_esc(cat)}</text>')
        for gi, (name, g) in enumerate(groups.items()):
            v = g[ci] if ci < len(g) else None
            if v is None or v != v:
                continue
            x = _ML + slot * ci + slot * 0.1 + bw * gi
            y = ypix(float(v))
            out.append(f'<rect x="{x:.1f}" y="{y:.1f}" width="{bw:.1f}" height="{_H - _MB - y:.1f}" fill="{_COLORS[gi % len(_COLORS)]}"/>')
    _legend(out, list(groups))
    out.append("</svg>")
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_text("\n".join(out))
    return path


def compose(path, panels: list, cols: int = 2, remove: bool = True) -> Path:
    """One figure out of several chart files (the reference's 2x2 ``plt.subplots`` pages): every panel becomes a nested
    ``<svg>`` at its grid position.  ``remove`` deletes the panel files afterwards."""
    rows = (len(panels) + cols - 1) // cols
    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{_W * cols}" height="{_H * rows}">']
    for i, p in enumerate(panels):
        body = Path(p).read_text().replace("<svg ", f'<svg x="{_W * (i % cols)}" y="{_H * (i // cols)}" ', 1)
        out.append(body)
        if remove:
            Path(p).unlink()
    out.append("</svg>")
    path = Path(path)
    path.parent.mkdir(parents=True, exist_ok=True)
    path.write_text("\n".join(out))
    return path
