"""launch_breakdown_svg.py <launch summary .txt> <out.svg> [top]: bar chart of where a training step's GPU time goes, from the per-kernel
summary of a launch list (profiles/launches_*_summary.txt: "<ms> <pct> n=<count> avg=<us> <kernel>").  Bars of this library's kernels and
of library (at::, cudnn, nccl ...) kernels are drawn as two series so the share still on ATen is visible at a glance."""
import re
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from hyperion_b200.utils.svgplot import bar_chart  # noqa: E402

src, out = Path(sys.argv[1]), sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
lines = src.read_text().splitlines()
head = lines[0]
rows = []
for l in lines[1:]:
    m = re.match(r"\s*([\d.]+) ms\s+([\d.]+)%\s+n=\s*(\d+)\s+avg=\s*([\d.]+) us\s+(.*)", l)
    if m:
        rows.append((float(m.group(1)), int(m.group(3)), m.group(5).strip()))
own = lambda k: k.startswith(("hyp::", "gemm::", "attn::", "norm::", "conv::", "comm::", "ew::", "bn::", "ce::"))
short = lambda k: re.sub(r"<.*", "", k).split("::")[-1][:16]
rows.sort(reverse=True)
shown, rest = rows[:top], rows[top:]
cats = [f"{short(k)} x{n}" for _, n, k in shown] + ([f"other x{sum(n for _, n, _ in rest)}"] if rest else [])
mine = [ms if own(k) else None for ms, _, k in shown] + ([sum(ms for ms, _, k in rest if own(k)) or None] if rest else [])
lib = [None if own(k) else ms for ms, _, k in shown] + ([sum(ms for ms, _, k in rest if not own(k)) or None] if rest else [])
total_own = sum(ms for ms, _, k in rows if own(k))
total = sum(ms for ms, _, _ in rows)
bar_chart(out, cats, {f"this library ({total_own / total:.1%})": mine, f"ATen / other ({1 - total_own / total:.1%})": lib},
          title=f"{src.stem}: {head}", ylabel="ms per step")
print(f"{out}: {len(rows)} kernels, {total:.2f} ms, {total_own / total:.1%} in this library's kernels")
