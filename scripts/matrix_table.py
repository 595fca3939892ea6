"""matrix_table.py [dir] [--svg out.svg]: markdown table of the bench matrix
(profiles/bench/r2/matrix_n<N>_<config>_<impl>[_flags].json); --svg also draws throughput relative to the same arm at N=1."""
import json
import re
import sys
from pathlib import Path

argv = sys.argv[1:]
svg = None
if "--svg" in argv:
    i = argv.index("--svg")
    svg = argv[i + 1]
    del argv[i:i + 2]
d = Path(argv[0] if argv else "profiles/bench/r2")
rows = {}
for f in sorted(d.glob("matrix_n*_*.json")):
    m = re.match(r"matrix_n(\d+)_(.+?)_(hyperion|torch|reference)(?:_(.+))?$", f.stem)
    if not m:
        continue
    n, cfg, impl, flags = int(m.group(1)), m.group(2), m.group(3), m.group(4) or ""
    txt = f.read_text().strip()
    if not txt:
        continue
    j = json.loads(txt.splitlines()[0])
    key = (cfg, impl + ("+" + flags if flags else ""))
    if "unavailable" in j:
        rows.setdefault(key, {})[n] = "unavailable"
        continue
    clk = j.get("clocks") or {}
    note = ""
    if clk.get("reasons"):
        note = " [" + ",".join(clk["reasons"]) + "]"
    e2e = (j.get("e2e") or {}).get("value")
    rows.setdefault(key, {})[n] = (j["value"], j["ms_per_step"], e2e, clk.get("sm_mhz"), note, j.get("unit", ""), j.get("peak_mem_gb"))
ns = sorted({n for v in rows.values() for n in v})
print("| config | arm | " + " | ".join(f"N={n}" for n in ns) + " |")
print("|---|---|" + "---|" * len(ns))
for (cfg, arm), v in sorted(rows.items()):
    cells = []
    for n in ns:
        c = v.get(n)
        if c is None:
            cells.append("--")
        elif c == "unavailable":
            cells.append("unavailable")
        else:
            val, ms, e2e, mhz, note, unit, mem = c
            s = f"{val:,.0f} ({ms:.2f} ms"
            if e2e and abs(e2e - val) / val > 1e-6:
                s += f"; e2e {e2e:,.0f}"
            if n > 1 and 1 in v and v[1] != "unavailable":
                s += f"; eff {val / (n * v[1][0]):.0%}"
            s += f"; {mhz:.0f} MHz{note}" if mhz else ""
            s += f"; {mem:.1f} GB" if mem else ""
            cells.append(s + ")")
    print(f"| {cfg} | {arm} | " + " | ".join(cells) + " |")

if svg:
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    from hyperion_b200.utils.svgplot import line_chart

    series = {"ideal": (ns, [float(n) for n in ns])}
    for (cfg, arm), v in sorted(rows.items()):
        pts = {n: c[0] for n, c in v.items() if c != "unavailable"}
        if 1 in pts and len(pts) > 2 and "+" not in arm:
            series[f"{cfg} {arm}"] = (sorted(pts), [pts[n] / pts[1] for n in sorted(pts)])
    line_chart(svg, series, title="Weak scaling: throughput vs the same arm at N=1", xlabel="GPUs", ylabel="x N=1")
