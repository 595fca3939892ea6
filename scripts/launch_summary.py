"""launch_summary.py <ncu launch csv> [top] [marker]: per-kernel totals of an `ncu --metrics gpu__time_duration.sum --csv` log.

With a marker (a kernel-name substring that occurs once per training step, e.g. "adamw") only the launches after the
second-to-last and up to the last occurrence are counted, i.e. exactly one step."""
import collections
import csv
import re
import sys

path = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
lines = [l for l in open(path) if not l.startswith("==")]
marker = sys.argv[3] if len(sys.argv) > 3 else None
rows = list(csv.DictReader(lines))
if marker:
    hits = [i for i, r in enumerate(rows) if marker in r.get("Kernel Name", "")]
    if len(hits) >= 2:
        rows = rows[hits[-2] + 1: hits[-1] + 1]
agg = collections.defaultdict(lambda: [0, 0.0])
for row in rows:
    name = row.get("Kernel Name", "")
    try:
        val = float(row.get("Metric Value", "0").replace(",", ""))
    except ValueError:
        continue
    unit = row.get("Metric Unit", "ns")
    val_us = val / 1e3 if unit in ("ns", "nsecond") else (val if unit in ("us", "usecond") else val * 1e3)
    name = re.sub(r"\(.*", "", name)
    name = re.sub(r"^void ", "", name)
    a = agg[name[:110]]
    a[0] += 1
    a[1] += val_us
tot = sum(v[1] for v in agg.values())
n = sum(v[0] for v in agg.values())
print(f"{n} launches, {tot / 1e3:.2f} ms total")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{v[1] / 1e3:8.3f} ms {100 * v[1] / tot:5.1f}%  n={v[0]:4d}  avg={v[1] / v[0]:8.1f} us  {k}")
