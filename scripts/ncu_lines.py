"""ncu_lines.py <report.ncu-rep> [top]: warp-stall samples of the profiled kernel aggregated per CUDA source line."""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h = next(r for r in rows if "# Samples" in r)
isamp, iexec = h.index("# Samples"), h.index("Instructions Executed")
stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not" not in x]
tot, acc = 0, []
for r in rows[rows.index(h) + 1:]:
    if len(r) <= isamp or r[0] == "":
        continue
    try:
        n, e = int(r[isamp]), int(r[iexec])
    except ValueError:
        continue
    tot += n
    st = sorted(((h[i][6:], int(r[i] or 0)) for i in stall if (r[i] or "0").isdigit() and int(r[i]) > 0), key=lambda kv: -kv[1])[:2]
    acc.append((n, e, r[0], r[1], st))
print("total samples", tot)
for n, e, ln, src, st in sorted(acc, key=lambda x: -x[0])[:top]:
    print(f"{ln:>5} {n:6d} {100 * n / max(tot, 1):5.1f}% exec={e:9d}  {src.strip()[:90]:90s} {st}")
