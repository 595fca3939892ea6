"""ncu_summary.py <report.ncu-rep> <out.txt>: key raw metrics + the top stall locations of the profiled kernel."""
import csv
import io
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ["Kernel Name", "gpu__time_duration.sum", "sm__cycles_elapsed.max", "launch__registers_per_thread", "launch__block_size",
        "launch__grid_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
lines = [f"# ncu --set full --clock-control none, report {rep.split('/')[-1]}", ""]
for h, u, v in zip(hdr, units, vals):
    if any(h == w or h.endswith("." + w) or h.endswith(w) for w in want):
        lines.append(f"{h} [{u}] = {v}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
srows = list(csv.reader(io.StringIO(src)))
if len(srows) > 2:
    h2, data = srows[1], srows[2:]
    isrc, isamp = h2.index("Source"), h2.index("# Samples")
    stall_cols = [i for i, h in enumerate(h2) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[isamp] or 0) for r in data) or 1
    agg = {}
    for r in data:
        for i in stall_cols:
            agg[h2[i]] = agg.get(h2[i], 0) + int(r[i] or 0)
    lines += ["", f"SASS instructions: {len(data)}; warp samples: {tot}",
              "stall totals: " + ", ".join(f"{k}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]), "",
              "top stall locations (samples, %, SASS, dominant stalls):"]
    for r in sorted(data, key=lambda r: -int(r[isamp] or 0))[:15]:
        st = sorted(((h2[i], int(r[i] or 0)) for i in stall_cols if int(r[i] or 0) > 0), key=lambda kv: -kv[1])[:2]
        lines.append(f"  {int(r[isamp]):6d} {100 * int(r[isamp]) / tot:5.1f}%  {r[isrc][:70]:70s} {st}")
open(out, "w").write("\n".join(lines) + "\n")
print("\n".join(lines[:25]))
