"""ncu_roofline.py <raw.csv> [out.md]: one line per profiled kernel from `ncu --page raw --csv` -- duration, DRAM bytes and
achieved DRAM bandwidth vs the measured copy peak, tensor-pipe active %, registers, grid -- the per-kernel roofline table
of BASELINE.md.  Peaks: MEASURED_PEAKS.json (6544 GB/s copy, 1637 / 1394 TFLOPS cuBLAS bf16 burst / sustained)."""
import csv
import re
import sys

PEAK_BW = 6544.0   # GB/s
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}


def col(r, name):
    i = idx.get(name)
    try:
        return float(r[i].replace(",", "")) if i is not None and r[i] else None
    except ValueError:
        return None


out = ["| kernel (ncu --set full, one launch, shape of the BASELINE configs) | time us | DRAM read MB | DRAM write MB | DRAM GB/s (% of 6544) | tensor pipe active % | L2 hit % | regs | grid x block |",
       "|---|---|---|---|---|---|---|---|---|"]
for r in data:
    name = re.sub(r"\(.*", "", r[idx["Kernel Name"]]).replace("void ", "")
    t = col(r, "gpu__time_duration.sum")          # ns
    rd, wr = col(r, "dram__bytes_read.sum"), col(r, "dram__bytes_write.sum")
    unit_t = rows[1][idx["gpu__time_duration.sum"]]
    t_us = t * {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6}.get(unit_t, 1.0)
    ur, uw = rows[1][idx["dram__bytes_read.sum"]], rows[1][idx["dram__bytes_write.sum"]]
    scale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}
    rd_mb = rd * scale.get(ur, 1e-6) if rd is not None else 0.0
    wr_mb = wr * scale.get(uw, 1e-6) if wr is not None else 0.0
    bw = (rd_mb + wr_mb) / 1e3 / (t_us / 1e6) if t_us else 0.0
    tens = col(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed")
    if tens is None:
        tens = col(r, "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed")
    l2 = col(r, "lts__t_sector_hit_rate.pct")
    regs = col(r, "launch__registers_per_thread")
    grid, block = col(r, "launch__grid_size"), col(r, "launch__block_size")
    out.append(f"| `{name[:64]}` | {t_us:.1f} | {rd_mb:.1f} | {wr_mb:.1f} | {bw:.0f} ({100 * bw / PEAK_BW:.0f} %) | "
               f"{tens if tens is not None else 0:.1f} | {l2 if l2 is not None else 0:.0f} | {int(regs or 0)} | {int(grid or 0)} x {int(block or 0)} |")
txt = "\n".join(out) + "\n"
if len(sys.argv) > 2:
    open(sys.argv[2], "w").write(txt)
print(txt)
