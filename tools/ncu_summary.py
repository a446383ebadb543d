#!/usr/bin/env python
"""Per-launch summary of an ncu report: python tools/ncu_summary.py report.ncu-rep > profiles/summary.csv
(duration, DRAM bytes read/written, grid, block, registers, occupancy, L2/L1 hit rates, SM and DRAM throughput)."""
import csv
import re
import subprocess
import sys

COLS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__block_size", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9,
         "usecond": 1e3, "msecond": 1e6, "nsecond": 1, "second": 1e9}


def main(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(out.splitlines()))
    head, units, data = rows[0], rows[1], rows[2:]
    ix = {n: i for i, n in enumerate(head)}
    w = csv.writer(sys.stdout)
    w.writerow(["kernel"] + COLS + ["dram_MB", "dram_GBps"])
    for r in data:
        name = re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("void ", "").replace("<unnamed>::", "")
        vals = []
        for c in COLS:
            try:
                v = float(r[ix[c]].replace(",", "")) * SCALE.get(units[ix[c]], 1)
            except (KeyError, ValueError, IndexError):
                v = float("nan")      # metric not in this report
            vals.append(v)
        mb = (vals[1] + vals[2]) / 1e6
        w.writerow([name] + [f"{v:.6g}" for v in vals] + [f"{mb:.1f}", f"{mb * 1e6 / vals[0]:.0f}"])


if __name__ == "__main__":
    main(sys.argv[1])
