#!/usr/bin/env python
"""One line per profiled launch of an ncu report: duration, DRAM traffic, instructions, issue
utilisation, occupancy, top stall reasons.   python tools/ncu_summary.py <report.ncu-rep> [--csv out]"""
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h = rows[0]
    col = {name: i for i, name in enumerate(h)}

    def get(r, name, default=""):
        return r[col[name]] if name in col else default

    units = rows[1]
    scale = {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
    tscale = scale.get(units[col["gpu__time_duration.sum"]], 1.0) if "gpu__time_duration.sum" in col else 1.0
    bscale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}

    def mb(r, name):
        return float(get(r, name) or 0) * bscale.get(units[col[name]], 1.0) if name in col else 0.0

    stall_cols = [c for c in h if c.startswith("smsp__average_warps_issue_stalled_") and c.endswith("_per_issue_active.ratio")]
    lines = []
    for r in rows[2:]:
        name = get(r, "Kernel Name").split("(")[0].replace("<unnamed>::", "").replace("void ", "")
        stalls = sorted(((float(r[col[c]] or 0), c.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")) for c in stall_cols), reverse=True)
        stalls = [(v, n) for v, n in stalls if n not in ("selected",)][:4]
        rec = dict(
            kernel=name, grid=get(r, "Grid Size"), block=get(r, "Block Size"),
            ms=float(get(r, "gpu__time_duration.sum") or 0) * tscale,
            dram_rd_mb=mb(r, "dram__bytes_read.sum"), dram_wr_mb=mb(r, "dram__bytes_write.sum"),
            dram_pct=float(get(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed") or 0),
            minst=float(get(r, "smsp__inst_executed.sum") or 0) / 1e6,
            issue_pct=float(get(r, "smsp__issue_active.avg.pct_of_peak_sustained_active") or 0),
            warps_pct=float(get(r, "sm__warps_active.avg.pct_of_peak_sustained_active") or 0),
            regs=get(r, "launch__registers_per_thread"),
            stalls=" ".join(f"{n}={v:.2f}" for v, n in stalls))
        lines.append(rec)
    for rec in lines:
        print("{kernel:38s} {grid:>14s} {block:>12s} {ms:8.4f} ms  dram r/w {dram_rd_mb:8.1f}/{dram_wr_mb:8.1f} MB ({dram_pct:4.1f}%)  "
              "{minst:8.1f} Minst  issue {issue_pct:4.1f}%  warps {warps_pct:4.1f}%  regs {regs}  | {stalls}".format(**rec))
    if "--csv" in sys.argv:
        path = sys.argv[sys.argv.index("--csv") + 1]
        with open(path, "w", newline="") as f:
            w = csv.DictWriter(f, fieldnames=list(lines[0].keys()))
            w.writeheader()
            w.writerows(lines)


if __name__ == "__main__":
    main()
