#!/usr/bin/env python
"""Summarise ncu output for profiles/ (run here, no GPU needed).

  python tools/ncu_summary.py launches gpurun_out/x_launches.csv > profiles/rN_launches_summary.csv
  python tools/ncu_summary.py full gpurun_out/x_prof.ncu-rep > profiles/rN_full_summary.csv

`launches`: per-kernel aggregation (count, total / mean device time, share) of an
`ncu --metrics gpu__time_duration.sum --csv` launch list.
`full`: one row per captured launch of a `--set full` report with the metrics DESIGN.md cites.
"""
import csv
import io
import re
import subprocess
import sys
from collections import OrderedDict

FULL_METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_tensor.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "l1tex__m_xbar2l1tex_read_bytes.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "sm__inst_executed_pipe_xu.sum",
]


def short(name):
    name = re.sub(r"\(.*", "", name)
    return name.replace("ppi::", "").strip()


def read_csv_rows(text):
    lines = [l for l in text.splitlines() if l.startswith('"')]
    return list(csv.DictReader(io.StringIO("\n".join(lines))))


def launches(path):
    rows = read_csv_rows(open(path, errors="replace").read())
    agg = OrderedDict()
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
        a = agg.setdefault(short(r["Kernel Name"]), [0, 0.0])
        a[0] += 1
        a[1] += v
    total = sum(a[1] for a in agg.values())
    out = csv.writer(sys.stdout)
    out.writerow(["kernel", "launches", "total_ms", "mean_us", "share"])
    for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.writerow([k, n, f"{ms:.3f}", f"{ms / n * 1e3:.1f}", f"{ms / total:.4f}"])
    out.writerow(["TOTAL", sum(a[0] for a in agg.values()), f"{total:.3f}", "", "1.0"])


def full(path):
    text = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(text)))
    header = rows[0]
    units = rows[1]
    idx = {h: i for i, h in enumerate(header)}
    cols = [m for m in FULL_METRICS if m in idx]
    out = csv.writer(sys.stdout)
    out.writerow(["id", "kernel"] + [f"{m} [{units[idx[m]]}]" for m in cols])
    for r in rows[2:]:
        if len(r) < len(header):
            continue
        out.writerow([r[idx["ID"]], short(r[idx["Kernel Name"]])] + [r[idx[m]] for m in cols])


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
