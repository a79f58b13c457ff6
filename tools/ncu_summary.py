#!/usr/bin/env python
"""Condenses `ncu -i x.ncu-rep --page raw --csv` into one line per launch: duration, registers, occupancy, active
lanes per instruction, issue utilisation, L1 / L2 hit rates and DRAM bytes (profiles/r02_*_summary.txt).

  python tools/ncu_summary.py gpurun_out/r02_build3_full_raw.csv [--group]     --group: mean per kernel name"""
import csv
import re
import sys
from collections import OrderedDict

COLS = OrderedDict([
    ("us", "gpu__time_duration.sum"), ("regs", "launch__registers_per_thread"), ("grid", "launch__grid_size"),
    ("occ%", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("lanes", "smsp__thread_inst_executed_per_inst_executed.ratio"),
    ("issue%", "smsp__issue_active.avg.pct_of_peak_sustained_active"), ("l1hit%", "l1tex__t_sector_hit_rate.pct"),
    ("l2hit%", "lts__t_sector_hit_rate.pct"), ("dram_rd_MB", "dram__bytes_read.sum"), ("dram_wr_MB", "dram__bytes_write.sum"),
    ("dram%", "dram__throughput.avg.pct_of_peak_sustained_elapsed"), ("winst_M", "smsp__inst_executed.sum"),
])
SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6,
         "usecond": 1.0, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6}


def load(path):
    rows = list(csv.reader(open(path)))
    start = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[start], rows[start + 1]
    out = []
    for r in rows[start + 2:]:
        if len(r) != len(hdr):
            continue
        d = {"name": re.sub(r"\(.*", "", r[hdr.index("Kernel Name")]).split("::")[-1]}
        for key, col in COLS.items():
            if col not in hdr:
                d[key] = None
                continue
            i = hdr.index(col)
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                d[key] = None
                continue
            u = units[i]
            if key in ("us", "dram_rd_MB", "dram_wr_MB"):
                v *= SCALE.get(u, 1.0)
            if key == "winst_M":
                v *= 1e-6
            d[key] = v
        out.append(d)
    return out


def main():
    path = sys.argv[1]
    rows = load(path)
    if "--group" in sys.argv:
        groups = OrderedDict()
        for d in rows:
            groups.setdefault(d["name"], []).append(d)
        rows = []
        for name, ds in groups.items():
            g = {"name": "%s x%d" % (name, len(ds))}
            for k in COLS:
                vals = [d[k] for d in ds if d[k] is not None]
                g[k] = (sum(vals) if k in ("us", "dram_rd_MB", "dram_wr_MB", "winst_M") else sum(vals) / len(vals)) if vals else None
            rows.append(g)
    keys = list(COLS)
    print("%-34s " % "kernel" + " ".join("%10s" % k for k in keys))
    for d in rows:
        print("%-34s " % d["name"][:34] + " ".join("%10s" % ("-" if d[k] is None else ("%.1f" % d[k] if abs(d[k]) < 1e6 else "%.3g" % d[k])) for k in keys))


if __name__ == "__main__":
    main()
