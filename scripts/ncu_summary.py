#!/usr/bin/env python
"""Per-kernel table from an `ncu --page raw --csv` export (profiles/*_full_*_raw.csv) or a launch list
(`--metrics gpu__time_duration.sum` CSV).  Usage: ncu_summary.py raw.csv [launches.csv]"""
import csv
import sys
from collections import OrderedDict, defaultdict

METRICS = [
    ("duration us", "gpu__time_duration.sum", 1e-3),
    ("dram read MB", "dram__bytes_read.sum", None),
    ("dram write MB", "dram__bytes_write.sum", None),
    ("regs", "launch__registers_per_thread", 1),
    ("warps active %", "sm__warps_active.avg.pct_of_peak_sustained_active", 1),
    ("pipe alu %", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1),
    ("pipe fma %", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", 1),
    ("sm thr %", "sm__throughput.avg.pct_of_peak_sustained_elapsed", 1),
    ("issue active %", "sm__inst_issued.avg.pct_of_peak_sustained_active", 1),
    ("warp inst M", "smsp__inst_executed.sum", 1e-6),
    ("l2 hit %", "lts__t_sector_hit_rate.pct", 1),
    ("dram thr %", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1),
]


def to_float(s):
    try:
        return float(s.replace(",", ""))
    except Exception:
        return None


def unit_scale(unit):
    u = unit.lower()
    return {"byte": 1e-6, "kbyte": 1e-3, "mbyte": 1.0, "gbyte": 1e3, "ns": 1.0, "us": 1e3, "ms": 1e6}.get(u)


def raw_table(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    kn = col["Kernel Name"]
    out = []
    for r in data:
        d = OrderedDict(kernel=r[kn].split("(")[0])
        for label, m, scale in METRICS:
            if m not in col:
                continue
            v = to_float(r[col[m]])
            if v is None:
                continue
            if scale is None:                      # bytes: normalise by the unit row
                v *= unit_scale(units[col[m]]) or 1.0
            elif m == "gpu__time_duration.sum":
                v *= (unit_scale(units[col[m]]) or 1.0) * 1e-3
            else:
                v *= scale
            d[label] = v
        # top stall reasons (warps per issue)
        stalls = []
        for h, i in col.items():
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio") \
                    and "not_issued" not in h:
                v = to_float(r[i])
                if v:
                    stalls.append((v, h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        stalls.sort(reverse=True)
        d["top stalls"] = ", ".join(f"{n} {v:.1f}" for v, n in stalls[:3])
        out.append(d)
    return out


def launch_table(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = rows[0]
    col = {h: i for i, h in enumerate(hdr)}
    agg = defaultdict(list)
    for r in rows[1:]:
        if r[col["Metric Name"]] != "gpu__time_duration.sum":
            continue
        v = to_float(r[col["Metric Value"]])
        u = unit_scale(r[col["Metric Unit"]]) or 1.0
        agg[r[col["Kernel Name"]].split("(")[0]].append(v * u * 1e-3)
    return agg


if __name__ == "__main__":
    for d in raw_table(sys.argv[1]):
        print(" | ".join(f"{k}: {v:.2f}" if isinstance(v, float) else f"{k}: {v}" for k, v in d.items()))
    if len(sys.argv) > 2:
        print("-- launch list (us: n, median, min) --")
        for k, v in launch_table(sys.argv[2]).items():
            v = sorted(v)
            print(f"{k}: n={len(v)} median={v[len(v)//2]:.2f} min={v[0]:.2f}")
