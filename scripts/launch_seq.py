#!/usr/bin/env python
"""Prints the kernel launches of the LAST frame found in an ncu launch list (--metrics gpu__time_duration.sum --csv):
name, grid, duration.  Usage: launch_seq.py launches.csv [first_kernel_substring]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
hdr = rows[h]
kn, mv, gs = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
first = sys.argv[2] if len(sys.argv) > 2 else "light_transform"
recs = [(r[kn].split("(")[0].replace("void ", ""), r[gs], float(r[mv].replace(",", "")) / 1e3) for r in rows[h + 1:] if len(r) > mv]
starts = [i for i, r in enumerate(recs) if first in r[0]]
if not starts:
    sys.exit("no frame found")
seq = recs[starts[-1]:]
# cut at the next kernel that does not belong to a frame
names = ("light_transform", "hier", "shard_", "peer_")
out = []
for r in seq:
    if not any(n in r[0] for n in names):
        break
    out.append(r)
tot = sum(r[2] for r in out)
for r in out:
    print(f"{r[0][:48]:48s} grid {r[1]:>16s} {r[2]:9.1f} us")
print(f"{'sum':48s} {'':21s} {tot:9.1f} us over {len(out)} launches")
