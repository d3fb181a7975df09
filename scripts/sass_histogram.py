#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of the shipped library (cuobjdump -sass hellovulkan_b200/lib/libhvcf.so): the mnemonics
that show which hardware paths a kernel uses -- UBLKCP (TMA bulk copy, either direction), SYNCS (mbarrier), LDGSTS (cp.async),
FADD2 / FMUL2 (packed FP32), FFMA (must be 0 in every kernel on the parity path: --fmad=false + _rn intrinsics), REDUX / VOTE /
SHFL, ATOM / RED, MEMBAR, UCGABAR / CCTL (cluster barrier), ACQBULK (griddepcontrol.wait).  No GPU needed.
usage: sass_histogram.py [lib.so] > profiles/<round>_sass_histogram.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "hellovulkan_b200", "lib", "libhvcf.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
names = dict(zip(re.findall(r"Function : (\S+)", out), demangle))
WATCH = ["UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "FADD2", "FMUL2", "FFMA2", "FFMA", "FADD", "FMUL", "FMNMX", "REDUX", "VOTE", "SHFL", "ATOM", "RED",
         "MEMBAR", "UCGABAR", "ACQBULK", "LDG", "STG", "LDS", "STS", "BAR"]
rows = []
cur, hist = None, None
for line in out.split("\n"):
    m = re.search(r"Function : (\S+)", line)
    if m:
        if cur:
            rows.append((cur, hist))
        cur, hist = m.group(1), collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and cur:
        op = m.group(1)
        hist["total"] += 1
        for w in WATCH:
            if op == w or op.startswith(w + "."):
                hist[w] += 1
                break
        else:
            for w in ("UBLKCP", "UTMALDG", "UTMASTG", "SYNCS", "LDGSTS", "REDUX", "UCGABAR", "ACQBULK", "ATOM", "RED"):
                if op.startswith(w):
                    hist[w] += 1
                    break
if cur:
    rows.append((cur, hist))
cols = ["total", "UBLKCP", "SYNCS", "LDGSTS", "FADD2", "FMUL2", "FFMA", "FFMA2", "FADD", "FMUL", "FMNMX", "REDUX", "VOTE", "SHFL", "ATOM", "RED", "MEMBAR", "UCGABAR", "ACQBULK"]
print(f"# {os.path.relpath(lib, ROOT)}: {len(rows)} kernels; cuobjdump -sass, instructions per kernel by mnemonic")
print(f"{'kernel':78s} " + " ".join(f"{c:>7s}" for c in cols))
for mangled, h in sorted(rows, key=lambda r: names.get(r[0], r[0])):
    n = names.get(mangled, mangled)
    n = re.sub(r"^void ", "", n)
    n = re.sub(r"\([^()]*\)$", "", n).replace("hvcf::", "").replace("(int)", "").replace("(bool)", "")
    print(f"{n[:78]:78s} " + " ".join(f"{h.get(c, 0):7d}" for c in cols))
tot = collections.Counter()
for _, h in rows:
    tot.update(h)
print(f"{'ALL':78s} " + " ".join(f"{tot.get(c, 0):7d}" for c in cols))
