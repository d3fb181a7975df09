#!/usr/bin/env python
"""What the rounds of a frame look like on the device: live groups per round, survivors per (group, macro tile) slot of each round.
usage: rounds_probe.py [WORKLOAD]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hellovulkan_b200 import hvcf  # noqa: E402

W = {"A1M": ((16, 9, 24), (2560, 1600), 1 << 20), "B1M": ((32, 18, 48), (3840, 2160), 1 << 20), "C16M": ((64, 36, 96), (3840, 2160), 1 << 24)}
grid, screen, n = W[sys.argv[1] if len(sys.argv) > 1 else "C16M"]
hvcf.load()
with hvcf.Context(grid=grid, max_lights=n) as ctx:
    ctx.set_lights(hvcf.host_init_lights(n))
    ctx.set_ubo(0, hvcf.host_camera_ubo(grid, screen))
    ctx.cull(0)
    ncb, K, tiles_cap, rounds = ctx.rounds_plan()
    alive = ctx.rounds_state("alive")
    bc = ctx.rounds_state("bin_count")
    st = ctx.stats(0)
print("groups", ncb, "K", K, "rounds", len(rounds), "stats", st)
for r, rd in enumerate(rounds):
    M = 1 << rd["macro_shift"]
    a = np.ones(ncb, bool) if r == 0 else alive[r] != 0
    heads = np.arange(rd["tile_begin"], rd["tile_end"], M)
    c = bc[a][:, heads].astype(np.int64)
    cap = 512 * M
    nz = c[c > 0]
    print(f"round {r}: tiles [{rd['tile_begin']},{rd['tile_end']}) tps {rd['tiles_per_seg']} segs {rd['seg_end'] - rd['seg_begin']} M {M} pipe {rd['pipe']}: "
          f"live groups {int(a.sum())}/{ncb}; slots {c.size}, non-empty {nz.size}, records {int(np.minimum(c, cap).sum())} (+{int((c > cap).sum())} raw), "
          f"per non-empty slot mean {nz.mean() if nz.size else 0:.1f} p50 {np.median(nz) if nz.size else 0:.0f} max {nz.max() if nz.size else 0}; "
          f"per live group records/tile {np.minimum(c, cap).sum(axis=1).mean() / max(1, rd['tile_end'] - rd['tile_begin']):.1f}")
    if r:
        g = np.nonzero(a)[0]
        per = np.minimum(c, cap).sum(axis=1) / max(1, rd["tile_end"] - rd["tile_begin"])
        order = np.argsort(-per)[:12]
        print("   busiest live groups (group: records per tile):", ", ".join(f"{g[i]}:{per[i]:.0f}" for i in order))
