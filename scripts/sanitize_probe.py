#!/usr/bin/env python
"""Small frames of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck):
  single   grid A x 65,536 lights, hierarchical (fused transform + binning, sub-tile segments) and brute force
  rounds   grid B x 60,000 lights in forced rounds with the two-level binning kernel and macro tiles
  routed   3 shard contexts in this process (peer mailboxes, route kernel, two-phase routing), grid A x 30,000 lights, 3 frames
usage: sanitize_probe.py [single|rounds|routed|all]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hellovulkan_b200 import hvcf  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "all"
hvcf.load()
A, SA = (16, 9, 24), (2560, 1600)
B, SB = (32, 18, 48), (3840, 2160)


def frame(grid, screen, n, frames=2, **kw):
    with hvcf.Context(grid=grid, max_lights=n, **kw) as ctx:
        ctx.set_lights(hvcf.host_init_lights(n))
        for f in range(frames):
            ctx.set_ubo(f & 1, hvcf.host_camera_ubo(grid, screen))
            ctx.cull(f & 1)
        return ctx.download_grid((frames - 1) & 1)


if what in ("single", "all"):
    h = frame(A, SA, 65536)
    b = frame(A, SA, 65536, mode=hvcf.MODE_BRUTE)
    assert h[2] == b[2] and h[1].tobytes() == b[1].tobytes()
    print("single ok", h[2])
if what in ("rounds", "all"):
    os.environ.update({"HVCF_ROUNDS": "0"})
    one = frame(B, SB, 60000)
    os.environ.update({"HVCF_ROUNDS": "4", "HVCF_ROUNDS_HEAD": "4", "HVCF_ROUNDS_GROWTH": "2", "HVCF_ROUNDS_MACRO": "4", "HVCF_BIN2": "1", "HVCF_HIER_PARTS": "1"})
    got = frame(B, SB, 60000)
    for k in ("HVCF_ROUNDS", "HVCF_ROUNDS_HEAD", "HVCF_ROUNDS_GROWTH", "HVCF_ROUNDS_MACRO", "HVCF_BIN2", "HVCF_HIER_PARTS"):
        os.environ.pop(k)
    assert one[2] == got[2] and one[1].tobytes() == got[1].tobytes()
    print("rounds ok", got[2])
if what in ("routed", "all"):
    import torch
    n, cuts = 30000, (0, 15, 19, 24)
    lights = hvcf.host_init_lights(n)
    want = frame(A, SA, n)
    os.environ.update({"HVCF_ROUNDS": "3", "HVCF_ROUNDS_HEAD": "4", "HVCF_ROUNDS_GROWTH": "4"})
    world = len(cuts) - 1
    ctxs = [hvcf.Context(grid=A, max_lights=n, shard=(r, world, n, cuts, 4)) for r in range(world)]
    streams = [torch.cuda.Stream() for _ in range(world)]
    for c in ctxs:
        b, e = c.shard_light_range()
        c.set_lights_range(lights[b:e], b)
        for f in range(2):
            c.set_ubo(f, hvcf.host_camera_ubo(A, SA))
    torch.cuda.synchronize()
    for r, c in enumerate(ctxs):
        c.peer_attach_local(r, ctxs)
    for f in range(3):
        for c, s in zip(ctxs, streams):
            c.cull_routed(f & 1, stream=s.cuda_stream)
        for c, s in zip(ctxs, streams):
            assert c.peer_status(stream=s.cuda_stream) == 0
    torch.cuda.synchronize()
    for c in ctxs:
        g = c.download_grid(0)
        assert g[2] == want[2] and g[1].tobytes() == want[1].tobytes()
    for c in ctxs:
        c.peer_detach()
    for c in ctxs:
        c.close()
    print("routed ok", want[2])
