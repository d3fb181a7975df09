"""world = 1 routed frame (route to self): per-kernel times; used alone and under ncu (-k regex:shard_route)."""
import os, sys
os.environ["HVCF_KERNEL_TIMES"] = "1"
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np, torch
from hellovulkan_b200 import hvcf
name = sys.argv[1] if len(sys.argv) > 1 else "C16M"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 0
W = {"B1M": ((32, 18, 48), (3840, 2160), 1 << 20), "C16M": ((64, 36, 96), (3840, 2160), 1 << 24), "C2M": ((64, 36, 96), (3840, 2160), 1 << 21)}
grid, screen, n = W[name]
lights = hvcf.host_init_lights(n)
ubo = hvcf.host_camera_ubo(grid, screen)
with hvcf.Context(grid=grid, max_lights=n, shard=(0, 1, n, (0, grid[2]), B)) as c:
    c.set_lights_range(lights, 0)
    c.set_ubo(0, ubo); c.set_ubo(1, ubo)
    for f in range(4):
        c.cull_routed(f & 1)
    torch.cuda.synchronize()
    print(name, "B", B, "kernel ms (route+wait, bin, count, finalize, write):", [round(x, 4) for x in c.kernel_times(1)], "route alone:", round(c.route_ms(1), 4),
          "inbox", c.shard_inbox_records(), "total", c.download_total(1))
with hvcf.Context(grid=grid, max_lights=n) as c:
    c.set_lights(lights); c.set_ubo(0, ubo); c.set_ubo(1, ubo)
    for f in range(4):
        c.cull(f & 1)
    torch.cuda.synchronize()
    print(name, "plain hvcf_cull kernel ms (transform, bin, count, finalize, write):", [round(x, 4) for x in c.kernel_times(1)], "total", c.download_total(1))
