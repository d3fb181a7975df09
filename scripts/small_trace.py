import os, sys, ctypes
os.environ["HVCF_HIER_TRACE"] = "1"
sys.path.insert(0, "/root/repo")
import numpy as np
from hellovulkan_b200 import hvcf
hvcf.load()
grid, screen, n = (16, 9, 24), (2560, 1600), 1000
with hvcf.Context(grid=grid, max_lights=n) as ctx:
    ctx.set_lights(hvcf.host_init_lights(n)); ctx.set_ubo(0, hvcf.host_camera_ubo(grid, screen))
    for _ in range(5):
        ctx.cull(0)
    buf = np.zeros(36 * 8 + 64, np.uint64)
    nn = ctypes.c_size_t(0)
    hvcf._check(ctx.lib.hvcf_debug_trace(ctx.handle, buf.ctypes.data_as(ctypes.c_void_p), 36 * 8, ctypes.byref(nn)))
    t = buf[:36 * 8].reshape(36, 8).astype(np.int64)
    t0 = t[:, 0].min()
    print("phase ends relative to the first CTA start, us: start, transform, test, scan, write, end")
    for b in (0, 1, 5, 17, 35):
        print(b, [round((x - t0) / 1e3, 2) for x in t[b, :6]])
    d = np.diff(t[:, :6], axis=1) / 1e3
    print("mean phase us:", d.mean(axis=0).round(2), "max:", d.max(axis=0).round(2))
