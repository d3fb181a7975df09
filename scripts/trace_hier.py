"""Per-CTA timeline of the hierarchical kernels (debug; HVCF_HIER_TRACE=1)."""
import os, sys
os.environ["HVCF_HIER_TRACE"] = "1"
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from hellovulkan_b200 import hvcf
W = {"A65k": ((16, 9, 24), (2560, 1600), 65536), "B1M": ((32, 18, 48), (3840, 2160), 1 << 20)}
for name in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["A65k"]):
    grid, screen, n = W[name]
    lights = hvcf.host_init_lights(n); ubo = hvcf.host_camera_ubo(grid, screen)
    with hvcf.Context(grid=grid, max_lights=n, mode=2) as ctx:
        ctx.set_lights(lights); ctx.set_ubo(0, ubo)
        for _ in range(3): ctx.cull(0)
        torch.cuda.synchronize()
        for pi, t in enumerate(ctx.debug_trace()):
            t = t.astype(np.int64); t0 = t[:, 0].min()
            st, en, dur = t[:, 0] - t0, t[:, 1] - t0, t[:, 1] - t[:, 0]
            ncta = len(t)
            print(f"{name} pass{pi+1}: {ncta} CTAs, kernel span {en.max()/1e3:.1f} us, CTA dur mean {dur.mean()/1e3:.2f} p50 {np.median(dur)/1e3:.2f} p90 {np.percentile(dur,90)/1e3:.2f} max {dur.max()/1e3:.2f} us; sum {dur.sum()/1e3:.0f} us = {dur.sum()/1e3/148:.1f} us/SM")
            # by start-time decile
            order = np.argsort(st)
            for d in range(0, 10):
                sel = order[d * ncta // 10:(d + 1) * ncta // 10]
                print(f"   decile {d}: start {st[sel].min()/1e3:6.1f}-{st[sel].max()/1e3:6.1f} us  dur mean {dur[sel].mean()/1e3:6.2f} max {dur[sel].max()/1e3:6.2f}  l2tests(w0) mean {t[sel,3].mean():8.0f}")
            ncb = ncta // max(1, (ncta // 18 if name == 'A65k' else ncta // 60))
            # by cb (cluster group)
            per = (18 if name == "A65k" else 60) * (1 if os.environ.get("HVCF_HIER_V1") else 2)
            cb = np.arange(ncta) % per
            print("   by cb (mean dur us):", " ".join(f"{dur[cb==c].mean()/1e3:.1f}" for c in range(per)))
            conc = [(np.sum((st <= x) & (en > x))) for x in np.linspace(0, en.max(), 12)[1:-1]]
            print("   concurrent CTAs over time:", conc)
