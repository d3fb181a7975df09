"""Step time of both culling modes on the BASELINE workloads (CUDA events, L2 not flushed; quick tuning aid)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from hellovulkan_b200 import hvcf
W = {"A1k": ((16, 9, 24), (2560, 1600), 1000), "A65k": ((16, 9, 24), (2560, 1600), 65536), "A1M": ((16, 9, 24), (2560, 1600), 1 << 20),
     "B1M": ((32, 18, 48), (3840, 2160), 1 << 20)}
names = sys.argv[1].split(",") if len(sys.argv) > 1 else list(W)
modes = sys.argv[2].split(",") if len(sys.argv) > 2 else ["brute", "hier"]
st = torch.cuda.current_stream(); sp = st.cuda_stream
for name in names:
    grid, screen, n = W[name]
    lights = hvcf.host_init_lights(n); ubo = hvcf.host_camera_ubo(grid, screen)
    for m in modes:
        with hvcf.Context(grid=grid, max_lights=n, mode={"brute": 1, "hier": 2}[m]) as ctx:
            ctx.set_timing(True); ctx.set_lights(lights, sp); ctx.set_ubo(0, ubo, sp)
            for _ in range(5): ctx.cull(0, sp)
            torch.cuda.synchronize()
            K = 20; a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
            a.record(st)
            for _ in range(K): ctx.cull(0, sp)
            b.record(st); torch.cuda.synchronize()
            km, tm = ctx.timing(0); s = ctx.stats(0, sp)
            print(f"{name} {m}: step {a.elapsed_time(b)/K*1e3:.1f} us  main kernel {km*1e3:.1f} us  lib total {tm*1e3:.1f} us  "
                  f"exec {s['executed_tests']:.3e} pre {s['prefilter_tests']:.3e} nominal {s['nominal_tests']:.3e} total_idx {s['total_indices']} launches {s['kernel_launches']}", flush=True)
