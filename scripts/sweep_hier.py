"""Scratch tuning tool (not bench.py): hierarchical mode, tiles per segment x blocks per warp.
Usage: python scripts/sweep_hier.py [A65k|A1M|B1M] ...   (TPS=1,2,4 K=1,2,4 in the environment narrow the sweep)"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hellovulkan_b200 import hvcf  # noqa: E402

CONFIGS = {"A1k": ((16, 9, 24), (2560, 1600), 1000), "A65k": ((16, 9, 24), (2560, 1600), 65536),
           "A1M": ((16, 9, 24), (2560, 1600), 1 << 20), "B1M": ((32, 18, 48), (3840, 2160), 1 << 20)}


def main():
    names = sys.argv[1:] or ["A65k", "A1M", "B1M"]
    tps_list = [int(v) for v in os.environ.get("TPS", "0,1,2,4,8,16,32").split(",")]
    k_list = [int(v) for v in os.environ.get("K", "0,1,2,4").split(",")]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for name in names:
        grid, screen, n = CONFIGS[name]
        ubo = hvcf.host_camera_ubo(grid, screen)
        lights = hvcf.host_init_lights(n)
        for k in k_list:
            for tps in tps_list:
                os.environ["HVCF_HIER_TPS"] = str(tps)
                os.environ["HVCF_HIER_K"] = str(k)
                try:
                    with hvcf.Context(grid=grid, max_lights=n, mode=hvcf.MODE_HIERARCHICAL) as ctx:
                        ctx.set_lights(lights)
                        ctx.set_ubo(0, ubo)
                        for _ in range(3):
                            ctx.cull(0)
                        torch.cuda.synchronize()
                        times = []
                        for _ in range(20):
                            flush.zero_()
                            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                            e0.record()
                            ctx.cull(0)
                            e1.record()
                            torch.cuda.synchronize()
                            times.append(e0.elapsed_time(e1))
                        st = ctx.stats(0)
                        kt = None
                        if os.environ.get("HVCF_KERNEL_TIMES") == "1":
                            acc = np.zeros(5)
                            for _ in range(10):
                                flush.zero_()
                                ctx.cull(0)
                                acc += np.array(ctx.kernel_times(0))
                            kt = [round(float(v) * 100, 1) for v in acc]  # us, mean of 10
                except Exception as e:
                    print(json.dumps({"config": name, "K": k, "tps": tps, "error": str(e)[:100]}), flush=True)
                    continue
                print(json.dumps({"config": name, "K": k, "tps": tps, "ms": round(float(np.median(times)), 4), "min_ms": round(min(times), 4),
                                  "executed": st["executed_tests"], "prefilter": st["prefilter_tests"], "total": st["total_indices"],
                                  "kernel_us(transform,bin,count,finalize,write)": kt}), flush=True)


if __name__ == "__main__":
    main()
