"""Scratch tuning tool (not bench.py): time every cull-kernel variant x segment count on the GPU.
Usage: python scripts/time_variants.py [A65k|B1M|A1k] ..."""
import os
import sys
import json

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from cpu_checkers import Oracle  # noqa: E402  (inputs only; scratch tool)
from hellovulkan_b200 import hvcf  # noqa: E402

CONFIGS = {"A1k": ((16, 9, 24), (2560, 1600), 1000), "A65k": ((16, 9, 24), (2560, 1600), 65536),
           "B1M": ((32, 18, 48), (3840, 2160), 1 << 20)}


def main():
    names = sys.argv[1:] or ["A65k", "B1M"]
    O = Oracle()
    variants = [int(v) for v in os.environ.get("VARIANTS", "0,1,2,3,4,5").split(",")]
    seg_list = [int(v) for v in os.environ.get("SEGS", "0").split(",")]
    for name in names:
        grid, screen, n = CONFIGS[name]
        ubo = O.camera_ubo(grid, screen)
        lights = O.init_lights(n)
        for variant in variants:
            for segs in seg_list:
                os.environ["HVCF_CULL_VARIANT"] = str(variant)
                os.environ["HVCF_SEGMENTS"] = str(segs)
                with hvcf.Context(grid=grid, max_lights=n) as ctx:
                    ctx.set_lights(lights)
                    ctx.set_ubo(0, ubo)
                    for _ in range(3):
                        ctx.cull(0)
                    torch.cuda.synchronize()
                    times = []
                    for _ in range(10):
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record()
                        ctx.cull(0)
                        e1.record()
                        torch.cuda.synchronize()
                        times.append(e0.elapsed_time(e1))
                    st = ctx.stats(0)
                ms = float(np.median(times))
                print(json.dumps({"config": name, "variant": variant, "segs": segs, "ms": round(ms, 4), "min_ms": round(min(times), 4),
                                  "nominal_Gtests_s": round(st["nominal_tests"] / ms / 1e6, 2),
                                  "executed_Gtests_s": round(st["executed_tests"] / ms / 1e6, 2),
                                  "executed_frac": round(st["executed_tests"] / max(1, st["nominal_tests"]), 3),
                                  "total": st["total_indices"]}), flush=True)


if __name__ == "__main__":
    main()
