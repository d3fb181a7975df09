"""Generate tests/golden/golden_next.{npz,json}: fixtures of the rows next to the culling path from the
REFERENCE BUILD (oracle/_ref/libhvcf_ref.so: Camera.cpp, FrustumCulling.comp, Radiance.glsl, PBRHeader.glsl and the
light-loop lines of Scene.frag compiled as C++; oracle/refbuild/build_ref.sh).  Inputs that have no reference
counterpart (G-buffer, instance boxes) come from this repo's seeded host generators and are pinned by hash.
Run in the build container, where /root/reference exists:   python tests/golden/make_golden_next.py
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from cpu_checkers import Reference  # noqa: E402
from hellovulkan_b200 import hvcf  # noqa: E402  (host-side generators only; no GPU needed)

SHADE_SCREEN = (640, 400)
SHADE_STRIDE = 61
OFFAXIS = {"pos": (4.0, 3.0, -5.0), "target": (-1.0, 0.5, 2.0)}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    R = Reference()
    arrays, meta = {}, {"source": "oracle/_ref/libhvcf_ref.so (reference sources compiled as C++)"}
    # ---- frustum culling ------------------------------------------------------------------------------
    boxes = hvcf.host_init_boxes(10000, 60.0, seed=4242)
    meta["frustum"] = {"boxes": 10000, "extent": 60.0, "seed": 4242, "boxes_sha256": sha(boxes), "cases": {}}
    for name, cam in (("ref_camera", {}), ("offaxis", OFFAXIS)):
        f = R.frustum_ubo((2560, 1600), **cam)
        cmds = R.frustum_cull(f, boxes, threads=1)
        arrays[f"frustum_ubo_{name}"] = f
        arrays[f"frustum_visible_{name}"] = np.packbits(cmds["instance_count"].astype(np.uint8))
        meta["frustum"]["cases"][name] = {"visible": int(cmds["instance_count"].sum()), "instance_count_sha256": sha(cmds["instance_count"])}
    # ---- consumer: Scene.frag light loop ----------------------------------------------------------------
    grid = (16, 9, 24)
    ubo = R.camera_ubo(grid, SHADE_SCREEN)
    aabb = R.generate_aabbs(ubo)
    lights = R.init_lights(1000)
    cells, idx, total = R.cull(ubo, aabb, lights, threads=1)
    g = hvcf.host_init_gbuffer(*SHADE_SCREEN, seed=7)
    pcs = {"default": np.array([1.0, 0.04, 4.0, 1.0, 0.0], np.float32), "tuned": np.array([2.5, 0.06, 4.0, 1.7, 0.05], np.float32)}
    meta["shade"] = {"grid": grid, "screen": SHADE_SCREEN, "lights": 1000, "stride": SHADE_STRIDE, "gbuffer_seed": 7,
                     "gbuffer_sha256": {k: sha(g[k]) for k in ("frag_depth", "world_pos", "normal_rough", "albedo_metal")},
                     "total_indices": total, "cases": {}}
    for name, pc in pcs.items():
        color, cluster = R.shade(ubo, cells, idx, lights, g, pc=pc, threads=1)
        arrays[f"shade_color_{name}"] = color[::SHADE_STRIDE].copy()
        meta["shade"]["cases"][name] = {"pc": [float(v) for v in pc], "cluster_sha256": sha(cluster),
                                        "color_sum": [float(v) for v in color[:, :3].astype(np.float64).sum(0)]}
    arrays["shade_cluster"] = cluster.astype(np.uint16)
    np.savez_compressed(os.path.join(HERE, "golden_next.npz"), **arrays)
    with open(os.path.join(HERE, "golden_next.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps(meta, indent=1)[:2500])


if __name__ == "__main__":
    main()
