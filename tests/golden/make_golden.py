"""Generate tests/golden/golden_ref.{npz,json} from the REFERENCE BUILD (oracle/_ref/libhvcf_ref.so =
the reference's Camera.cpp + AABBGenerator.comp + LightCulling.comp compiled as C++ against its
vendored glm; see oracle/refbuild/build_ref.sh).  Run in the build container, where /root/reference
exists:   python tests/golden/make_golden.py
The fixtures travel to the GPU box; /root/reference does not.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from cpu_checkers import Reference, SCENE_SEED  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gathered(cells, idx):
    parts = [idx[int(o): int(o) + int(c)] for o, c in zip(cells["offset"], cells["count"])]
    return np.concatenate(parts) if parts else np.zeros(0, np.uint32)


def main():
    R = Reference()
    arrays, meta = {}, {"seed": SCENE_SEED, "source": "oracle/_ref/libhvcf_ref.so (reference sources compiled as C++)",
                        "max_lights_per_cluster": 150, "configs": {}}
    grids = {"A": ((16, 9, 24), (2560, 1600)), "B": ((32, 18, 48), (3840, 2160)), "C": ((64, 36, 96), (3840, 2160)),
             "A_default_screen": ((16, 9, 24), (1600, 1200))}
    for name, (grid, screen) in grids.items():
        ubo = R.camera_ubo(grid, screen)
        aabb = R.generate_aabbs(ubo)
        arrays[f"ubo_{name}"] = ubo
        entry = {"grid": grid, "screen": screen, "aabb_sha256": sha(aabb)}
        if name.startswith("A"):
            arrays[f"aabb_{name}"] = aabb.view(np.float32).reshape(-1, 8)
        else:
            arrays[f"aabb_{name}_head"] = aabb[:16].view(np.float32).reshape(-1, 8)
            arrays[f"aabb_{name}_tail"] = aabb[-16:].view(np.float32).reshape(-1, 8)
        meta["configs"][name] = entry
    # an off-axis camera too (yaw/pitch path of SetPositionAndTarget)
    ubo2 = R.camera_ubo((16, 9, 24), (2560, 1600), pos=(4.0, 3.0, -5.0), target=(-1.0, 0.5, 2.0))
    arrays["ubo_A_offaxis"] = ubo2
    meta["configs"]["A_offaxis"] = {"grid": (16, 9, 24), "screen": (2560, 1600), "pos": (4.0, 3.0, -5.0),
                                    "target": (-1.0, 0.5, 2.0), "aabb_sha256": sha(R.generate_aabbs(ubo2))}

    uboA = arrays["ubo_A"]
    aabbA = R.generate_aabbs(uboA)
    for n in (1000, 65536):
        lights = R.init_lights(n)
        cells, idx, total = R.cull(uboA, aabbA, lights, threads=1)
        g = gathered(cells, idx)
        m = {"lights_sha256": sha(lights), "total": total, "counts_sha256": sha(cells["count"]),
             "gathered_indices_sha256": sha(g), "overflow_clusters": int((cells["count"] == 150).sum()),
             "max_count": int(cells["count"].max())}
        arrays[f"lights_{n}_head"] = lights[:8].view(np.float32).reshape(-1, 12)
        if n == 1000:
            arrays["cull_A_1000_counts"] = cells["count"].copy()
            arrays["cull_A_1000_gathered"] = g
        meta[f"cull_A_{n}"] = m
    lights = R.init_lights(1000)
    cells, idx, total = R.cull(ubo2, R.generate_aabbs(ubo2), lights, threads=1)
    meta["cull_A_offaxis_1000"] = {"total": total, "counts_sha256": sha(cells["count"]),
                                   "gathered_indices_sha256": sha(gathered(cells, idx))}
    np.savez_compressed(os.path.join(HERE, "golden_ref.npz"), **arrays)
    with open(os.path.join(HERE, "golden_ref.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(json.dumps(meta, indent=1)[:1500])


if __name__ == "__main__":
    main()
