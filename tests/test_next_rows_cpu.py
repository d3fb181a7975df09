"""CPU tests of the rows next to the culling path: the oracle's restatement of the Scene.frag light
loop and of FrustumCulling.comp against fixtures generated from the compiled reference (tests/golden/golden_next.*),
against the compiled reference itself where oracle/_ref exists, and the host-side producers."""
import hashlib
import json
import os

import numpy as np
import pytest

from cpu_checkers import Oracle, Reference

HERE = os.path.dirname(os.path.abspath(__file__))
OFFAXIS = {"pos": (4.0, 3.0, -5.0), "target": (-1.0, 0.5, 2.0)}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(HERE, "golden", "golden_next.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(HERE, "golden", "golden_next.npz"))


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()  # host-side producers only: no GPU is touched
    return hvcf


@pytest.fixture(scope="module")
def shade_scene(oracle, gold, hv):
    meta, _ = gold
    m = meta["shade"]
    grid, screen = tuple(m["grid"]), tuple(m["screen"])
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(m["lights"])
    cells, idx, total, _ = oracle.cull(ubo, aabb, lights)
    assert total == m["total_indices"]
    g = hv.host_init_gbuffer(*screen, seed=m["gbuffer_seed"])
    return ubo, cells, idx, lights, g


def test_host_generators_are_pinned(gold, hv):
    meta, _ = gold
    g = hv.host_init_gbuffer(*meta["shade"]["screen"], seed=meta["shade"]["gbuffer_seed"])
    for k, h in meta["shade"]["gbuffer_sha256"].items():
        assert sha(g[k]) == h, k
    assert g["frag_depth"].min() >= 0.0 and g["frag_depth"].max() < 1.0
    n = g["normal_rough"][:, :3].astype(np.float64)
    assert np.allclose((n * n).sum(1), 1.0, atol=1e-5)
    f = meta["frustum"]
    assert sha(hv.host_init_boxes(f["boxes"], f["extent"], seed=f["seed"])) == f["boxes_sha256"]


@pytest.mark.parametrize("case", ["ref_camera", "offaxis"])
def test_frustum_oracle_matches_golden(oracle, gold, hv, case):
    meta, arrays = gold
    f = meta["frustum"]
    cam = OFFAXIS if case == "offaxis" else {}
    ubo = oracle.frustum_ubo((2560, 1600), **cam)
    assert ubo.tobytes() == arrays[f"frustum_ubo_{case}"].tobytes()
    assert hv.host_frustum_ubo((2560, 1600), **cam).tobytes() == ubo.tobytes()  # include/hv/Camera.h GetFrustumUBO
    boxes = hv.host_init_boxes(f["boxes"], f["extent"], seed=f["seed"])
    cmds = oracle.frustum_cull(ubo, boxes)
    want = np.unpackbits(arrays[f"frustum_visible_{case}"])[: f["boxes"]]
    assert np.array_equal(cmds["instance_count"], want)
    assert int(cmds["instance_count"].sum()) == f["cases"][case]["visible"] > 0
    assert sha(cmds["instance_count"]) == f["cases"][case]["instance_count_sha256"]
    # only instanceCount is written (FrustumCulling.comp:50-57)
    seeded = np.zeros(len(boxes), cmds.dtype)
    seeded["vertex_count"], seeded["first_vertex"], seeded["first_instance"] = 36, 7, np.arange(len(boxes))
    seeded["instance_count"] = 99
    out = oracle.frustum_cull(ubo, boxes, commands=seeded)
    assert np.array_equal(out["instance_count"], want) and np.array_equal(out["first_instance"], seeded["first_instance"])
    assert np.all(out["vertex_count"] == 36) and np.all(out["first_vertex"] == 7)


def test_frustum_edge_cases(oracle):
    ubo = oracle.frustum_ubo((2560, 1600))
    assert len(oracle.frustum_cull(ubo, np.zeros(0, [("min_point", "<f4", 4), ("max_point", "<f4", 4)]))) == 0
    boxes = np.zeros(4, [("min_point", "<f4", 4), ("max_point", "<f4", 4)])
    boxes["min_point"][0], boxes["max_point"][0] = (-1000, -1000, -1000, 1), (1000, 1000, 1000, 1)  # contains the frustum
    boxes["min_point"][1], boxes["max_point"][1] = (0, 2, 1, 1), (0, 2, 1, 1)                       # a point in view
    boxes["min_point"][2], boxes["max_point"][2] = (0, 1, 20, 1), (1, 2, 21, 1)                     # behind the camera
    boxes["min_point"][3], boxes["max_point"][3] = (np.nan, 0, 0, 1), (1, 1, 1, 1)                  # NaN never counts as outside
    got = oracle.frustum_cull(ubo, boxes)["instance_count"]
    assert got.tolist() == [1, 1, 0, 1]


@pytest.mark.parametrize("case", ["default", "tuned"])
def test_shade_oracle_matches_golden(oracle, gold, shade_scene, case):
    meta, arrays = gold
    m = meta["shade"]
    ubo, cells, idx, lights, g = shade_scene
    pc = np.array(m["cases"][case]["pc"], np.float32)
    color, cluster = oracle.shade(ubo, cells, idx, lights, g, pc=pc)
    assert sha(cluster) == m["cases"][case]["cluster_sha256"]
    assert np.array_equal(cluster, arrays["shade_cluster"].astype(np.uint32))
    want = arrays[f"shade_color_{case}"]
    got = color[:: m["stride"]]
    # same libm, same operation order: the restatement reproduces the compiled shader bit for bit
    assert got.tobytes() == want.tobytes()
    assert np.allclose(color[:, :3].astype(np.float64).sum(0), m["cases"][case]["color_sum"], rtol=1e-9)
    assert np.all(color[:, 3] == 1.0)


def test_shade_cluster_lookup_properties(oracle, shade_scene):
    ubo, cells, idx, lights, g = shade_scene
    color, cluster = oracle.shade(ubo, cells, idx, lights, g)
    X, Y, Z = 16, 9, 24
    W, H = g["width"], g["height"]
    p = np.arange(W * H)
    # tile x/y follow the pixel grid (Scene.frag:66-72)
    assert np.array_equal(cluster % X, ((p % W) + 0.5) // (W / X))
    assert np.array_equal((cluster // X) % Y, ((p // W) + 0.5) // (H / Y))
    assert cluster.max() < X * Y * Z
    # a cluster without lights leaves Lo = albedo * albedoMultipler = 0 (Scene.frag:117)
    empty = cells["count"][cluster] == 0
    assert empty.any() and np.all(color[empty, :3] == 0.0)
    assert np.all(color[:, :3] >= 0.0) and np.isfinite(color).all()
    # stride / range sampling evaluates exactly the sampled pixels
    part, _ = oracle.shade(ubo, cells, idx, lights, g, begin=1000, end=50000, stride=7)
    sel = np.arange(1000, 50000, 7)
    assert np.array_equal(part[sel], color[sel])
    mask = np.ones(W * H, bool)
    mask[sel] = False
    assert np.all(part[mask] == 0)


@pytest.mark.skipif(not Reference.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_next_rows_oracle_equals_compiled_reference(oracle, hv):
    R = Reference()
    for grid, screen, n in (((16, 9, 24), (512, 320), 4000), ((32, 18, 48), (384, 216), 20000)):
        ubo = R.camera_ubo(grid, screen, **OFFAXIS)
        assert ubo.tobytes() == oracle.camera_ubo(grid, screen, **OFFAXIS).tobytes()
        aabb = R.generate_aabbs(ubo)
        lights = R.init_lights(n)
        cells, idx, total, _ = oracle.cull(ubo, aabb, lights)
        g = hv.host_init_gbuffer(*screen, seed=11, **OFFAXIS)
        pc = np.array([1.5, 0.04, 4.0, 2.0, 0.1], np.float32)
        a = oracle.shade(ubo, cells, idx, lights, g, pc=pc, camera_position=OFFAXIS["pos"])
        b = R.shade(ubo, cells, idx, lights, g, pc=pc, camera_position=OFFAXIS["pos"])
        assert np.array_equal(a[1], b[1])
        assert a[0].tobytes() == b[0].tobytes()
    boxes = hv.host_init_boxes(50000, 80.0, seed=99)
    for screen, cam in (((2560, 1600), {}), ((1600, 1200), OFFAXIS)):
        fo, fr = oracle.frustum_ubo(screen, **cam), R.frustum_ubo(screen, **cam)
        assert fo.tobytes() == fr.tobytes()
        assert np.array_equal(oracle.frustum_cull(fo, boxes), R.frustum_cull(fr, boxes))
