"""The oracle (oracle/hvcf_oracle.c) against the golden vectors generated from the reference build,
against the reference build itself when it is present, and on hand-made scenes
(SURVEY.md section 4).  CPU only."""
import hashlib

import numpy as np
import pytest

from cpu_checkers import (AABB_DT, LIGHT_DT, Oracle, Reference, same_light_sets, MAX_LIGHTS_PER_CLUSTER)

GRIDS = {"A": ((16, 9, 24), (2560, 1600)), "B": ((32, 18, 48), (3840, 2160)), "C": ((64, 36, 96), (3840, 2160)),
         "A_default_screen": ((16, 9, 24), (1600, 1200))}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gathered(cells, idx):
    parts = [idx[int(o): int(o) + int(c)] for o, c in zip(cells["offset"], cells["count"])]
    return np.concatenate(parts) if parts else np.zeros(0, np.uint32)


def test_layouts():
    assert LIGHT_DT.itemsize == 48 and LIGHT_DT.fields["color"][1] == 16 and LIGHT_DT.fields["radius"][1] == 32
    assert AABB_DT.itemsize == 32 and AABB_DT.fields["max_point"][1] == 16


@pytest.mark.parametrize("name", list(GRIDS))
def test_camera_and_aabb_match_golden(oracle, golden, name):
    meta, arrs = golden
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    assert np.array_equal(ubo, arrs[f"ubo_{name}"])
    aabb = oracle.generate_aabbs(ubo)
    assert sha(aabb) == meta["configs"][name]["aabb_sha256"]
    f = aabb.view(np.float32).reshape(-1, 8)
    if name.startswith("A"):
        assert np.array_equal(f.view(np.uint32), arrs[f"aabb_{name}"].view(np.uint32))
    else:
        assert np.array_equal(f[:16].view(np.uint32), arrs[f"aabb_{name}_head"].view(np.uint32))
        assert np.array_equal(f[-16:].view(np.uint32), arrs[f"aabb_{name}_tail"].view(np.uint32))


def test_camera_known_answers(oracle):
    # inverse projection of the reference camera at 2560x1600
    f = oracle.camera_ubo((16, 9, 24), (2560, 1600)).view(np.float32)
    assert f[0] == np.float32(0.662741601) and f[5] == np.float32(-0.414213449)
    assert f[11] == np.float32(-9.98999977) and f[14] == np.float32(-1.0) and f[15] == np.float32(10.0)
    aabb = oracle.generate_aabbs(oracle.camera_ubo((16, 9, 24), (2560, 1600)))
    assert np.allclose(aabb[0]["min_point"][:3], [-0.088378, 0.0322166, -0.133352], rtol=1e-5)
    assert np.allclose(aabb[0]["max_point"][:3], [-0.0579899, 0.0552363, -0.1], rtol=1e-5)
    assert aabb[-1]["min_point"][2] == np.float32(-100.0)
    assert abs(aabb[-1]["max_point"][2] - (-74.9894)) < 1e-3
    assert np.all(aabb["min_point"][:, 3] == 0) and np.all(aabb["max_point"][:, 3] == 0)
    assert np.all(aabb["min_point"][:, :3] <= aabb["max_point"][:, :3])


def test_offaxis_camera(oracle, golden):
    meta, arrs = golden
    c = meta["configs"]["A_offaxis"]
    ubo = oracle.camera_ubo(c["grid"], c["screen"], pos=c["pos"], target=c["target"])
    assert np.array_equal(ubo, arrs["ubo_A_offaxis"])
    aabb = oracle.generate_aabbs(ubo)
    assert sha(aabb) == c["aabb_sha256"]
    lights = oracle.init_lights(1000)
    cells, idx, total, _ = oracle.cull(ubo, aabb, lights, threads=2)
    g = meta["cull_A_offaxis_1000"]
    assert total == g["total"] and sha(cells["count"]) == g["counts_sha256"]
    assert sha(gathered(cells, idx)) == g["gathered_indices_sha256"]


@pytest.mark.parametrize("n", [1000, 65536])
def test_scene_and_cull_match_golden(oracle, golden, n):
    meta, arrs = golden
    lights = oracle.init_lights(n)
    g = meta[f"cull_A_{n}"]
    assert sha(lights) == g["lights_sha256"]
    assert np.array_equal(lights[:8].view(np.float32).reshape(-1, 12).view(np.uint32), arrs[f"lights_{n}_head"].view(np.uint32))
    ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600))
    aabb = oracle.generate_aabbs(ubo)
    cells, idx, total, executed = oracle.cull(ubo, aabb, lights)
    assert total == g["total"] and int(cells["count"].max()) == g["max_count"]
    assert int((cells["count"] == 150).sum()) == g["overflow_clusters"]
    assert sha(cells["count"]) == g["counts_sha256"]
    assert sha(gathered(cells, idx)) == g["gathered_indices_sha256"]
    if n == 1000:
        assert np.array_equal(cells["count"], arrs["cull_A_1000_counts"])
        assert np.array_equal(gathered(cells, idx), arrs["cull_A_1000_gathered"])
        assert executed == len(aabb) * n  # no cluster reaches the cap
    else:
        assert executed < len(aabb) * n
    # this repo's own invariant: offsets are the exclusive scan of counts in cluster order
    assert np.array_equal(cells["offset"], np.concatenate([[0], np.cumsum(cells["count"])[:-1]]).astype(np.uint32))


def test_serial_and_threaded_agree(oracle):
    ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600))
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(4000)
    a = oracle.cull(ubo, aabb, lights, threads=1)
    b = oracle.cull(ubo, aabb, lights, threads=4)
    assert a[2] == b[2] and a[3] == b[3] and a[0].tobytes() == b[0].tobytes() and np.array_equal(a[1], b[1])


def test_literal_per_pair_transform_equals_hoisted(oracle):
    """hvo_cull hoists cameraView*position out of the pair loop; hvo_cull_one keeps it inside like
    TestSphereAABB (LightCulling.comp:49-57).  Same bits."""
    ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600))
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(3000)
    cells, idx, _, _ = oracle.cull(ubo, aabb, lights)
    for c in (0, 7, 500, 1234, 2000, 3455):
        one, _ = oracle.cull_one(ubo, aabb[c:c + 1], lights)
        assert np.array_equal(one, idx[cells["offset"][c]: cells["offset"][c] + cells["count"][c]])


def _unit_scene(oracle):
    """identity view matrix, one hand-made cluster box [-1,1]^3 shifted to z in [-3,-1]"""
    ubo = oracle.camera_ubo((1, 1, 1), (100, 100))
    f = ubo.view(np.float32)
    f[16:32] = np.eye(4, dtype=np.float32).reshape(-1)
    aabb = np.zeros(1, AABB_DT)
    aabb[0]["min_point"] = (-1, -1, -3, 0)
    aabb[0]["max_point"] = (1, 1, -1, 0)
    return ubo, aabb


def _light(x, y, z, r):
    l = np.zeros(1, LIGHT_DT)
    l[0]["position"] = (x, y, z, 1)
    l[0]["color"] = (1, 1, 1, 1)
    l[0]["radius"] = r
    return l


def test_handmade_hit_miss_strictness(oracle):
    ubo, aabb = _unit_scene(oracle)
    cases = [
        (_light(0, 0, -2, 0.5), True),      # centre of the box
        (_light(0, 0, -2, 0.0), False),     # radius 0 never hits: 0 < 0 is false (LightCulling.comp:56)
        (_light(3, 0, -2, 2.0), False),     # exactly `radius` away on one axis: strict '<' -> miss
        (_light(3, 0, -2, np.nextafter(np.float32(2.0), np.float32(3.0))), True),
        (_light(0, 0, 5, 1.0), False),      # behind the camera, far away
        (_light(2, 2, 0, 1.75), True),      # corner distance sqrt(1+1+1)=1.732
        (_light(2, 2, 0, 1.7), False),
    ]
    for l, expect in cases:
        idx, executed = oracle.cull_one(ubo, aabb, l)
        assert (len(idx) == 1) == expect
        assert executed == 1


def test_cap_keeps_ascending_prefix(oracle):
    ubo, aabb = _unit_scene(oracle)
    n = 400
    lights = np.zeros(n, LIGHT_DT)
    lights["position"] = (0, 0, -2, 1)
    lights["radius"] = 1.0
    lights["radius"][::3] = 0.0  # every third light misses
    idx, executed = oracle.cull_one(ubo, aabb, lights)
    hits = np.nonzero(lights["radius"] > 0)[0].astype(np.uint32)
    assert len(idx) == MAX_LIGHTS_PER_CLUSTER and np.array_equal(idx, hits[:150])
    assert executed == int(hits[149]) + 1  # the loop stops right after the 150th hit
    cells, gi, total, ex = oracle.cull(ubo, aabb, lights, cap=64)
    assert total == 64 and np.array_equal(gi, hits[:64]) and ex == int(hits[63]) + 1
    cells, gi, total, ex = oracle.cull(ubo, aabb, lights[:0])
    assert total == 0 and cells["count"][0] == 0 and ex == 0


def test_boundary_report_small(oracle):
    ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600))
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(1000)
    near1, flips = oracle.boundary_report(ubo, aabb, lights)
    assert near1 >= 0 and flips >= 0 and flips < 100  # boundary pairs are rare
    a = oracle.cull(ubo, aabb, lights, use_fma=False)
    b = oracle.cull(ubo, aabb, lights, use_fma=True)
    assert abs(a[2] - b[2]) <= flips


def test_cluster_index_convention(oracle):
    ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600))
    assert oracle.cluster_index(ubo, 0.5, 0.5, 0.0) == 0          # near plane, top-left tile
    assert oracle.cluster_index(ubo, 2559.5, 0.5, 0.0) == 15      # x fastest
    assert oracle.cluster_index(ubo, 0.5, 1599.5, 0.0) == 8 * 16  # then y
    # depth 1.0 is the far plane (GLM_FORCE_DEPTH_ZERO_TO_ONE): log2(100)*scaling+bias = 24 -> slice index 24
    # (one past the grid: the reference shader does not clamp, Scene.frag:65)
    assert oracle.cluster_index(ubo, 0.5, 0.5, 0.999) // (16 * 9) == 21   # linear depth 50.0
    assert oracle.cluster_index(ubo, 0.5, 0.5, 0.9999) // (16 * 9) == 23  # linear depth 90.9


@pytest.mark.skipif(not Reference.available(), reason="oracle/_ref not built (needs /root/reference)")
class TestAgainstReferenceBuild:
    def test_ubo_aabb_lights(self, oracle):
        R = Reference()
        for grid, screen in GRIDS.values():
            uo, ur = oracle.camera_ubo(grid, screen), R.camera_ubo(grid, screen)
            assert np.array_equal(uo, ur)
            assert oracle.generate_aabbs(uo).tobytes() == R.generate_aabbs(ur).tobytes()
        for pos, tgt in [((4, 3, -5), (-1, 0.5, 2)), ((0, 8, 0.5), (0, 0, 0)), ((-7, 0.2, 1), (3, 4, -2))]:
            uo, ur = oracle.camera_ubo((16, 9, 24), (1920, 1080), pos, tgt), R.camera_ubo((16, 9, 24), (1920, 1080), pos, tgt)
            assert np.array_equal(uo, ur)
        for seed in (1, 12345, 987654321):
            assert oracle.init_lights(5000, seed).tobytes() == R.init_lights(5000, seed).tobytes()

    @pytest.mark.parametrize("pos,tgt,n", [((0, 1, 6), (0, 2.5, 0), 5000), ((4, 3, -5), (-1, 0.5, 2), 3000),
                                           ((0, 12, 0.5), (0, 0, 0), 20000)])
    def test_cull(self, oracle, pos, tgt, n):
        R = Reference()
        ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600), pos, tgt)
        aabb = oracle.generate_aabbs(ubo)
        lights = oracle.init_lights(n, seed=n)
        co, io, tot, _ = oracle.cull(ubo, aabb, lights)
        cr, ir, totr = R.cull(ubo, aabb, lights, threads=1)
        assert tot == totr and same_light_sets(co, io, cr, ir)
        assert np.array_equal(co["offset"], cr["offset"])  # serial reference = cluster-order scan
        # the threaded reference races on atomicAdd: offsets differ, per-cluster lists do not
        cr2, ir2, totr2 = R.cull(ubo, aabb, lights, threads=4)
        assert totr2 == tot and same_light_sets(co, io, cr2, ir2)

    @pytest.mark.parametrize("n", [0, 1, 4096])
    def test_cull_edge_scenes(self, oracle, n):
        """the scene of tests/test_gpu_parity.py::test_handmade_boundaries (radius 0 / tiny / huge, NaN and +-Inf positions, w != 1)
        and the empty and one-light scenes: the restatement and the compiled shader source agree on every list, so the GPU
        parity tests on these scenes are pinned to the reference itself"""
        R = Reference()
        ubo = oracle.camera_ubo((16, 9, 24), (2560, 1600))
        aabb = oracle.generate_aabbs(ubo)
        lights = oracle.init_lights(n, seed=11)
        if n > 1:
            rng = np.random.default_rng(5)
            lights["radius"][::7] = 0.0
            lights["position"][5::97, 0] = np.nan
            lights["position"][9::101, 1] = np.inf
            lights["position"][3::89, 2] = -np.inf
            lights["radius"][13::53] = np.float32(1e-20)
            lights["radius"][17::59] = np.float32(1e18)
            lights["position"][:, 3] = 1.0
            lights["position"][21::61, 3] = rng.uniform(0.5, 2.0, size=len(lights["position"][21::61, 3])).astype(np.float32)
        co, io, tot, _ = oracle.cull(ubo, aabb, lights)
        cr, ir, totr = R.cull(ubo, aabb, lights, threads=1)
        assert tot == totr and np.array_equal(co["offset"], cr["offset"]) and same_light_sets(co, io, cr, ir)
        if n > 1:
            assert tot > 0 and (co["count"] == 150).any()      # the huge radii fill clusters to the cap

