"""GPU parity of the rows next to the culling path through the C ABI:
  hvcf_shade         cluster index per pixel EXACT; colour within the float tolerance below
  hvcf_frustum_cull  bit-exact instanceCount, other command words untouched
Tolerance of the light loop (float, not reproducible bit for bit: the shader's pow/normalize/divide have no
fixed rounding even between Vulkan drivers): |gpu - oracle| <= 2e-4 * |oracle| + 2e-5 per channel, where the oracle
equals the reference shader compiled as C++ bit for bit (tests/test_next_rows_cpu.py)."""
import json
import os

import numpy as np
import pytest

from cpu_checkers import AABB_DT, CMD_DT

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
OFFAXIS = {"pos": (4.0, 3.0, -5.0), "target": (-1.0, 0.5, 2.0)}
RTOL, ATOL = 2e-4, 2e-5


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(HERE, "golden", "golden_next.json")) as f:
        meta = json.load(f)
    return meta, np.load(os.path.join(HERE, "golden", "golden_next.npz"))


def gpu_shade(hv, grid, ubo, lights, g, pc, cam, cap=150, mode=0):
    P = g["width"] * g["height"]
    with hv.Context(grid=grid, max_lights=max(len(lights), 1), max_lights_per_cluster=cap, mode=mode) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo)
        ctx.cull(0)
        dg = hv.DeviceGBuffer(g)
        try:
            ctx.shade(0, dg, pc=hv.PushConstPBR(*[float(v) for v in pc]), camera_position=cam)
            color = dg.out_color.download(np.float32).reshape(P, 4)
            cluster = dg.out_cluster.download(np.uint32)
            cells = ctx.download_light_cells()
            idx = ctx.download_light_indices(ctx.download_total(0))
        finally:
            dg.close()
    return color, cluster, cells, idx


def assert_colors_close(got, want):
    err = np.abs(got.astype(np.float64) - want.astype(np.float64))
    bound = RTOL * np.abs(want.astype(np.float64)) + ATOL
    bad = err > bound
    assert not bad.any(), f"{int(bad.sum())} of {bad.size} channels off; worst {err.max():.3e} vs {want[np.unravel_index(err.argmax(), err.shape)]:.3e}"


@pytest.mark.parametrize("case", ["default", "tuned"])
def test_shade_matches_golden_and_oracle(hv, oracle, gold, case):
    meta, arrays = gold
    m = meta["shade"]
    grid, screen = tuple(m["grid"]), tuple(m["screen"])
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(m["lights"])
    g = hv.host_init_gbuffer(*screen, seed=m["gbuffer_seed"])
    pc = np.array(m["cases"][case]["pc"], np.float32)
    color, cluster, cells, idx = gpu_shade(hv, grid, ubo, lights, g, pc, (0.0, 1.0, 6.0))
    assert np.array_equal(cluster, arrays["shade_cluster"].astype(np.uint32))          # exact, against the compiled reference
    assert_colors_close(color[:: m["stride"]], arrays[f"shade_color_{case}"])
    want, wcluster = oracle.shade(ubo, cells, idx, lights, g, pc=pc)                  # every pixel, against the oracle
    assert np.array_equal(cluster, wcluster)
    assert_colors_close(color, want)
    assert np.all(color[:, 3] == 1.0)


@pytest.mark.parametrize("grid,screen,n", [((16, 9, 24), (2560, 1600), 65536), ((32, 18, 48), (3840, 2160), 200000),
                                           ((16, 9, 24), (1000, 563), 3000)])
def test_shade_full_size_sampled_against_oracle(hv, oracle, grid, screen, n):
    """BASELINE sizes (2560x1600 with 65,536 lights: lists at the 150 cap; 3840x2160 on grid B; an odd size whose
    tiles do not divide the screen): cluster index of every pixel exact, colours checked on a strided sample."""
    ubo = oracle.camera_ubo(grid, screen, **OFFAXIS) if screen[0] == 1000 else oracle.camera_ubo(grid, screen)
    cam = OFFAXIS["pos"] if screen[0] == 1000 else (0.0, 1.0, 6.0)
    lights = oracle.init_lights(n)
    g = hv.host_init_gbuffer(*screen, seed=3, **(OFFAXIS if screen[0] == 1000 else {}))
    pc = np.array([1.0, 0.04, 4.0, 1.0, 0.0], np.float32)
    color, cluster, cells, idx = gpu_shade(hv, grid, ubo, lights, g, pc, cam)
    P = screen[0] * screen[1]
    stride = max(1, P // 60000) | 1
    want, wcluster = oracle.shade(ubo, cells, idx, lights, g, pc=pc, camera_position=cam, stride=stride)
    sel = np.arange(0, P, stride)
    assert np.array_equal(cluster[sel], wcluster[sel])
    assert_colors_close(color[sel], want[sel])
    # every pixel's cluster index against the oracle's lookup alone (hvo_cluster_index, Scene.frag:55-78)
    ys, xs = np.divmod(np.arange(P), screen[0])
    probe = np.arange(0, P, 7)
    ref_idx = np.array([oracle.cluster_index(ubo, float(xs[p]) + 0.5, float(ys[p]) + 0.5, float(g["frag_depth"][p])) for p in probe[:20000]],
                       np.uint32)
    assert np.array_equal(cluster[probe[:20000]], ref_idx)
    assert cells["count"][cluster].max() > 0


def test_shade_slice_boundaries_exact(hv, oracle):
    """Depths placed on and one ulp around every z-slice boundary: the kernel's table lookup must agree with
    uint(max(log2(linDepth) * scaling + bias, 0)) evaluated by the oracle's libm."""
    grid, screen = (16, 9, 24), (256, 100)
    ubo = oracle.camera_ubo(grid, screen)
    near, far, Z = 0.1, 100.0, 24
    W, H = screen
    g = hv.host_init_gbuffer(W, H, seed=5)
    depths = []
    for k in range(0, Z + 1):
        lin = near * (far / near) ** (k / Z)
        z = (near * far / lin - far) / (near - far)  # invert LinearDepth
        z32 = np.float32(min(max(z, 0.0), 0.99999))
        for step in range(-40, 41):
            v = z32
            for _ in range(abs(step)):
                v = np.nextafter(v, np.float32(2.0 if step > 0 else -1.0), dtype=np.float32)
            depths.append(min(max(float(v), 0.0), 0.99999))
    d = np.resize(np.array(depths, np.float32), W * H)
    g["frag_depth"] = d
    lights = oracle.init_lights(64)
    color, cluster, cells, idx = gpu_shade(hv, grid, ubo, lights, g, np.array([1, 0.04, 4, 1, 0], np.float32), (0.0, 1.0, 6.0))
    _, want = oracle.shade(ubo, cells, idx, lights, g)
    assert np.array_equal(cluster, want)
    assert len(np.unique(cluster // (16 * 9))) == Z


def test_shade_errors(hv, oracle):
    grid = (16, 9, 24)
    g = hv.host_init_gbuffer(64, 40)
    with hv.Context(grid=grid, max_lights=16) as ctx:
        dg = hv.DeviceGBuffer(g)
        try:
            with pytest.raises(hv.HvcfError):  # no UBO for the frame yet
                ctx.shade(0, dg)
            ctx.set_ubo(0, oracle.camera_ubo(grid, (64, 40)))
            with pytest.raises(hv.HvcfError):
                ctx.shade(5, dg)
        finally:
            dg.close()


@pytest.mark.parametrize("count", [0, 1, 255, 10000, 1 << 20])
def test_frustum_cull_bit_exact(hv, oracle, count):
    boxes = hv.host_init_boxes(count, 60.0, seed=4242)
    for screen, cam in (((2560, 1600), {}), ((1600, 1200), OFFAXIS)):
        f = oracle.frustum_ubo(screen, **cam)
        assert hv.host_frustum_ubo(screen, **cam).tobytes() == f.tobytes()
        seeded = np.zeros(count, CMD_DT)
        seeded["vertex_count"], seeded["first_vertex"], seeded["instance_count"] = 36, 7, 99
        seeded["first_instance"] = np.arange(count)
        want = oracle.frustum_cull(f, boxes, commands=seeded)
        db, dc = hv.DeviceBuffer.from_array(boxes), hv.DeviceBuffer.from_array(seeded)
        try:
            hv.frustum_cull(f, db.ptr, count, dc.ptr)
            got = dc.download(CMD_DT, count)
        finally:
            db.close(), dc.close()
        assert got.tobytes() == want.tobytes()


def test_frustum_cull_golden_and_edges(hv, oracle, gold):
    meta, arrays = gold
    f = meta["frustum"]
    boxes = hv.host_init_boxes(f["boxes"], f["extent"], seed=f["seed"])
    for case, cam in (("ref_camera", {}), ("offaxis", OFFAXIS)):
        ubo = arrays[f"frustum_ubo_{case}"]
        db, dc = hv.DeviceBuffer.from_array(boxes), hv.DeviceBuffer.from_array(np.zeros(len(boxes), CMD_DT))
        try:
            hv.frustum_cull(ubo, db.ptr, len(boxes), dc.ptr)
            got = dc.download(CMD_DT, len(boxes))["instance_count"]
        finally:
            db.close(), dc.close()
        assert np.array_equal(got, np.unpackbits(arrays[f"frustum_visible_{case}"])[: len(boxes)])
        assert int(got.sum()) == f["cases"][case]["visible"]
    edge = np.zeros(4, AABB_DT)
    edge["min_point"][0], edge["max_point"][0] = (-1000, -1000, -1000, 1), (1000, 1000, 1000, 1)
    edge["min_point"][1], edge["max_point"][1] = (0, 2, 1, 1), (0, 2, 1, 1)
    edge["min_point"][2], edge["max_point"][2] = (0, 1, 20, 1), (1, 2, 21, 1)
    edge["min_point"][3], edge["max_point"][3] = (np.nan, 0, 0, 1), (1, 1, 1, 1)
    ubo = oracle.frustum_ubo((2560, 1600))
    db, dc = hv.DeviceBuffer.from_array(edge), hv.DeviceBuffer.from_array(np.zeros(4, CMD_DT))
    try:
        hv.frustum_cull(ubo, db.ptr, 4, dc.ptr)
        assert dc.download(CMD_DT, 4)["instance_count"].tolist() == [1, 1, 0, 1]
        nan_f = ubo.copy()
        nan_f.view(np.float32)[24] = np.nan  # corners[0].x: the "frustum outside box" x tests can no longer fire
        hv.frustum_cull(nan_f, db.ptr, 4, dc.ptr)
        assert np.array_equal(dc.download(CMD_DT, 4)["instance_count"], oracle.frustum_cull(nan_f, edge)["instance_count"])
    finally:
        db.close(), dc.close()


def test_headless_app_end_to_end(hv, oracle, tmp_path):
    """hellovulkan_b200/lib/hv_cluster_forward_app: the C++ drop-in classes (include/hv/ClusterForward.h) driven in the
    order of AppPBRClusterForward::Init / UpdateUBOs / FillCommandBuffer for 3 frames (moving light 0 and the camera);
    every buffer it dumps is compared with the oracle."""
    import subprocess
    from cpu_checkers import CELL_DT, LIGHT_DT
    app = os.path.join(os.path.dirname(HERE), "hellovulkan_b200", "lib", "hv_cluster_forward_app")
    assert os.path.exists(app), "build() did not produce the headless driver"
    W, H, n = 640, 400, 5000
    prefix = str(tmp_path / "run")
    r = subprocess.run([app, "16", "9", "24", str(W), str(H), str(n), "12345", "3", prefix], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    ubo = np.fromfile(prefix + ".ubo", np.uint8)
    cam = np.fromfile(prefix + ".camera", np.float32)
    assert np.allclose(cam, (1.0, 1.0, 6.0))                                  # frame 2 of MainLoop
    assert ubo.tobytes() == oracle.camera_ubo((16, 9, 24), (W, H), pos=tuple(cam)).tobytes()
    lights = np.fromfile(prefix + ".lights", LIGHT_DT)
    want_lights = oracle.init_lights(n)
    want_lights["position"][0] = (0.5, 2.0, 1.0, 1.0)                          # UpdateLightPosition of frame 2
    assert lights.tobytes() == want_lights.tobytes()
    aabb = np.fromfile(prefix + ".aabb", AABB_DT)
    assert aabb.tobytes() == oracle.generate_aabbs(ubo).tobytes()
    cells, idx = np.fromfile(prefix + ".cells", CELL_DT), np.fromfile(prefix + ".indices", np.uint32)
    wc, wi, total, _ = oracle.cull(ubo, aabb, lights)
    assert cells.tobytes() == wc.tobytes() and idx.tobytes() == wi.tobytes()
    g = hv.host_init_gbuffer(W, H, seed=7)
    color = np.fromfile(prefix + ".color", np.float32).reshape(-1, 4)
    cluster = np.fromfile(prefix + ".cluster", np.uint32)
    want, wcluster = oracle.shade(ubo, cells, idx, lights, g, camera_position=tuple(cam))
    assert np.array_equal(cluster, wcluster)
    assert_colors_close(color, want)
    frustum = np.fromfile(prefix + ".frustum", np.uint8)
    assert frustum.tobytes() == oracle.frustum_ubo((W, H), pos=tuple(cam)).tobytes()
    cmds = np.fromfile(prefix + ".commands", CMD_DT)
    boxes = hv.host_init_boxes(10000, 60.0, seed=4242)
    seeded = np.zeros(10000, CMD_DT)
    seeded["vertex_count"], seeded["instance_count"], seeded["first_instance"] = 36, 1, np.arange(10000)
    assert cmds.tobytes() == oracle.frustum_cull(frustum, boxes, commands=seeded).tobytes()
