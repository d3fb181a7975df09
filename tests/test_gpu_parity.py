"""GPU parity: libhvcf.so (through its C ABI) against the CPU oracle on the same seeded inputs.
Bit-exact: AABB bytes equal; per-cluster (count, index sequence) equal; offsets = exclusive scan."""
import os

import numpy as np
import pytest

from cpu_checkers import AABB_DT, LIGHT_DT, Oracle, same_light_sets

pytestmark = pytest.mark.gpu

MODES = {"brute": 1, "hier": 2, "small": 2}  # hvcf_mode: HVCF_MODE_BRUTE, HVCF_MODE_HIERARCHICAL; "small" = hierarchical with the
# single-kernel path for frames of <= 2,048 lights switched on (conftest.py), as it is by default outside the tests
GRIDS = {"A": ((16, 9, 24), (2560, 1600)), "B": ((32, 18, 48), (3840, 2160)), "C": ((64, 36, 96), (3840, 2160))}


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


def check_invariants(cells, indices, total, n_lights, cap):
    cnt = cells["count"].astype(np.int64)
    assert np.all(cnt <= cap)
    assert int(cnt.sum()) == total
    assert np.array_equal(cells["offset"].astype(np.int64), np.concatenate([[0], np.cumsum(cnt)[:-1]]))
    idx = indices[:total]
    assert idx.size == 0 or int(idx.max()) < n_lights
    # strictly ascending inside each cluster
    if total > 1:
        d = np.diff(idx.astype(np.int64))
        starts = cells["offset"][cnt > 0].astype(np.int64)
        boundary = np.zeros(total, bool)
        boundary[starts] = True
        assert np.all((d > 0) | boundary[1:])


@pytest.mark.parametrize("name", list(GRIDS))
def test_aabb_bit_exact(hv, oracle, golden, name):
    meta, _ = golden
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    with hv.Context(grid=grid, max_lights=1024) as ctx:
        ctx.set_ubo(0, ubo)
        assert ctx.is_aabb_dirty()
        ctx.generate_aabbs(0)
        assert not ctx.is_aabb_dirty()
        got = ctx.download_aabbs()
    want = oracle.generate_aabbs(ubo)
    assert got.tobytes() == want.tobytes()
    import hashlib
    assert hashlib.sha256(got.tobytes()).hexdigest() == meta["configs"][name]["aabb_sha256"]


def run_cull(hv, grid, ubo, lights, cap=150, frame=0, **kw):
    with hv.Context(grid=grid, max_lights=max(len(lights), 1), max_lights_per_cluster=cap, **kw) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(frame, ubo)
        ctx.cull(frame)
        cells = ctx.download_light_cells()
        total = ctx.download_total(frame)
        idx = ctx.download_light_indices(total)
        stats = ctx.stats(frame)
    return cells, idx, total, stats


@pytest.mark.parametrize("mode", list(MODES))
@pytest.mark.parametrize("n", [0, 1, 31, 64, 255, 256, 257, 1000, 1025, 2047, 2048, 2049, 65536])
def test_cull_config_A(hv, oracle, n, mode):
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(n)
    cells, idx, total, stats = run_cull(hv, grid, ubo, lights, mode=MODES[mode])
    oc, oi, ototal, oexec = oracle.cull(ubo, aabb, lights)
    assert total == ototal and stats["total_indices"] == ototal
    assert np.array_equal(cells["count"], oc["count"])
    assert same_light_sets(cells, idx, oc, oi)
    assert np.array_equal(cells["offset"], oc["offset"]) and np.array_equal(idx, oi)
    check_invariants(cells, idx, total, n, 150)
    assert stats["nominal_tests"] == len(aabb) * n
    assert stats["overflow_clusters"] == int((oc["count"] == 150).sum())
    if mode == "brute":
        assert stats["executed_tests"] >= oexec or n == 0  # tile-granular early exit never does less than the exact bound
        assert stats["prefilter_tests"] == 0
    elif mode == "small" and n <= 2048:
        assert stats["plan_flags"] == 64 and stats["kernel_launches"] <= 2      # HVCF_PLAN_SINGLE_KERNEL (+ AABB generation)
        assert stats["executed_tests"] == stats["nominal_tests"] and stats["prefilter_tests"] == 0
    elif n:
        assert stats["prefilter_tests"] > 0
        assert stats["executed_tests"] + stats["prefilter_tests"] < stats["nominal_tests"] or n < 4096


@pytest.mark.parametrize("mode", list(MODES))
def test_cull_matches_golden_reference_lists(hv, oracle, golden, mode):
    meta, arrs = golden
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(1000)
    cells, idx, total, _ = run_cull(hv, grid, ubo, lights, mode=MODES[mode])
    assert total == meta["cull_A_1000"]["total"]
    assert np.array_equal(cells["count"], arrs["cull_A_1000_counts"])
    assert np.array_equal(idx, arrs["cull_A_1000_gathered"])


@pytest.mark.parametrize("segs", ["1", "3", "0"])
def test_brute_force_segment_counts_agree(hv, oracle, segs, monkeypatch):
    monkeypatch.setenv("HVCF_SEGMENTS", segs)
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen, pos=(4, 3, -5), target=(-1, 0.5, 2))
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(20000, seed=7)
    cells, idx, total, _ = run_cull(hv, grid, ubo, lights, mode=MODES["brute"])
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert total == ototal and np.array_equal(cells["count"], oc["count"]) and np.array_equal(idx, oi)


@pytest.mark.parametrize("k", ["1", "2", "4"])
@pytest.mark.parametrize("tps", ["0", "1", "2", "8", "64"])
@pytest.mark.parametrize("shape", ["", "4,4,2", "32,1,1", "2,2,8"])
def test_hier_plans_agree(hv, oracle, k, tps, shape, monkeypatch):
    """every block shape / blocks-per-warp / tiles-per-segment plan of the hierarchical mode gives the same bytes"""
    monkeypatch.setenv("HVCF_HIER_K", k)
    monkeypatch.setenv("HVCF_HIER_TPS", tps)
    if shape:
        monkeypatch.setenv("HVCF_HIER_SHAPE", shape)
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen, pos=(4, 3, -5), target=(-1, 0.5, 2))
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(20000, seed=7)
    cells, idx, total, _ = run_cull(hv, grid, ubo, lights, mode=MODES["hier"])
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert total == ototal and np.array_equal(cells["count"], oc["count"]) and np.array_equal(idx, oi)
    assert np.array_equal(cells["offset"], oc["offset"])


@pytest.mark.parametrize("grid", [(1, 1, 1), (3, 2, 5), (5, 7, 3), (33, 2, 2), (16, 9, 1)])
@pytest.mark.parametrize("mode", list(MODES))
def test_odd_grids(hv, oracle, grid, mode):
    ubo = oracle.camera_ubo(grid, (1280, 720))
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(5000, seed=17)
    cells, idx, total, _ = run_cull(hv, grid, ubo, lights, mode=MODES[mode])
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert total == ototal and np.array_equal(cells["count"], oc["count"]) and np.array_equal(idx, oi)
    assert np.array_equal(cells["offset"], oc["offset"])


@pytest.mark.parametrize("mode", list(MODES))
@pytest.mark.parametrize("cap", [1, 7, 64, 4096])
def test_cap_values(hv, oracle, cap, mode):
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(9000, seed=3)
    cells, idx, total, _ = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES[mode])
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights, cap=cap)
    assert total == ototal and np.array_equal(cells["count"], oc["count"]) and np.array_equal(idx, oi)
    check_invariants(cells, idx, total, len(lights), cap)


@pytest.mark.parametrize("mode", list(MODES))
def test_handmade_boundaries(hv, oracle, mode):
    """strict '<', radius 0, light behind the camera, NaN/Inf positions."""
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    rng = np.random.default_rng(5)
    n = 4096
    lights = np.zeros(n, LIGHT_DT)
    # lights placed exactly at cluster corners / face distances in view space need world positions:
    # use the oracle's own random scene and then overwrite special cases
    lights[:] = oracle.init_lights(n, seed=11)
    lights["radius"][::7] = 0.0
    lights["position"][5::97, 0] = np.nan
    lights["position"][9::101, 1] = np.inf
    lights["position"][3::89, 2] = -np.inf
    lights["radius"][13::53] = np.float32(1e-20)
    lights["radius"][17::59] = np.float32(1e18)
    lights["position"][:, 3] = 1.0
    lights["position"][21::61, 3] = rng.uniform(0.5, 2.0, size=len(lights["position"][21::61, 3])).astype(np.float32)
    cells, idx, total, _ = run_cull(hv, grid, ubo, lights, mode=MODES[mode])
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert total == ototal and np.array_equal(cells["count"], oc["count"]) and np.array_equal(idx, oi)


@pytest.mark.parametrize("mode", list(MODES))
def test_frames_update_light_and_dirty_tracking(hv, oracle, mode):
    grid, screen = GRIDS["A"]
    ubo0 = oracle.camera_ubo(grid, screen)
    ubo1 = oracle.camera_ubo(grid, screen, pos=(2, 1.5, 5), target=(0, 2, 0))
    lights = oracle.init_lights(3000, seed=21)
    with hv.Context(grid=grid, max_lights=4000, mode=MODES[mode]) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo0)
        ctx.cull(0)
        assert not ctx.is_aabb_dirty()
        ctx.set_ubo(1, ubo1)  # same projection, new view: AABBs stay valid
        assert not ctx.is_aabb_dirty()
        ctx.update_light(0, (0.5, 2.0, 1.0))  # ResourcesLight::UpdateLightPosition
        ctx.update_light(10_000_000, (0, 0, 0))  # out of range: ignored like the reference
        ctx.cull(1)
        cells = ctx.download_light_cells()
        total = ctx.download_total(1)
        idx = ctx.download_light_indices(total)
        assert ctx.download_total(0) != 0
        # a new screen size changes the projection -> dirty
        ctx.set_ubo(0, oracle.camera_ubo(grid, (1600, 1200)))
        assert ctx.is_aabb_dirty()
    lights2 = lights.copy()
    lights2["position"][0] = (0.5, 2.0, 1.0, 1.0)
    aabb = oracle.generate_aabbs(ubo1)
    oc, oi, ototal, _ = oracle.cull(ubo1, aabb, lights2)
    assert total == ototal and np.array_equal(cells["count"], oc["count"]) and np.array_equal(idx, oi)


@pytest.mark.parametrize("mode", list(MODES))
def test_sharded_contexts_concatenate(hv, oracle, mode):
    """z-slab shards: each writes its cells/indices at GLOBAL positions given the base (sum of the
    lower shards' totals), so segments concatenate without a fix-up pass."""
    import torch
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(30000, seed=9)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    shards = [(0, 6), (6, 12), (12, 24)]
    ctxs = [hv.Context(grid=grid, max_lights=len(lights), z_begin=a, z_end=b, mode=MODES[mode]) for a, b in shards]
    try:
        totals = []
        for c in ctxs:
            c.set_lights(lights)
            c.set_ubo(0, ubo)
            c.cull_begin(0)
            c.cull_finish(0, None)  # base 0: only to read this shard's total back
            b = c.buffers()
            totals.append(int(c.download_light_cells()[b.cluster_begin:b.cluster_end]["count"].sum()))
            assert c.stats(0)["total_indices"] == totals[-1]
        bases = np.concatenate([[0], np.cumsum(totals)[:-1]])
        all_cells = np.zeros(len(aabb), oc.dtype)
        all_idx = np.zeros(ototal, np.uint32)
        for c, base, tot in zip(ctxs, bases, totals):
            base_t = torch.tensor([int(base)], dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            c.cull_finish(0, base_t.data_ptr())
            b = c.buffers()
            all_cells[b.cluster_begin:b.cluster_end] = c.download_light_cells()[b.cluster_begin:b.cluster_end]
            all_idx[int(base):int(base) + tot] = c.download_light_indices(int(base) + tot)[int(base):]
        assert sum(totals) == ototal
        assert np.array_equal(all_cells["count"], oc["count"]) and np.array_equal(all_cells["offset"], oc["offset"])
        assert np.array_equal(all_idx, oi)
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("mode", list(MODES))
@pytest.mark.parametrize("name,n,sample", [("B", 1 << 20, 96), ("C", 200_000, 64)])
def test_large_configs_sampled(hv, oracle, name, n, sample, mode):
    """Full-size grids: the oracle checks a seeded sample of clusters in full; invariants on everything."""
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(n)
    cells, idx, total, stats = run_cull(hv, grid, ubo, lights, mode=MODES[mode])
    check_invariants(cells, idx, total, n, 150)
    rng = np.random.default_rng(123)
    for c in rng.choice(len(aabb), size=sample, replace=False):
        want, _ = oracle.cull_one(ubo, aabb[c:c + 1], lights)
        got = idx[cells["offset"][c]: cells["offset"][c] + cells["count"][c]]
        assert np.array_equal(got, want), f"cluster {c}"
    assert stats["executed_tests"] <= stats["nominal_tests"] + len(aabb) * 1024


def test_modes_give_identical_bytes(hv, oracle):
    """brute force and hierarchical: the same light cells and the same index buffer, byte for byte (grid B)"""
    grid, screen = GRIDS["B"]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(300_000, seed=5)
    a = run_cull(hv, grid, ubo, lights, mode=MODES["brute"])
    b = run_cull(hv, grid, ubo, lights, mode=MODES["hier"])
    assert a[2] == b[2] and a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()


@pytest.mark.parametrize("name,n,z", [("A", 70_000, None), ("B", 300_000, None), ("C", 200_000, None), ("C", 150_000, (40, 52))])
def test_one_and_two_level_binning_give_identical_bytes(hv, oracle, monkeypatch, name, n, z):
    """level 0 of the hierarchy as one kernel per (8 groups, tile) or as the two-level kernel per tile (super-groups first):
    the same survivor slots, hence the same light cells and index buffer, byte for byte; both equal brute force"""
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=11)
    kw = {} if z is None else {"z_begin": z[0], "z_end": z[1]}
    out = {}
    for lvl in ("0", "1"):
        monkeypatch.setenv("HVCF_BIN2", lvl)
        out[lvl] = run_cull(hv, grid, ubo, lights, mode=MODES["hier"], **kw)
    monkeypatch.delenv("HVCF_BIN2")
    ref = run_cull(hv, grid, ubo, lights, mode=MODES["brute"], **kw)
    for lvl in ("0", "1"):
        assert out[lvl][2] == ref[2]
        assert out[lvl][0].tobytes() == ref[0].tobytes() and out[lvl][1].tobytes() == ref[1].tobytes(), f"HVCF_BIN2={lvl}"
    assert out["1"][3]["prefilter_tests"] != out["0"][3]["prefilter_tests"]   # the two kernels really ran


@pytest.mark.parametrize("name,n,cap", [("A", 70_000, 150), ("A", 65_536, 150), ("A", 1, 150), ("A", 0, 150), ("A", 3_000, 4), ("B", 20_000, 150),
                                        ("A", 262_144, 150)])
def test_fused_and_split_transform_binning_give_identical_bytes(hv, oracle, monkeypatch, name, n, cap):
    """frames with few lights run transform + level-0 binning as one kernel (hier_xbin_kernel, HVCF_XBIN_MAX lights at most);
    with the fusion switched off the two kernels run: same light cells and index buffer, same binning test count, one launch more"""
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=23)
    out = {}
    for mx in ("0", "262144"):
        monkeypatch.setenv("HVCF_XBIN_MAX", mx)
        out[mx] = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["hier"])
    monkeypatch.delenv("HVCF_XBIN_MAX")
    a, b = out["0"], out["262144"]
    assert a[2] == b[2] and a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()
    for k in ("bin_tests", "total_indices", "overflow_clusters"):   # the other counters depend on when early-exit knowledge arrives
        assert a[3][k] == b[3][k], k
    assert b[3]["prefilter_tests"] >= b[3]["bin_tests"]
    fused = name == "A"   # grid B has enough cluster groups for the two-level binning kernel, which is not fused
    assert a[3]["kernel_launches"] - b[3]["kernel_launches"] == (1 if fused else 0)
    ref = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["brute"])
    assert ref[2] == b[2] and ref[0].tobytes() == b[0].tobytes() and ref[1].tobytes() == b[1].tobytes()


@pytest.mark.parametrize("name,n,cap,z,tps", [("A", 70_000, 150, None, 0), ("A", 65_536, 150, None, 8), ("A", 3_000, 4, None, 0), ("B", 300_000, 150, None, 0),
                                             ("B", 300_000, 150, None, 32), ("C", 150_000, 150, (40, 52), 16), ("A", 0, 150, None, 0),
                                             ("A", 200_000, 4096, None, 16), ("A", 1, 150, None, 0)])
def test_pipelined_and_plain_slot_walk_give_identical_bytes(hv, oracle, monkeypatch, name, n, cap, z, tps):
    """count / write pass with the survivor slots fetched one batch ahead by cp.async into a double buffer (hier2_kernel<..., PIPE>,
    chosen by the host for long light segments; forced here) against the plain walk: same light cells and index buffer, equal to
    brute force.  tps = tiles per light segment (0 = the plan's own); cap 4096 keeps dense slots (raw-tile batches) alive."""
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=31)
    kw = {} if z is None else {"z_begin": z[0], "z_end": z[1]}
    if tps:
        monkeypatch.setenv("HVCF_HIER_TPS", str(tps))
    out = {}
    for sw in ("0", "1"):
        monkeypatch.setenv("HVCF_BATCH_PIPE", sw)
        out[sw] = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["hier"], **kw)
    monkeypatch.delenv("HVCF_BATCH_PIPE")
    monkeypatch.delenv("HVCF_HIER_TPS", raising=False)
    ref = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["brute"], **kw)
    for sw in ("0", "1"):
        assert out[sw][2] == ref[2], f"HVCF_BATCH_PIPE={sw}"
        assert out[sw][0].tobytes() == ref[0].tobytes() and out[sw][1].tobytes() == ref[1].tobytes(), f"HVCF_BATCH_PIPE={sw}"


@pytest.mark.parametrize("name,n,cap", [("A", 70_000, 150), ("A", 65_536, 150), ("A", 1, 150), ("A", 0, 150), ("A", 3_000, 4), ("B", 100_000, 150),
                                        ("A", 40_000, 4096), ("A", 200_000, 150)])
def test_sub_tile_light_segments_give_identical_bytes(hv, oracle, monkeypatch, name, n, cap):
    """frames of few tiles cut each tile's survivor slot into HVCF_HIER_PARTS batch ranges, one light segment each (shorter serial
    walks in the dense cluster groups): same light cells and index buffer for 1, 2, 3, 4 and 8 parts, equal to brute force"""
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=37)
    ref = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["brute"])
    for parts in ("1", "2", "3", "4", "8"):
        monkeypatch.setenv("HVCF_HIER_PARTS", parts)
        got = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["hier"])
        assert got[2] == ref[2], f"HVCF_HIER_PARTS={parts}"
        assert got[0].tobytes() == ref[0].tobytes() and got[1].tobytes() == ref[1].tobytes(), f"HVCF_HIER_PARTS={parts}"
    monkeypatch.delenv("HVCF_HIER_PARTS")


def test_stats_are_split_per_kernel(hv, oracle):
    """hvcf_stats: the binning kernel's tests are a part of prefilter_tests, the write pass's repeats are counted apart, and the
    count pass's exact tests equal the oracle's executed-test count rounded up to what a tile-granular early exit runs"""
    grid, screen = GRIDS["B"]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(100_000, seed=3)
    _, _, total, st = run_cull(hv, grid, ubo, lights, mode=MODES["hier"])
    assert 0 < st["bin_tests"] < st["prefilter_tests"]
    assert st["write_pass_tests"] > 0 and st["executed_tests"] > 0
    assert st["executed_tests"] + st["prefilter_tests"] < st["nominal_tests"]
    _, _, total_b, sb = run_cull(hv, grid, ubo, lights, mode=MODES["brute"])
    assert total_b == total and sb["bin_tests"] == 0 and sb["write_pass_tests"] == 0 and sb["prefilter_tests"] == 0


@pytest.mark.parametrize("name,n,cap,z", [("A", 70_000, 150, None), ("A", 3_000, 4, None), ("B", 300_000, 150, None), ("C", 150_000, 150, (40, 52)),
                                         ("B", 0, 150, None), ("A", 65_536, 1, None), ("A", 1_025, 150, None), ("B", 40_000, 7, None)])
@pytest.mark.parametrize("plan", ["8:1:2", "2:3:2", "3:16:4", "8:64:4", "8:1:2:2", "4:2:2:64", "3:1:4:8"])
def test_rounds_give_identical_bytes(hv, oracle, monkeypatch, name, n, cap, z, plan):
    """Frames of many tiles are binned and counted in rounds; between rounds hier_alive_kernel marks the cluster groups whose clusters
    all hold their cap already, and the next round's binning skips them.  Forced here on small frames with several round plans
    (rounds : head tiles : growth [: tiles per macro tile of the rounds after the first two, with the two-level binning kernel]):
    the bytes equal the single-launch frame's and brute force's, and no round does more level-0 work."""
    grid, screen = GRIDS[name]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=17)
    kw = {} if z is None else {"z_begin": z[0], "z_end": z[1]}
    monkeypatch.setenv("HVCF_ROUNDS", "0")
    monkeypatch.setenv("HVCF_HIER_PARTS", "1")
    one = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["hier"], **kw)
    rounds, head, growth, macro = (plan.split(":") + [None])[:4]
    if macro:
        monkeypatch.setenv("HVCF_ROUNDS_MACRO", macro)
        monkeypatch.setenv("HVCF_BIN2", "1")
    monkeypatch.setenv("HVCF_ROUNDS", rounds)
    monkeypatch.setenv("HVCF_ROUNDS_HEAD", head)
    monkeypatch.setenv("HVCF_ROUNDS_GROWTH", growth)
    got = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["hier"], **kw)
    ref = run_cull(hv, grid, ubo, lights, cap=cap, mode=MODES["brute"], **kw)
    for out in (one, got):
        assert out[2] == ref[2]
        assert out[0].tobytes() == ref[0].tobytes() and out[1].tobytes() == ref[1].tobytes()
    tiles = (n + 1023) // 1024
    assert one[3]["plan_rounds"] == 1 and one[3]["device_bytes"] > one[3]["scratch_bytes"] > 0
    if tiles >= 2 * int(head):
        assert got[3]["plan_rounds"] >= 2 and got[3]["plan_segments"] >= got[3]["plan_rounds"]
        assert got[3]["kernel_launches"] > one[3]["kernel_launches"]          # the rounds really ran
        assert macro or got[3]["bin_tests"] <= one[3]["bin_tests"]


def test_download_grid_in_one_call(hv, oracle):
    grid, screen = GRIDS["A"]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(5000)
    with hv.Context(grid=grid, max_lights=5000) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo)
        ctx.cull(0)
        cells, idx, total = ctx.download_grid(0)
        assert total == ctx.download_total(0)
        assert cells.tobytes() == ctx.download_light_cells().tobytes()
        assert idx.tobytes() == ctx.download_light_indices(total).tobytes()
        import ctypes
        small = np.zeros(4, np.uint32)
        t = ctypes.c_uint32()
        rc = ctx.lib.hvcf_download_grid(ctx.handle, 0, cells.ctypes.data, cells.nbytes, small.ctypes.data, small.nbytes, ctypes.byref(t), None)
        assert rc == -4 and t.value == total          # HVCF_ERR_CAPACITY, the total is still reported


@pytest.mark.parametrize("n", [0, 1, 5000, 70_000])
def test_download_grid_into_pinned_memory_uses_one_kernel_and_gives_the_same_bytes(hv, oracle, monkeypatch, n):
    """pinned destinations: count, cells and indices[0, count) are stored across PCIe by grid_download_kernel (no host round
    trip for the count); pageable destinations and HVCF_DOWNLOAD_KERNEL=0 take the two-copy path.  Same bytes either way."""
    import ctypes
    grid, screen = GRIDS["A"]
    C = grid[0] * grid[1] * grid[2]
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=29)
    got = {}
    for sw in ("1:4", "1:1", "1:0", "0:4"):   # kernel on/off : MB of indices the kernel may store (the copy engine takes the rest)
        monkeypatch.setenv("HVCF_DOWNLOAD_KERNEL", sw[0])
        monkeypatch.setenv("HVCF_DOWNLOAD_KERNEL_MAX_MB", sw[2])
        with hv.Context(grid=grid, max_lights=max(n, 1)) as ctx:
            ctx.set_lights(lights)
            ctx.set_ubo(0, ubo)
            ctx.cull(0)
            want_cells, want_idx, want_total = ctx.download_grid(0)     # pageable numpy destinations
            ch, ih = hv.PinnedBuffer(8 * C), hv.PinnedBuffer(4 * C * 150)
            ih.array[:] = 0xAB
            t = ctypes.c_uint32(0xFFFFFFFF)
            rc = ctx.lib.hvcf_download_grid(ctx.handle, 0, ch.ptr, 8 * C, ih.ptr, 4 * C * 150, ctypes.byref(t), None)
            assert rc == 0 and t.value == want_total
            assert ch.array.tobytes() == want_cells.tobytes()
            assert ih.array[: 4 * want_total].tobytes() == want_idx.tobytes()
            assert (ih.array[4 * want_total: 4 * want_total + 64] == 0xAB).all()     # nothing past the count is touched
            if want_total > 8:
                small = hv.PinnedBuffer(32)
                small.array[:] = 0xCD
                rc = ctx.lib.hvcf_download_grid(ctx.handle, 0, ch.ptr, 8 * C, small.ptr, 32, ctypes.byref(t), None)
                assert rc == -4 and t.value == want_total and (small.array == 0xCD).all()
                small.close()
            got[sw] = (ch.array.tobytes(), ih.array[: 4 * want_total].tobytes(), want_total)
            ch.close(), ih.close()
    assert got["0:4"] == got["1:4"] == got["1:1"] == got["1:0"]
    assert n < 70_000 or got["1:1"][2] > (1 << 18)   # the 1 MB case really splits the list


@pytest.mark.parametrize("mode", ["small"])
@pytest.mark.parametrize("cap", [1, 150, 4096])
@pytest.mark.parametrize("grid,screen,n", [((16, 9, 24), (2560, 1600), 1000), ((16, 9, 24), (2560, 1600), 2048), ((1, 1, 1), (640, 480), 300), ((33, 2, 2), (1280, 720), 777),
                                           ((5, 7, 3), (1280, 720), 1), ((32, 18, 48), (3840, 2160), 1500), ((64, 36, 96), (3840, 2160), 2048)])
def test_small_frames_run_as_one_kernel(hv, oracle, grid, screen, n, cap, mode):
    """csrc/hvcf_small.cu: transform + test + capped counts + scan over the clusters + index lists in ONE kernel for frames of at
    most 2,048 lights (the reference's own scene is 1,000 lights in one dispatch, PipelineLightCulling.cpp:74-77).  Four frames in a
    row on both frame slots with a moving camera and a moved light (the look-back serial lives on the device), a scene with NaN /
    Inf / zero-radius lights, bytes against the oracle every time."""
    ubos = [oracle.camera_ubo(grid, screen), oracle.camera_ubo(grid, screen, pos=(1.5, 2.0, 5.0), target=(0, 2, -1))]
    lights = oracle.init_lights(n, seed=5)
    if n >= 300:
        lights["position"][7] = (np.nan, 1, 1, 1)
        lights["position"][8] = (np.inf, 0, 0, 1)
        lights["radius"][9] = 0.0
        lights["radius"][10] = np.inf
    aabb = oracle.generate_aabbs(ubos[0])
    with hv.Context(grid=grid, max_lights=n, max_lights_per_cluster=cap, mode=MODES[mode]) as ctx:
        ctx.set_lights(lights)
        for f in range(4):
            if f == 2:
                lights["position"][0] = (0.25, 2.0, 1.0, 1.0)
                ctx.update_light(0, (0.25, 2.0, 1.0))
            ctx.set_ubo(f & 1, ubos[f & 1])
            ctx.cull(f & 1)
            cells, idx, total = ctx.download_grid(f & 1)
            st = ctx.stats(f & 1)
            oc, oi, ototal, _ = oracle.cull(ubos[f & 1], aabb, lights, cap=cap)
            assert total == ototal and st["total_indices"] == ototal, f
            assert cells.tobytes() == oc.tobytes() and idx.tobytes() == oi.tobytes(), f
            assert st["plan_flags"] == 64 and st["kernel_launches"] == (2 if f == 0 else 1)
            assert st["overflow_clusters"] == int((oc["count"] == cap).sum())
            check_invariants(cells, idx, total, n, cap)
