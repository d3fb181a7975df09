"""GPU parity at the FULL sizes of BASELINE.json's configs, against the CPU oracle (not against brute force):
  A1M   north_star's target shape (16x9x24 x 2^20 lights): every cluster, every list.
  B1M   configs[2] (32x18x48 x 2^20): every cluster, every list, default (hierarchical) mode.
  C16M  configs[3] (64x36x96 x 2^24): >= 2,048 seeded clusters spread over all 96 z-slices, each checked in full.
  fly   configs[4] at B1M: frames 0, 300, 1023 of the orbit with animated lights, 256 sampled clusters each.
The oracle stops scanning a cluster at its 150th hit exactly as LightCulling.comp:71 does, so the full B1M frame is
~8.5e9 oracle tests (seconds on the GPU box's host cores)."""
import numpy as np
import pytest

from cpu_checkers import same_light_sets

pytestmark = pytest.mark.gpu

GRIDS = {"A": ((16, 9, 24), (2560, 1600)), "B": ((32, 18, 48), (3840, 2160)), "C": ((64, 36, 96), (3840, 2160))}


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


def gpu_frame(hv, grid, ubo, lights, frame=0):
    with hv.Context(grid=grid, max_lights=len(lights)) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(frame, ubo)
        ctx.cull(frame)
        cells = ctx.download_light_cells()
        total = ctx.download_total(frame)
        idx = ctx.download_light_indices(total)
    return cells, idx, total


def check_sample(oracle, ubo, aabb, lights, cells, idx, sample):
    """oracle.cull on the gathered boxes of `sample` (all host threads), list by list against the GPU's."""
    sub = np.ascontiguousarray(aabb[sample])
    oc, oi, _, _ = oracle.cull(ubo, sub, lights)
    for k, c in enumerate(sample):
        want = oi[int(oc["offset"][k]): int(oc["offset"][k]) + int(oc["count"][k])]
        got = idx[int(cells["offset"][c]): int(cells["offset"][c]) + int(cells["count"][c])]
        assert np.array_equal(got, want), f"cluster {c}: {len(got)} vs {len(want)} lights"


@pytest.mark.parametrize("name", ["A", "B"])
def test_one_million_lights_every_cluster_against_the_oracle(hv, oracle, name):
    grid, screen = GRIDS[name]
    n = 1 << 20
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(n)
    cells, idx, total = gpu_frame(hv, grid, ubo, lights)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert total == ototal
    assert np.array_equal(cells["count"], oc["count"])
    assert np.array_equal(cells["offset"], oc["offset"])   # both sides: exclusive scan in cluster order
    assert np.array_equal(idx, oi)
    assert same_light_sets(cells, idx, oc, oi)


def test_sixteen_million_lights_2048_clusters_over_all_slices(hv, oracle):
    grid, screen = GRIDS["C"]
    n = 1 << 24
    X, Y, Z = grid
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(n)
    cells, idx, total = gpu_frame(hv, grid, ubo, lights)
    cnt = cells["count"].astype(np.int64)
    assert int(cnt.sum()) == total and int(cnt.max()) <= 150
    assert np.array_equal(cells["offset"].astype(np.int64), np.concatenate([[0], np.cumsum(cnt)[:-1]]))
    rng = np.random.default_rng(2024)
    per = -(-2048 // Z)                                   # 22 clusters of every slice: 2,112 in all
    sample = np.concatenate([z * X * Y + rng.choice(X * Y, size=per, replace=False) for z in range(Z)])
    assert len(sample) >= 2048 and len(np.unique(sample // (X * Y))) == Z
    check_sample(oracle, ubo, aabb, lights, cells, idx, sample)


def test_flythrough_frames_at_full_size(hv, oracle):
    grid, screen = GRIDS["B"]
    n = 1 << 20
    base = oracle.init_lights(n)
    table = hv.fly_table()
    rng = np.random.default_rng(77)
    aabb = None
    with hv.Context(grid=grid, max_lights=n) as ctx:
        ctx.set_lights(base)
        ctx.set_light_animation(table, hv.FLY_AMPLITUDE)
        for k, f in enumerate((0, 300, 1023)):
            frame = k & 1
            ubo = oracle.camera_ubo(grid, screen, pos=hv.fly_camera_pos(f))
            if aabb is None:
                aabb = oracle.generate_aabbs(ubo)
            ctx.set_ubo(frame, ubo)
            ctx.set_animation_step(frame, hv.fly_step(f))
            ctx.cull(frame)
            want = hv.fly_lights_host(base, f, table)
            assert ctx.download_lights(n).tobytes() == want.tobytes(), f"animated light buffer differs at frame {f}"
            cells, total = ctx.download_light_cells(), ctx.download_total(frame)
            idx = ctx.download_light_indices(total)
            sample = rng.choice(len(aabb), size=256, replace=False)
            check_sample(oracle, ubo, aabb, want, cells, idx, sample)
