"""GPU parity of the light upload / animation path: batched UpdateLightPosition and the device-side
animation fused into the per-frame light transform, against the oracle run on host-moved lights.  Bit-exact."""
import numpy as np
import pytest

from cpu_checkers import LIGHT_DT, same_light_sets

pytestmark = pytest.mark.gpu
MODES = {"brute": 1, "hier": 2}


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


@pytest.mark.parametrize("mode", list(MODES))
def test_update_lights_batch_matches_sequential_updates(hv, oracle, mode):
    grid, screen, n = (16, 9, 24), (2560, 1600), 20000
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    base = oracle.init_lights(n)
    rng = np.random.default_rng(5)
    with hv.Context(grid=grid, max_lights=n, mode=MODES[mode]) as ctx:
        ctx.set_lights(base)
        ctx.set_ubo(0, ubo)
        want = base.copy()
        for batch in (1, 17, 5000, 3000):  # growing and shrinking batches reuse / regrow the pinned staging
            idx = rng.choice(n + 50, size=batch, replace=False).astype(np.uint32)  # a few out of range: ignored
            xyz = rng.uniform(-8, 8, size=(batch, 3)).astype(np.float32)
            ctx.update_lights(idx, xyz)
            ok = idx < n
            want["position"][idx[ok], :3] = xyz[ok]
            want["position"][idx[ok], 3] = 1.0
        ctx.cull(0)
        got_lights = ctx.download_lights(n)
        assert got_lights.tobytes() == want.tobytes()
        cells, total = ctx.download_light_cells(), ctx.download_total(0)
        idxs = ctx.download_light_indices(total)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, want)
    assert total == ototal and same_light_sets(cells, idxs, oc, oi)


@pytest.mark.parametrize("mode", list(MODES))
def test_device_animation_matches_host_animation(hv, oracle, mode):
    grid, screen, n = (16, 9, 24), (2560, 1600), 30011  # not a multiple of the 1,024-light tile
    aabb = None
    base = oracle.init_lights(n)
    table = hv.fly_table()
    with hv.Context(grid=grid, max_lights=n, mode=MODES[mode]) as ctx:
        ctx.set_lights(base)
        ctx.set_light_animation(table, hv.FLY_AMPLITUDE)
        for k, f in enumerate((0, 1, 7, 300, 1023)):
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
            idxs = ctx.download_light_indices(total)
            oc, oi, ototal, _ = oracle.cull(ubo, aabb, want)
            assert total == ototal and same_light_sets(cells, idxs, oc, oi), f"frame {f}"
        # a position update moves the rest position of that light; the animation continues around it
        ctx.update_lights(np.array([5], np.uint32), np.array([[1.0, 2.0, 3.0]], np.float32))
        base2 = base.copy()
        base2["position"][5] = (1.0, 2.0, 3.0, 1.0)
        ctx.set_animation_step(0, hv.fly_step(11))
        ctx.cull(0)
        assert ctx.download_lights(n).tobytes() == hv.fly_lights_host(base2, 11, table).tobytes()
        # off: positions stay where the last animated frame left them
        ctx.set_light_animation(None, 0.0)
        ctx.cull(0)
        assert ctx.download_lights(n).tobytes() == hv.fly_lights_host(base2, 11, table).tobytes()


def test_animation_argument_errors(hv):
    with hv.Context(grid=(16, 9, 24), max_lights=64) as ctx:
        with pytest.raises(hv.HvcfError):
            ctx.set_light_animation(np.zeros(1000, np.float32), 0.5)  # not a power of two
        with pytest.raises(hv.HvcfError):
            ctx.set_animation_step(9, 0)
        ctx.update_lights(np.zeros(0, np.uint32), np.zeros((0, 3), np.float32))  # empty batch is a no-op
