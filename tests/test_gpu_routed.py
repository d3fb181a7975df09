"""GPU: the sharded frame with the LIGHTS sharded too (hvcf_cull_routed; csrc/hvcf_shard.cu, hvcf_shard_api.cu).

Rank r owns a z-slab of the cluster grid and a contiguous range of the lights; per frame it transforms its own lights,
tests them against every rank's slab box and stores the survivors straight into the owning rank's inbox, in light
order; the owner bins its inbox only.  As in test_gpu_peer_exchange.py the shard contexts live in THIS process on
cuda:0 (hvcf_peer_attach_local), each on its own stream, so the wait kernels of one shard overlap the work of the
others as they do across GPUs (across processes the same kernels run over CUDA IPC mappings: bench.py under torchrun
checks those bytes).  Afterwards EVERY context must hold the full light grid, byte-identical to the single-context
frame over all the lights and equal to the oracle's lists."""
import numpy as np
import pytest

from cpu_checkers import same_light_sets

pytestmark = pytest.mark.gpu

GRID_A, SCREEN_A = (16, 9, 24), (2560, 1600)
GRID_B, SCREEN_B = (32, 18, 48), (3840, 2160)
INBOX_SEEN = {}


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    return torch


def run_single(hv, grid, ubo, lights, cap=150):
    with hv.Context(grid=grid, max_lights=max(len(lights), 1), max_lights_per_cluster=cap) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo)
        ctx.cull(0)
        return ctx.download_grid(0)


def run_routed(hv, torch, grid, ubos, lights, cuts, frames=1, cap=150, block_tiles=0, check=None):
    """one context per rank; returns [(cells, idx, total)] per rank after the last frame; `check(frame, results)` per frame"""
    world, n = len(cuts) - 1, len(lights)
    ctxs = [hv.Context(grid=grid, max_lights=max(n, 1), max_lights_per_cluster=cap, shard=(r, world, n, cuts, block_tiles))
            for r in range(world)]
    streams = [torch.cuda.Stream() for _ in range(world)]
    try:
        covered = 0
        for c in ctxs:
            b, e = c.shard_light_range()
            assert b == min(covered, n) and b <= e <= n
            covered = max(covered, e)
            c.set_lights_range(lights[b:e], b)       # every rank holds ONLY its own lights
            for f in range(2):
                c.set_ubo(f, ubos[f % len(ubos)])
        assert covered == n
        torch.cuda.synchronize()
        if world > 1:
            for r, c in enumerate(ctxs):
                c.peer_attach_local(r, ctxs)
        out = None
        for f in range(frames):
            for c, s in zip(ctxs, streams):
                c.cull_routed(f & 1, stream=s.cuda_stream)
            for c, s in zip(ctxs, streams):
                assert c.peer_status(stream=s.cuda_stream) == 0
            torch.cuda.synchronize()
            out = [c.download_grid(f & 1) for c in ctxs]
            if check:
                check(f, out)
        inbox = [c.shard_inbox_records() for c in ctxs]
        if world > 1:
            for c in ctxs:
                c.peer_detach()
        return out, inbox
    finally:
        for c in ctxs:
            c.close()


def assert_all_equal(out, want, what=""):
    want_cells, want_idx, want_total = want
    for r, (cells, idx, total) in enumerate(out):
        assert total == want_total, f"{what} rank {r}: total {total} != {want_total}"
        assert cells.tobytes() == want_cells.tobytes(), f"{what} rank {r}: light cells"
        assert idx.tobytes() == want_idx.tobytes(), f"{what} rank {r}: light indices"


@pytest.mark.parametrize("n", [0, 1, 1000, 1025, 9000, 70_000])
@pytest.mark.parametrize("cuts", [(0, 24), (0, 12, 24), (0, 15, 19, 24), (0, 3, 6, 9, 12, 15, 18, 21, 24)])
def test_routed_equals_single_context_and_oracle(hv, torch_mod, oracle, n, cuts):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(n)
    want = run_single(hv, GRID_A, ubo, lights)
    out, inbox = run_routed(hv, torch_mod, GRID_A, [ubo], lights, cuts, frames=2)
    assert_all_equal(out, want)
    aabb = oracle.generate_aabbs(ubo)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert want[2] == ototal and same_light_sets(want[0], want[1], oc, oi)
    assert all(r <= n for r in inbox)   # a light reaches a rank at most once


@pytest.mark.parametrize("block_tiles", [1, 2, 4, 16])
def test_route_block_sizes(hv, torch_mod, oracle, block_tiles):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(40_000, seed=3)
    want = run_single(hv, GRID_A, ubo, lights)
    out, _ = run_routed(hv, torch_mod, GRID_A, [ubo], lights, (0, 14, 18, 24), block_tiles=block_tiles)
    assert_all_equal(out, want, f"B={block_tiles}")


@pytest.mark.parametrize("cap", [1, 7, 4096])
def test_routed_cap_values(hv, torch_mod, oracle, cap):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(20_000, seed=9)
    want = run_single(hv, GRID_A, ubo, lights, cap=cap)
    out, _ = run_routed(hv, torch_mod, GRID_A, [ubo], lights, (0, 13, 17, 24), cap=cap)
    assert_all_equal(out, want, f"cap={cap}")


def test_routed_frames_alternate_cameras_and_moved_lights(hv, torch_mod, oracle):
    """four frames over the two frame slots with two cameras: the inbox is double-buffered on the frame parity and nothing
    of an earlier frame may leak into a later one"""
    ubos = [oracle.camera_ubo(GRID_A, SCREEN_A), oracle.camera_ubo(GRID_A, SCREEN_A, pos=(0.5, 1.0, 6.0))]
    lights = oracle.init_lights(30_011)
    wants = [run_single(hv, GRID_A, u, lights) for u in ubos]

    def check(f, out):
        assert_all_equal(out, wants[f & 1], f"frame {f}")
    run_routed(hv, torch_mod, GRID_A, ubos, lights, (0, 12, 16, 20, 24), frames=4, check=check)


def test_routed_dense_slab_overflows_slots(hv, torch_mod, oracle):
    """lights packed into a small volume: (group, tile) slots overflow kBinCap, so the count / write passes read the raw inbox tile"""
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(50_000, seed=11)
    rng = np.random.default_rng(5)
    lights["position"][:, 0] = rng.uniform(-0.5, 0.5, len(lights)).astype(np.float32)
    lights["position"][:, 1] = rng.uniform(2.0, 3.0, len(lights)).astype(np.float32)
    lights["position"][:, 2] = rng.uniform(-0.5, 0.5, len(lights)).astype(np.float32)
    want = run_single(hv, GRID_A, ubo, lights)
    out, _ = run_routed(hv, torch_mod, GRID_A, [ubo], lights, (0, 16, 20, 24))
    assert_all_equal(out, want)
    aabb = oracle.generate_aabbs(ubo)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert want[2] == ototal and same_light_sets(want[0], want[1], oc, oi)


def test_routed_nan_and_degenerate_lights(hv, torch_mod, oracle):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(5_000, seed=2)
    lights["radius"][::7] = 0.0
    lights["radius"][5::11] = np.inf
    lights["position"][3::13, 0] = np.nan
    lights["radius"][9::17] = np.nan
    want = run_single(hv, GRID_A, ubo, lights)
    out, _ = run_routed(hv, torch_mod, GRID_A, [ubo], lights, (0, 10, 18, 24))
    assert_all_equal(out, want)
    aabb = oracle.generate_aabbs(ubo)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert want[2] == ototal and same_light_sets(want[0], want[1], oc, oi)


def test_routed_grid_b_one_million_lights(hv, torch_mod, oracle):
    """configs[2]'s shape in 4 unequal slabs: routed bytes = single-context bytes (which test_gpu_full_size checks against the oracle)"""
    ubo = oracle.camera_ubo(GRID_B, SCREEN_B)
    lights = oracle.init_lights(1 << 20)
    want = run_single(hv, GRID_B, ubo, lights)
    out, inbox = run_routed(hv, torch_mod, GRID_B, [ubo], lights, (0, 26, 33, 38, 48))
    assert_all_equal(out, want)
    assert sum(inbox) < 4 * len(lights)   # the slab filter is selective


def test_routed_argument_errors(hv, oracle):
    with pytest.raises(hv.HvcfError):
        hv.Context(grid=GRID_A, max_lights=100, shard=(1, 2, 100, (0, 24, 24)))      # rank 1 owns no slice
    with pytest.raises(hv.HvcfError):
        hv.Context(grid=GRID_A, max_lights=100, shard=(0, 2, 100, (0, 12, 23)))      # cuts do not reach Z
    with pytest.raises(hv.HvcfError):
        hv.Context(grid=GRID_A, max_lights=100, shard=(0, 2, 100, (0, 12, 24), 3))   # block_tiles not a power of two
    with hv.Context(grid=GRID_A, max_lights=100, shard=(0, 2, 100, (0, 12, 24))) as c:
        c.set_ubo(0, oracle.camera_ubo(GRID_A, SCREEN_A))
        with pytest.raises(hv.HvcfError):
            c.cull_routed(0)                                                             # world 2, nobody attached
    with hv.Context(grid=GRID_A, max_lights=100) as c:
        with pytest.raises(hv.HvcfError):
            c.cull_routed(0)
        with pytest.raises(hv.HvcfError):
            c.shard_light_range()


# ---- row-granular shards: cuts inside a z-slice (row = y + z * Y), for load balance finer than whole slices ----
@pytest.mark.parametrize("cuts", [(0, 216), (0, 100, 216), (0, 37, 101, 130, 216), (0, 8, 16, 24, 32, 40, 48, 150, 216), (0, 1, 2, 216)])
@pytest.mark.parametrize("n", [1000, 70_000])
def test_routed_row_cuts(hv, torch_mod, oracle, cuts, n):
    """grid A has 9 x 24 = 216 cluster rows; cuts that are not multiples of the block height (4) leave partly masked blocks"""
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(n, seed=21)
    want = run_single(hv, GRID_A, ubo, lights)
    out, _ = run_routed(hv, torch_mod, GRID_A, [ubo], lights, cuts, frames=2)
    assert_all_equal(out, want, f"rows {cuts}")


def test_row_shards_of_plain_contexts_concatenate(hv, oracle):
    """hvcf_config.row_begin / row_end on plain contexts (no peers): each writes its own cells and its own index segment at
    shard-local offsets; concatenated in row order they are the whole grid"""
    ubo = oracle.camera_ubo(GRID_B, SCREEN_B)
    lights = oracle.init_lights(120_000, seed=4)
    want_cells, want_idx, want_total = run_single(hv, GRID_B, ubo, lights)
    X, Y, Z = GRID_B
    cuts = [0, 5, 300, 301, 555, Y * Z]
    got_counts, got_idx = [], []
    for a, b in zip(cuts[:-1], cuts[1:]):
        with hv.Context(grid=GRID_B, max_lights=len(lights), rows=(a, b)) as ctx:
            ctx.set_lights(lights)
            ctx.set_ubo(0, ubo)
            ctx.cull(0)
            cells = ctx.download_light_cells()[a * X: b * X]
            total = ctx.download_total(0)
            idx = ctx.download_light_indices(total)
            assert int(cells["count"].sum()) == total
            assert np.array_equal(cells["offset"].astype(np.int64), np.concatenate([[0], np.cumsum(cells["count"].astype(np.int64))[:-1]]))
            got_counts.append(cells["count"].copy())
            got_idx.append(idx.copy())
    assert np.array_equal(np.concatenate(got_counts), want_cells["count"])
    assert np.array_equal(np.concatenate(got_idx), want_idx)


@pytest.mark.parametrize("two_phase", ["1", "0"])
@pytest.mark.parametrize("plan", ["8:1:2", "3:4:4", "4:16:8", "5:8:2", "4:4:4:m0", "3:8:2:b8", "4:16:4:b16", "3:2:2:b1"])
@pytest.mark.parametrize("grid,screen,n,cuts,cap", [(GRID_A, SCREEN_A, 70_000, (0, 15, 19, 24), 150), (GRID_B, SCREEN_B, 120_000, (0, 30, 40, 48), 150),
                                                   (GRID_A, SCREEN_A, 30_000, (0, 12, 24), 5)])
def test_routed_frames_in_rounds(hv, torch_mod, oracle, monkeypatch, plan, grid, screen, n, cuts, cap, two_phase):
    """the inbox of a routed frame is in light order, so it takes rounds like any other frame (hier_alive_kernel between the rounds'
    binning launches); forced here on small frames: rounds : head tiles : growth.  With three rounds or more (and the third one
    starting at a whole route block) the frame is routed in two phases: the lights of the first two rounds against the slabs' union
    boxes, the rest against the boxes of the clusters still below their cap, which every rank publishes in between
    (slab_alive_kernel).  Every rank's full grid equals the single context's, two frames in a row (both mailbox parities)."""
    monkeypatch.setenv("HVCF_ROUTE_TWO_PHASE", two_phase)
    ubo = oracle.camera_ubo(grid, screen)
    lights = oracle.init_lights(n, seed=21)
    monkeypatch.setenv("HVCF_ROUNDS", "0")
    want = run_single(hv, grid, ubo, lights, cap=cap)
    rounds, head, growth, opt = (plan.split(":") + [""])[:4]
    monkeypatch.setenv("HVCF_ROUNDS", rounds)
    monkeypatch.setenv("HVCF_ROUNDS_HEAD", head)
    monkeypatch.setenv("HVCF_ROUNDS_GROWTH", growth)
    if opt == "m0":     # one slot per inbox tile in every round (otherwise the late rounds keep one slot per route block)
        monkeypatch.setenv("HVCF_ROUNDS_MACRO", "0")
    block_tiles = int(opt[1:]) if opt.startswith("b") else 0
    out, inbox = run_routed(hv, torch_mod, grid, [ubo], lights, cuts, frames=3, cap=cap, block_tiles=block_tiles)
    assert_all_equal(out, want, f"rounds {plan} two_phase {two_phase}")
    INBOX_SEEN.setdefault((plan, grid, n, cuts, cap), {})[two_phase] = sum(inbox)
    seen = INBOX_SEEN[(plan, grid, n, cuts, cap)]
    if len(seen) == 2:
        assert seen["1"] <= seen["0"]   # live boxes never route more than the union boxes


@pytest.mark.parametrize("args", [("16", "9", "24", "2560", "1600", "30000", "12345", "3", "3"),
                                  ("32", "18", "48", "3840", "2160", "1500", "7", "2", "2"),      # unsharded side: the single-kernel frame
                                  ("16", "9", "24", "2560", "1600", "5000", "3", "1", "2")])      # world 1: no attach
def test_routed_frame_driven_from_cpp_alone(args):
    """hellovulkan_b200/lib/hv_routed_shards_app (host/app_routed_shards.cpp): `world` shard contexts in one C++ process, nothing but
    the C ABI and one CUDA stream per shard; it compares every shard's full light grid with the unsharded hvcf_cull frame by frame
    (camera orbiting, light 0 moving) and prints OK."""
    import os
    import subprocess
    app = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "hellovulkan_b200", "lib", "hv_routed_shards_app")
    assert os.path.exists(app), "build() did not produce the routed-shards driver"
    env = {k: v for k, v in os.environ.items() if not k.startswith("HVCF_")}    # the library's defaults, as a C++ caller gets them
    r = subprocess.run([app, *args], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and r.stdout.startswith("OK "), r.stdout + r.stderr


def test_routed_frames_read_bound_light_buffers(hv, torch_mod, oracle):
    """hvcf_bind_lights on contexts of hvcf_create_sharded: every rank binds a device buffer holding ONLY the lights it owns
    (two buffers in turn, with different scenes, as a host does that uploads the next frame's lights while this one is culled);
    every rank's full grid equals the single context's frame over the scene of that frame."""
    n, cuts = 50_000, (0, 13, 18, 24)
    world = len(cuts) - 1
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    scenes = [oracle.init_lights(n, seed=4), oracle.init_lights(n, seed=8)]
    wants = [run_single(hv, GRID_A, ubo, sc) for sc in scenes]
    assert wants[0][1].tobytes() != wants[1][1].tobytes()
    ctxs = [hv.Context(grid=GRID_A, max_lights=n, shard=(r, world, n, cuts, 0)) for r in range(world)]
    streams = [torch_mod.cuda.Stream() for _ in range(world)]
    bufs = []
    try:
        for c in ctxs:
            b, e = c.shard_light_range()
            bufs.append([hv.DeviceBuffer.from_array(sc[b:e]) if e > b else None for sc in scenes])
            for f in range(2):
                c.set_ubo(f, ubo)
            with pytest.raises(hv.HvcfError):
                c.bind_lights(bufs[-1][0].ptr.value, e - b + 1)       # exactly the owned lights
        torch_mod.cuda.synchronize()
        for r, c in enumerate(ctxs):
            c.peer_attach_local(r, ctxs)
        for f in range(5):
            which = (f // 2) & 1                                       # frame slots and buffers alternate at different rates
            for r, (c, s) in enumerate(zip(ctxs, streams)):
                b, e = c.shard_light_range()
                if e > b:
                    c.bind_lights(bufs[r][which].ptr.value, e - b)
                c.cull_routed(f & 1, stream=s.cuda_stream)
            for c, s in zip(ctxs, streams):
                assert c.peer_status(stream=s.cuda_stream) == 0
            torch_mod.cuda.synchronize()
            assert_all_equal([c.download_grid(f & 1) for c in ctxs], wants[which], f"frame {f}")
        with pytest.raises(hv.HvcfError):
            ctxs[0].cull(0)                                            # only the routed frame reads an owned-lights buffer
        with pytest.raises(hv.HvcfError):
            ctxs[0].update_light(0, (0.0, 0.0, 0.0))
        for c in ctxs:
            c.bind_lights(None, 0)
            c.peer_detach()
    finally:
        for c in ctxs:
            c.close()
        for pair in bufs:
            for b in pair:
                if b is not None:
                    b.close()
