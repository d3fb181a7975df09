"""GPU: the exchange step of the z-slab sharded grid done by kernels over peer memory (csrc/hvcf_peer.cu).

Several shard contexts live in THIS process on cuda:0 (hvcf_peer_attach_local maps them onto each other with plain
pointers; across processes the same kernels run over CUDA IPC mappings, exercised by bench.py under torchrun).
Each context runs hvcf_cull_sharded on its own stream, so the wait kernels of one shard overlap the work of the
others exactly as they do across GPUs.  Afterwards EVERY context must hold the full light grid, byte-identical to
the single-context result (offsets are the exclusive scan in cluster order, so the bytes are comparable) and to
the oracle."""
import numpy as np
import pytest

from cpu_checkers import same_light_sets

pytestmark = pytest.mark.gpu

GRID_A, SCREEN_A = (16, 9, 24), (2560, 1600)


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


@pytest.fixture(scope="module")
def torch_mod():
    import torch
    return torch


def run_sharded(hv, torch, grid, ubo, lights, slabs, frames=1, cap=150, split_calls=False):
    world = len(slabs)
    ctxs = [hv.Context(grid=grid, max_lights=max(len(lights), 1), max_lights_per_cluster=cap, z_begin=a, z_end=b) for a, b in slabs]
    streams = [torch.cuda.Stream() for _ in slabs]
    try:
        for c in ctxs:
            c.set_lights(lights)
            for f in range(2):
                c.set_ubo(f, ubo)
        torch.cuda.synchronize()
        for r, c in enumerate(ctxs):
            c.peer_attach_local(r, ctxs)
        for f in range(frames):
            for c, s in zip(ctxs, streams):
                sp = s.cuda_stream
                if split_calls:
                    c.cull_begin(f & 1, stream=sp)
                    c.peer_exchange_totals(stream=sp)
                    c.cull_finish(f & 1, c.peer_index_base_ptr(), stream=sp)
                    c.peer_push_grid(stream=sp)
                    c.peer_wait_grid(stream=sp)
                else:
                    c.cull_sharded(f & 1, stream=sp)
        out = []
        for c, s in zip(ctxs, streams):
            assert c.peer_status(stream=s.cuda_stream) == 0
        torch.cuda.synchronize()
        for c in ctxs:
            cells = c.download_light_cells()
            total = int(cells["count"].astype(np.int64).sum())
            out.append((cells, c.download_light_indices(total), total))
        for c in ctxs:
            c.peer_detach()
        return out
    finally:
        for c in ctxs:
            c.close()


def run_single(hv, grid, ubo, lights, cap=150):
    with hv.Context(grid=grid, max_lights=max(len(lights), 1), max_lights_per_cluster=cap) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo)
        ctx.cull(0)
        cells = ctx.download_light_cells()
        total = ctx.download_total(0)
        return cells, ctx.download_light_indices(total), total


@pytest.mark.parametrize("slabs", [[(0, 24)], [(0, 12), (12, 24)], [(0, 5), (5, 6), (6, 24)], [(0, 1), (1, 2), (2, 3), (3, 24)],
                                   [(z, z + 3) for z in range(0, 24, 3)]])
@pytest.mark.parametrize("n", [1000, 65536])
def test_every_shard_holds_the_full_grid(hv, torch_mod, oracle, slabs, n):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(n)
    want_cells, want_idx, want_total = run_single(hv, GRID_A, ubo, lights)
    oc, oi, ototal, _ = oracle.cull(ubo, aabb, lights)
    assert want_total == ototal and same_light_sets(want_cells, want_idx, oc, oi)
    got = run_sharded(hv, torch_mod, GRID_A, ubo, lights, slabs)
    for r, (cells, idx, total) in enumerate(got):
        assert total == want_total, f"rank {r}"
        assert cells.tobytes() == want_cells.tobytes(), f"rank {r}: light cells differ from the unsharded run"
        assert idx.tobytes() == want_idx.tobytes(), f"rank {r}: light indices differ from the unsharded run"


def test_repeated_frames_and_split_calls(hv, torch_mod, oracle):
    """epochs advance in lockstep over many frames (double-buffered mailbox slots), also through the five separate calls"""
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(4096)
    want_cells, want_idx, want_total = run_single(hv, GRID_A, ubo, lights)
    for split in (False, True):
        got = run_sharded(hv, torch_mod, GRID_A, ubo, lights, [(0, 7), (7, 16), (16, 24)], frames=9, split_calls=split)
        for cells, idx, total in got:
            assert total == want_total
            assert cells.tobytes() == want_cells.tobytes() and idx.tobytes() == want_idx.tobytes()


def test_empty_scene_and_odd_segments(hv, torch_mod, oracle):
    """no lights at all (totals 0, nothing to push) and a cap that makes index segments start at odd word offsets"""
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    for n, cap in ((0, 150), (777, 7), (5000, 1)):
        lights = oracle.init_lights(n)
        want_cells, want_idx, want_total = run_single(hv, GRID_A, ubo, lights, cap=cap)
        got = run_sharded(hv, torch_mod, GRID_A, ubo, lights, [(0, 11), (11, 13), (13, 24)], cap=cap)
        for cells, idx, total in got:
            assert total == want_total
            assert cells.tobytes() == want_cells.tobytes() and idx.tobytes() == want_idx.tobytes()


def test_missing_peer_is_an_error_not_a_hang(hv, torch_mod, oracle):
    """a rank whose peer never shows up gets an error from hvcf_peer_status after the bounded wait"""
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(100)
    ctxs = [hv.Context(grid=GRID_A, max_lights=100, z_begin=a, z_end=b) for a, b in ((0, 12), (12, 24))]
    try:
        for c in ctxs:
            c.set_lights(lights)
            c.set_ubo(0, ubo)
        for r, c in enumerate(ctxs):
            c.peer_attach_local(r, ctxs)
        ctxs[0].cull_sharded(0)          # rank 1 never calls
        with pytest.raises(hv.HvcfError):
            ctxs[0].peer_status()
        for c in ctxs:
            c.peer_detach()
    finally:
        for c in ctxs:
            c.close()


def test_calls_without_peers_fail(hv):
    with hv.Context(grid=GRID_A, max_lights=16) as ctx:
        for call in (ctx.peer_exchange_totals, ctx.peer_push_grid, ctx.peer_wait_grid, lambda: ctx.cull_sharded(0)):
            with pytest.raises(hv.HvcfError):
                call()


# ---- hvcf_cull_begin in pieces (the light exchange pipelined with the first kernels of the frame) ----------------
def cull_in_ranges(hv, ctx, n, frame, chunks, flags=0):
    from hellovulkan_b200.sharded import ShardedCuller
    for b, e in ShardedCuller.chunk_bounds(n, chunks):
        ctx.cull_begin_range(frame, b, e, flags)


@pytest.mark.parametrize("chunks", [1, 3, 64])
@pytest.mark.parametrize("n", [0, 1, 1024, 5000, 65536])
def test_ranges_give_the_bytes_of_one_call(hv, oracle, n, chunks):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(n)
    want_cells, want_idx, want_total = run_single(hv, GRID_A, ubo, lights)
    with hv.Context(grid=GRID_A, max_lights=max(n, 1)) as ctx:
        ctx.set_lights(lights)
        for f in range(2):
            ctx.set_ubo(f, ubo)
        for f in (0, 1, 0):                      # several frames: the per-frame scratch is cleared by the first range
            cull_in_ranges(hv, ctx, n, f, chunks)
            ctx.cull_finish(f, None)
            cells = ctx.download_light_cells()
            total = ctx.download_total(f)
            assert total == want_total
            assert cells.tobytes() == want_cells.tobytes()
            assert ctx.download_light_indices(total).tobytes() == want_idx.tobytes()


def test_view_records_transformed_elsewhere(hv, torch_mod, oracle):
    """wire='view': one context transforms, the 16 B view records are copied, the other context only bins"""
    from hellovulkan_b200.sharded import as_tensor
    n = 20000
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(n)
    want_cells, want_idx, want_total = run_single(hv, GRID_A, ubo, lights)
    with hv.Context(grid=GRID_A, max_lights=n) as a, hv.Context(grid=GRID_A, max_lights=n) as b:
        a.set_lights(lights)
        b.set_lights(np.zeros(n, lights.dtype))     # b never sees the LightData; only the count matters
        for c in (a, b):
            c.set_ubo(0, ubo)
        dev = torch_mod.device("cuda", 0)
        va = as_tensor(*a.lights_view_ptr(), dev)
        vb = as_tensor(*b.lights_view_ptr(), dev)
        for frame_no in range(2):
            cull_in_ranges(hv, a, n, 0, 4, hv.RANGE_TRANSFORM)
            torch_mod.cuda.synchronize()
            vb.copy_(va)
            torch_mod.cuda.synchronize()
            cull_in_ranges(hv, b, n, 0, 5, hv.RANGE_BIN)
            b.cull_finish(0, None)
            total = b.download_total(0)
            assert total == want_total
            assert b.download_light_cells().tobytes() == want_cells.tobytes()
            assert b.download_light_indices(total).tobytes() == want_idx.tobytes()


def test_range_argument_errors(hv, oracle):
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(5000)
    with hv.Context(grid=GRID_A, max_lights=5000) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo)
        for args in ((0, 100, 1000), (0, 0, 1000), (0, 1024, 2048), (0, 0, 6000), (0, 2048, 1024), (0, 0, 2048, 8)):
            with pytest.raises(hv.HvcfError):
                ctx.cull_begin_range(*args)
        ctx.cull_begin_range(0, 0, 2048)
        with pytest.raises(hv.HvcfError):
            ctx.cull_begin_range(0, 3072, 5000)   # gap
        ctx.cull_begin_range(0, 2048, 5000)
        ctx.cull_finish(0, None)
    with hv.Context(grid=GRID_A, max_lights=5000, mode=1) as brute:
        brute.set_lights(lights)
        brute.set_ubo(0, ubo)
        with pytest.raises(hv.HvcfError):
            brute.cull_begin_range(0, 0, 5000)


def test_config3_full_size_sharded_equals_unsharded(hv, torch_mod, oracle):
    """BASELINE configs[3] at its full size, 64x36x96 clusters x 16,777,216 lights: invariants on everything, a seeded sample of
    clusters checked in full by the oracle, and the z-slab sharded result (4 unequal slabs, peer exchange, light buffer fed in
    chunks through hvcf_cull_begin_range) byte-identical to the unsharded one on every shard."""
    grid, screen, n = (64, 36, 96), (3840, 2160), 1 << 24
    ubo = oracle.camera_ubo(grid, screen)
    aabb = oracle.generate_aabbs(ubo)
    lights = oracle.init_lights(n)
    want_cells, want_idx, want_total = run_single(hv, grid, ubo, lights)
    cnt = want_cells["count"].astype(np.int64)
    assert np.all(cnt <= 150) and int(cnt.sum()) == want_total
    assert np.array_equal(want_cells["offset"].astype(np.int64), np.concatenate([[0], np.cumsum(cnt)[:-1]]))
    rng = np.random.default_rng(7)
    for c in rng.choice(len(aabb), size=12, replace=False):
        want, _ = oracle.cull_one(ubo, aabb[c:c + 1], lights)
        got = want_idx[want_cells["offset"][c]: want_cells["offset"][c] + want_cells["count"][c]]
        assert np.array_equal(got, want), f"cluster {c}"
    slabs = [(0, 40), (40, 61), (61, 70), (70, 96)]
    ctxs = [hv.Context(grid=grid, max_lights=n, z_begin=a, z_end=b) for a, b in slabs]
    streams = [torch_mod.cuda.Stream() for _ in slabs]
    try:
        for c in ctxs:
            c.set_lights(lights)
            c.set_ubo(0, ubo)
        torch_mod.cuda.synchronize()
        for r, c in enumerate(ctxs):
            c.peer_attach_local(r, ctxs)
        from hellovulkan_b200.sharded import ShardedCuller
        for c, s in zip(ctxs, streams):
            sp = s.cuda_stream
            for b, e in ShardedCuller.chunk_bounds(n, 4):
                c.cull_begin_range(0, b, e, 0, stream=sp)
            c.peer_exchange_totals(stream=sp)
            c.cull_finish(0, c.peer_index_base_ptr(), stream=sp)
            c.peer_push_grid(stream=sp)
            c.peer_wait_grid(stream=sp)
        for c, s in zip(ctxs, streams):
            assert c.peer_status(stream=s.cuda_stream) == 0
        for r, c in enumerate(ctxs):
            cells, idx, total = c.download_grid(0)
            assert total == want_total, f"shard {r}"
            assert cells.tobytes() == want_cells.tobytes(), f"shard {r}: light cells"
            assert idx.tobytes() == want_idx.tobytes(), f"shard {r}: light indices"
        for c in ctxs:
            c.peer_detach()
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("n", [0, 1, 1024, 5000, 70_000])
def test_transform_slices_in_any_order_then_bin(hv, oracle, n):
    """lights owned in shards (wire='gather'): the view records are produced slice by slice in any order by
    hvcf_transform_lights, then hvcf_cull_begin_range(BIN) alone completes the frame: the bytes of one hvcf_cull call"""
    ubo = oracle.camera_ubo(GRID_A, SCREEN_A)
    lights = oracle.init_lights(n)
    want_cells, want_idx, want_total = run_single(hv, GRID_A, ubo, lights)
    tiles = max(1, -(-n // 1024))
    with hv.Context(grid=GRID_A, max_lights=max(n, 1)) as ctx:
        ctx.set_lights(lights)
        ctx.set_ubo(0, ubo)
        order = list(range(tiles))
        np.random.default_rng(n).shuffle(order)
        for frame_no in range(2):
            for t in order:
                b = t * 1024
                ctx.transform_lights(0, b, min(b + 1024, max(n, b)))
            ctx.cull_begin_range(0, 0, n, hv.RANGE_BIN)
            ctx.cull_finish(0, None)
            cells, idx, total = ctx.download_grid(0)
            assert total == want_total
            assert cells.tobytes() == want_cells.tobytes() and idx.tobytes() == want_idx.tobytes()
        for args in ((0, 100, 2048), (0, 0, 99999999), (0, 2048, 1024)):
            if n >= 5000:
                with pytest.raises(hv.HvcfError):
                    ctx.transform_lights(*args)
