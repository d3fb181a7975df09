"""GPU: hvcf_cull_small_noalloc (include/hvcf.h; csrc/hvcf_noalloc.cu) -- the single-kernel culling frame with its scratch in a
file-scope __device__ array of the library, for callers that cannot allocate.  It has no counterpart in the reference; the
existing entry point hvcf_cull is the checker here: same lights, same UBO, same AABBs -> the light cells, the light indices and
the total must be the SAME BYTES (the tolerance of every culling test in this suite: none), on calls in a row that share the
static scratch -- a large grid first, then a small one, then the large one again, then a frame that hits the per-cluster cap.

Collected last on purpose (the name).  The file needs nothing from the tests before it and
passes alone in a fresh process as well as after them."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GRID_A, SCREEN_A = (16, 9, 24), (2560, 1600)
GRID_B, SCREEN_B = (32, 18, 48), (3840, 2160)


@pytest.fixture(scope="module")
def hv():
    from hellovulkan_b200 import hvcf
    hvcf.load()
    return hvcf


class Frame:
    """inputs on the device + the existing entry point's result for them"""

    def __init__(self, hv, grid, screen, n, cap, seed, pos=(0.0, 1.0, 6.0)):
        self.hv, self.n, self.cap = hv, n, cap
        self.clusters = grid[0] * grid[1] * grid[2]
        ubo = hv.host_camera_ubo(grid, screen, pos=pos)
        lights = hv.host_init_lights(n, seed=seed)
        self.ctx = hv.Context(grid=grid, max_lights=max(n, 1), max_lights_per_cluster=cap)
        self.ctx.set_lights(lights)
        self.ctx.set_ubo(0, ubo)
        self.ctx.cull(0)                                   # the existing entry point (also generates the AABBs)
        self.want = self.ctx.download_grid(0)
        b = self.ctx.buffers()
        self.aabbs_ptr, self.lights_ptr = b.aabb, b.lights   # device inputs, read in place
        self.ubo = hv.DeviceBuffer.from_array(np.ascontiguousarray(ubo).view(np.uint8))
        self.capacity = self.clusters * min(n, cap)
        # outputs start as garbage: everything compared below must have been written by the call
        self.cells = hv.DeviceBuffer.from_array(np.full(self.clusters * 2, 0xDEADBEEF, np.uint32))
        self.indices = hv.DeviceBuffer.from_array(np.full(max(self.capacity, 1), 0xDEADBEEF, np.uint32))
        self.total = hv.DeviceBuffer.from_array(np.full(4, 0xDEADBEEF, np.uint32))

    def run_and_check(self, what):
        hv = self.hv
        hv.cull_small_noalloc(self.ubo.ptr, self.lights_ptr, self.n, self.aabbs_ptr, self.clusters, self.cap, self.cells.ptr,
                              self.indices.ptr, self.capacity, self.total.ptr)
        want_cells, want_idx, want_total = self.want
        total = int(self.total.download(np.uint32, 1)[0])
        assert total == want_total, f"{what}: total {total} != {want_total}"
        cells = self.cells.download(hv.CELL_DT, self.clusters)
        assert cells.tobytes() == want_cells.tobytes(), f"{what}: light cells"
        idx = self.indices.download(np.uint32, total) if total else np.zeros(0, np.uint32)
        assert idx.tobytes() == want_idx.tobytes(), f"{what}: light indices"

    def close(self):
        for b in (self.ubo, self.cells, self.indices, self.total):
            b.close()
        self.ctx.close()


def test_no_alloc_entry_point_equals_hvcf_cull_on_calls_in_a_row(hv):
    frames = [Frame(hv, GRID_B, SCREEN_B, 2048, 150, seed=5),                       # 864 CTAs, the most lights the kernel takes
              Frame(hv, GRID_A, SCREEN_A, 1000, 150, seed=12345),                   # the reference's own scene: 108 CTAs
              Frame(hv, GRID_A, SCREEN_A, 1500, 7, seed=9, pos=(0.5, 1.0, 6.0)),    # most clusters at the cap
              Frame(hv, GRID_A, SCREEN_A, 1, 150, seed=3)]
    try:
        assert frames[2].want[2] > 0 and (frames[2].want[0]["count"] == 7).any()
        # every call shares the static scratch and finds the previous call's contents in it: large grid, small grid, large again...
        for rep in range(2):
            for i, f in enumerate(frames):
                f.run_and_check(f"pass {rep} frame {i}")
        for rep in range(3):                                                         # ... and the same frame back to back
            frames[1].run_and_check(f"back to back {rep}")
    finally:
        for f in frames:
            f.close()


def test_no_alloc_entry_point_rejects_what_the_static_scratch_cannot_hold(hv):
    f = Frame(hv, GRID_A, SCREEN_A, 100, 150, seed=1)
    try:
        args = dict(ubo_ptr=f.ubo.ptr, lights_ptr=f.lights_ptr, num_lights=f.n, aabbs_ptr=f.aabbs_ptr, cluster_count=f.clusters, cap=f.cap,
                    cells_ptr=f.cells.ptr, indices_ptr=f.indices.ptr, indices_capacity=f.capacity, total_ptr=f.total.ptr)
        for bad, code, word in [(dict(num_lights=2049, indices_capacity=1 << 30), hv.HVCF_ERR_CAPACITY, "lights"),
                                (dict(cluster_count=131073, indices_capacity=1 << 30), hv.HVCF_ERR_CAPACITY, "clusters"),
                                (dict(indices_capacity=f.capacity - 1), hv.HVCF_ERR_CAPACITY, "capacity"),
                                (dict(cluster_count=0), hv.HVCF_ERR_INVALID_ARGUMENT, "positive"),
                                (dict(total_ptr=None), hv.HVCF_ERR_INVALID_ARGUMENT, "null"),
                                (dict(aabbs_ptr=f.aabbs_ptr + 4), hv.HVCF_ERR_INVALID_ARGUMENT, "misaligned")]:
            with pytest.raises(hv.HvcfError) as e:
                hv.cull_small_noalloc(**{**args, **bad})
            assert e.value.code == code and word in str(e.value), (bad, str(e.value))
        f.run_and_check("after the rejected calls")                                  # nothing was launched, nothing is left behind
    finally:
        f.close()
