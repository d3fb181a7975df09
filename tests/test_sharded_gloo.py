"""Host logic of the z-slab sharded path (hellovulkan_b200/sharded.py) under torch.distributed/gloo with
world_size 2 and 3 on CPU.  The culling engine is a stand-in built on the CPU oracle (this is a TEST of the
plumbing: light broadcast, all-gather of totals, global offsets, gather of cells and index segments); the
product engine (HvcfEngine) has no CPU path and is covered by tests/test_gpu_parity.py and bench.py --gpus N."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

GRID, SCREEN, N_LIGHTS, CAP = (16, 9, 24), (2560, 1600), 3000, 150


class OracleEngine:
    """Duck-typed like sharded.HvcfEngine, on CPU tensors; culls its z-slab with the oracle."""

    def __init__(self, oracle, grid, max_lights, z_range, cap=CAP):
        self.oracle, self.grid, self.cap = oracle, grid, cap
        xy = grid[0] * grid[1]
        C = xy * grid[2]
        self.cluster_begin, self.cluster_end = z_range[0] * xy, z_range[1] * xy
        self.lights = torch.zeros(48 * max_lights, dtype=torch.uint8)
        self.cells = torch.zeros(C, 2, dtype=torch.int32)
        self.indices = torch.zeros(C * cap, dtype=torch.int32)
        self.shard_total = torch.zeros(1, dtype=torch.int32)
        self.num_lights = 0
        self.ubo = None

    def set_num_lights(self, n):
        self.num_lights = n

    def set_ubo(self, frame, ubo):
        self.ubo = ubo

    def cull_begin(self, frame):
        from cpu_checkers import LIGHT_DT
        lights = self.lights[: 48 * self.num_lights].numpy().view(LIGHT_DT)
        aabb = self.oracle.generate_aabbs(self.ubo)
        self._cells, self._idx, total, _ = self.oracle.cull(self.ubo, aabb, lights, cap=self.cap, cluster_begin=self.cluster_begin,
                                                            cluster_end=self.cluster_end)
        self.shard_total[0] = int(total)

    def cull_finish(self, frame, base):
        b = int(base)
        a, e = self.cluster_begin, self.cluster_end
        local = self._cells[a:e] if len(self._cells) == self.cells.shape[0] else self._cells
        self.cells[a:e, 0] = torch.from_numpy((local["offset"].astype(np.int64) + b).astype(np.int32))
        self.cells[a:e, 1] = torch.from_numpy(local["count"].astype(np.int32))
        n = int(self.shard_total[0])
        self.indices[b: b + n] = torch.from_numpy(self._idx[:n].astype(np.int32))

    def close(self):
        pass


UNEQUAL = {2: [(0, 17), (17, 24)], 3: [(0, 13), (13, 14), (14, 24)]}   # what balance_slabs produces: unequal slabs


def _worker(rank, world, port, out_dir, unequal=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cpu_checkers import Oracle
        from hellovulkan_b200.sharded import ShardedCuller, z_slabs
        O = Oracle()
        slabs = UNEQUAL[world] if unequal else z_slabs(GRID[2], world)
        eng = OracleEngine(O, GRID, N_LIGHTS, slabs[rank])
        sc = ShardedCuller(eng, GRID, world=world, rank=rank, slabs=slabs)
        assert sc.equal == (not unequal)
        if rank == 0:  # only rank 0 holds the scene; the others receive it by broadcast
            eng.lights[: 48 * N_LIGHTS] = torch.from_numpy(O.init_lights(N_LIGHTS).view(np.uint8).reshape(-1).copy())
        eng.set_ubo(0, O.camera_ubo(GRID, SCREEN))
        sc.broadcast_lights(N_LIGHTS)
        totals, bases = sc.frame(0, gather=True)
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), cells=eng.cells.numpy(), indices=eng.indices.numpy(),
                 totals=totals.numpy(), bases=bases.numpy(), lights=eng.lights.numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,unequal", [(2, False), (3, False), (2, True), (3, True)])
def test_sharded_frame_matches_single_process(world, unequal, oracle, tmp_path):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path), unequal), nprocs=world, join=True)
    ubo = oracle.camera_ubo(GRID, SCREEN)
    lights = oracle.init_lights(N_LIGHTS)
    oc, oi, ototal, _ = oracle.cull(ubo, oracle.generate_aabbs(ubo), lights)
    outs = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]
    for r, o in enumerate(outs):
        assert int(o["totals"].sum()) == ototal
        assert np.array_equal(o["bases"], np.concatenate([[0], np.cumsum(o["totals"])[:-1]]))
        assert np.array_equal(o["lights"][: 48 * N_LIGHTS], lights.view(np.uint8).reshape(-1)), f"rank {r}: broadcast lights"
        assert np.array_equal(o["cells"][:, 1].astype(np.uint32), oc["count"]), f"rank {r}: counts"
        assert np.array_equal(o["cells"][:, 0].astype(np.uint32), oc["offset"]), f"rank {r}: global offsets"
        assert np.array_equal(o["indices"][:ototal].astype(np.uint32), oi), f"rank {r}: gathered index list"


def test_z_slabs_cover_the_grid():
    from hellovulkan_b200.sharded import z_slabs
    for Z in (1, 5, 24, 96):
        for world in (1, 2, 3, 4, 8):
            s = z_slabs(Z, world)
            assert len(s) == world and s[0][0] == 0 and s[-1][1] == Z
            assert all(a[1] == b[0] for a, b in zip(s, s[1:]))
            sizes = [b - a for a, b in s]
            assert max(sizes) - min(sizes) <= 1
