"""z-slab sharding of the cluster grid across the GPUs of one box.

The reference has no multi-GPU path; this is new plumbing around the one exchange step the algorithm
has.  Clusters are independent given all lights (LightCulling.comp:59-88), and
tileIndex = x + y*X + z*X*Y makes a z-range a contiguous cluster range, so rank r owns slices
[z0_r, z1_r) and its light cells / index segment concatenate with its neighbours'.

Per frame:
  1. broadcast the light buffer (48 B x L) from rank 0                      -- dist.broadcast (NCCL)
  2. local: transform + cull + count + scan      (engine.cull_begin)        -- no communication
  3. all-gather the shard totals (4 B each); base_r = sum of totals below r -- dist.all_gather
  4. local: cells + indices written at GLOBAL offsets (engine.cull_finish)
  5. all-gather the light cells (equal slabs) and the index segments        -- dist.all_gather / broadcast
so that every GPU holds the full light grid, as BASELINE.json's north_star asks.

Two implementations of steps 3-5:
  exchange="nccl"  torch.distributed collectives (all-gather of totals, all-gather / per-owner broadcast of the grid);
  exchange="peer"  the GPUs do it themselves over NVLink peer memory (hvcf_cull_sharded, csrc/hvcf_peer.cu): totals are
                   posted into the peers' mailboxes, every rank pushes its cells + index segment into all peers with
                   one kernel and waits for theirs -- no host round trip and no collective call inside the frame.
Slabs need not be equal: far slices hold many more hits than near ones, so `balance_slabs` re-cuts the z-range from
measured per-rank times (a few iterations converge; the contexts are re-created at the new cuts).

torch.distributed is plumbing only.  `engine` is duck-typed (begin/finish/views) so the host logic can be
exercised on CPU tensors under gloo with a stand-in engine (tests/test_sharded_gloo.py); the product
engine is HvcfEngine below, which has no CPU path.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def z_slabs(Z: int, world: int):
    """Contiguous z-ranges, sizes differing by at most one slice; ranks beyond Z get empty ranges."""
    base, rem = divmod(Z, world)
    out, z = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((z, z + n))
        z += n
    return out


def balance_slabs(slabs, times, min_slices=1):
    """Re-cut contiguous z-slabs so that every rank gets the same share of the measured cost.

    slabs: [(z0, z1)] in rank order covering [0, Z); times: measured cost of each rank's slab (any unit).
    The cost is taken as uniform inside a slab (piecewise-constant density), cumulated over z and cut at k/N of the
    total; every rank keeps at least `min_slices` slices when Z allows.  Pure function: unit-tested on the CPU."""
    world = len(slabs)
    Z = slabs[-1][1]
    dens = np.zeros(Z, np.float64)
    for (a, b), t in zip(slabs, times):
        if b > a:
            dens[a:b] = max(float(t), 1e-9) / (b - a)
    if not dens.any():
        return list(slabs)
    cum = np.concatenate([[0.0], np.cumsum(dens)])
    cuts = [0]
    for k in range(1, world):
        target = cum[-1] * k / world
        z = int(np.searchsorted(cum, target, side="left"))
        # the cut nearer to the target among z-1, z
        if z > 0 and abs(cum[z - 1] - target) <= abs(cum[min(z, Z)] - target):
            z -= 1
        cuts.append(z)
    cuts.append(Z)
    if Z >= world * min_slices:
        for k in range(1, world):                       # forward: at least min_slices each
            cuts[k] = max(cuts[k], cuts[k - 1] + min_slices)
        for k in range(world - 1, 0, -1):               # backward: leave room for the ranks above
            cuts[k] = min(cuts[k], cuts[k + 1] - min_slices)
    else:
        for k in range(1, world):
            cuts[k] = min(max(cuts[k], cuts[k - 1]), Z)
    return [(cuts[k], cuts[k + 1]) for k in range(world)]


class CutSearch:
    """Iterative slab balancing from measured per-rank costs, without assuming anything about the cost inside a slab.

    Every measurement (cuts, times) gives the cumulative share of the cost at each of its cuts.  The shares of ALL measurements so
    far are kept as knots of one piecewise-linear, monotone estimate of the cumulative cost over the rows, and the next cuts are its
    k/N quantiles.  New cuts land where the cost is dense, so the estimate refines itself exactly where a piecewise-constant
    density (balance_slabs) overshoots -- a frame counted in rounds has nearly all of its late cost in the few slices at the edge
    of the lit region.  Pure host logic: unit-tested on the CPU (tests/test_sharded_host_logic.py)."""

    def __init__(self, total_units: int, world: int):
        self.total, self.world = int(total_units), int(world)
        self.samples = {}     # unit position -> [cumulative share]
        self.best = None      # (max time, cuts)

    def add(self, cuts, times, score=None):
        """times: per-rank cost of the slabs cut at `cuts` (what the estimate is built from); score: what `best` minimises
        (the measured frame time, when the caller has it: ranks also wait for each other inside a frame, so equal costs
        are not always the shortest frame), default max(times)."""
        times = [max(float(t), 1e-9) for t in times]
        score = max(times) if score is None else float(score)
        if self.best is None or score < self.best[0]:
            self.best = (score, list(cuts))
        cum = np.cumsum(times) / sum(times)
        for k in range(1, self.world):
            self.samples.setdefault(int(cuts[k]), []).append(float(cum[k - 1]))

    def next_cuts(self):
        pos = sorted(self.samples)
        xs = np.array([0] + pos + [self.total], np.float64)
        ys = np.array([0.0] + [float(np.mean(self.samples[p])) for p in pos] + [1.0])
        ys = np.maximum.accumulate(ys)                      # measurements of different cuts may disagree slightly
        cuts = [0]
        for k in range(1, self.world):
            t = k / self.world
            i = int(np.searchsorted(ys, t, side="left"))
            i = min(max(i, 1), len(xs) - 1)
            x0, x1, y0, y1 = xs[i - 1], xs[i], ys[i - 1], ys[i]
            x = x1 if y1 <= y0 else x0 + (x1 - x0) * (t - y0) / (y1 - y0)
            cuts.append(int(round(x)))
        cuts.append(self.total)
        for k in range(1, self.world):
            cuts[k] = max(cuts[k], cuts[k - 1] + 1)
        for k in range(self.world - 1, 0, -1):
            cuts[k] = min(cuts[k], cuts[k + 1] - 1)
        return cuts


class DevicePtrView:
    """Zero-copy torch view of a raw device pointer (via __cuda_array_interface__)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 3,
                                         "strides": None}


def as_tensor(ptr: int, nbytes: int, device) -> torch.Tensor:
    return torch.as_tensor(DevicePtrView(ptr, nbytes), device=device)


class HvcfEngine:
    """One hvcf context restricted to this rank's z-slab; exposes its device buffers as torch tensors."""

    def __init__(self, hvcf, grid, max_lights, z_range, device_index, cap=150, rows=None):
        self.hvcf = hvcf
        self.device = torch.device("cuda", device_index)
        if rows is not None:   # a row-granular shard (row = y + z * Y)
            self.ctx = hvcf.Context(grid=grid, max_lights=max_lights, max_lights_per_cluster=cap, device=device_index, rows=rows)
        else:
            self.ctx = hvcf.Context(grid=grid, max_lights=max_lights, max_lights_per_cluster=cap, device=device_index,
                                    z_begin=z_range[0], z_end=z_range[1])
        b = self.ctx.buffers()
        C = b.cluster_count
        self.cluster_begin, self.cluster_end = b.cluster_begin, b.cluster_end
        self.lights = as_tensor(b.lights, 48 * max_lights, self.device)
        self.cells = as_tensor(b.light_cells, 8 * C, self.device).view(torch.int32).view(C, 2)
        self.indices = as_tensor(b.light_indices, 4 * int(b.light_indices_capacity), self.device).view(torch.int32)
        self.shard_total = as_tensor(self.ctx.shard_total_ptr(), 4, self.device).view(torch.int32)
        self.base = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.num_lights = 0

    def _s(self):
        """torch's CURRENT stream: every hvcf call of the engine goes where the surrounding torch work goes (copies into the
        light buffer, NCCL waits, phase-timing events), also under torch.cuda.stream(...)."""
        return torch.cuda.current_stream(self.device).cuda_stream

    def set_num_lights(self, n):
        # the light bytes already sit in self.lights (broadcast target); just tell the context how many
        b = self.ctx.buffers()
        self.ctx.bind_lights(b.lights, n)
        self.num_lights = n

    def set_ubo(self, frame, ubo):
        self.ctx.set_ubo(frame, ubo, self._s())

    def cull_begin(self, frame):
        self.ctx.cull_begin(frame, self._s())

    def cull_finish(self, frame, base: torch.Tensor):
        self.base.copy_(base.view(1))
        self.ctx.cull_finish(frame, self.base.data_ptr(), self._s())

    # -- hvcf_cull_begin in pieces, for pipelining the light broadcast --
    def cull_begin_range(self, frame, begin, end, flags=0):
        self.ctx.cull_begin_range(frame, begin, end, flags, self._s())

    def transform_lights(self, frame, begin, end):
        self.ctx.transform_lights(frame, begin, end, self._s())

    def lights_view(self):
        if not hasattr(self, "_lights_view"):
            ptr, nbytes = self.ctx.lights_view_ptr()
            self._lights_view = as_tensor(ptr, nbytes, self.device)
        return self._lights_view

    # -- exchange over peer memory (csrc/hvcf_peer.cu) --
    def peer_export(self):
        return self.ctx.peer_export()

    def peer_attach(self, rank, world, handles):
        self.ctx.peer_attach(rank, world, handles)
        self.peer_totals = as_tensor(self.ctx.peer_totals_ptr(), 4 * world, self.device).view(torch.int32)
        self._index_base_ptr = self.ctx.peer_index_base_ptr()

    def peer_detach(self):
        self.ctx.peer_detach()

    def peer_exchange_totals(self):
        self.ctx.peer_exchange_totals(self._s())

    def cull_finish_peer(self, frame):
        self.ctx.cull_finish(frame, self._index_base_ptr, self._s())

    def peer_push_grid(self):
        self.ctx.peer_push_grid(self._s())

    def peer_wait_grid(self):
        self.ctx.peer_wait_grid(self._s())

    def cull_sharded(self, frame):
        self.ctx.cull_sharded(frame, self._s())

    def peer_status(self):
        return self.ctx.peer_status()

    def close(self):
        self.ctx.close()


class ShardedCuller:
    """Runs steps 1-5 above on `engine`; every tensor lives on engine's device."""

    def __init__(self, engine, grid, world=None, rank=None, group=None, slabs=None, exchange="nccl"):
        self.engine = engine
        self.group = group
        self.world = dist.get_world_size(group) if world is None else world
        self.rank = dist.get_rank(group) if rank is None else rank
        self.grid = tuple(grid)
        self.slabs = list(slabs) if slabs is not None else z_slabs(grid[2], self.world)
        if exchange not in ("nccl", "peer"):
            raise ValueError("exchange must be 'nccl' or 'peer'")
        self.exchange = exchange
        if exchange == "peer":
            self._attach_peers()
        xy = grid[0] * grid[1]
        self.cluster_ranges = [(a * xy, b * xy) for a, b in self.slabs]
        self.equal = len({b - a for a, b in self.slabs}) == 1
        self._marks = None          # [(phase, event)] of the frames run since enable_phase_timing()

    # ---- optional per-phase device timing (CUDA events on the current stream; bench / profiling only) ----
    def enable_phase_timing(self):
        self._marks = []

    def _mark(self, phase):
        if self._marks is not None and self.engine.lights.is_cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream())
            self._marks.append((phase, ev))

    def phase_ms(self):
        """{phase: total ms} between consecutive marks; call after a synchronize."""
        out = {}
        m = self._marks or []
        for (_, e0), (ph, e1) in zip(m[:-1], m[1:]):
            if ph != "start":
                out[ph] = out.get(ph, 0.0) + e0.elapsed_time(e1)
        return out

    def _attach_peers(self):
        """every rank's IPC handles to every rank (the only host-side exchange of the peer mode, once)."""
        mine = self.engine.peer_export()
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, mine, group=self.group)
        else:
            handles = [mine]
        self.engine.peer_attach(self.rank, self.world, handles)
        if self.world > 1:
            dist.barrier(group=self.group)   # nobody starts a frame before everybody has mapped everybody

    def close(self):
        if self.exchange == "peer":
            if self.world > 1:
                dist.barrier(group=self.group)   # nobody unmaps while a peer may still push
            self.engine.peer_detach()
            if self.world > 1:
                dist.barrier(group=self.group)

    def broadcast_lights(self, n):
        """step 1: rank 0's first 48*n bytes of engine.lights reach every rank."""
        self._mark("start")
        if self.world > 1:
            dist.broadcast(self.engine.lights[: 48 * n], src=0, group=self.group)
        self.engine.set_num_lights(n)
        self._mark("broadcast_lights")

    def frame(self, frame, gather=True):
        e = self.engine
        if self._marks is not None and (not self._marks or self._marks[-1][0] != "broadcast_lights"):
            self._mark("start")
        if self.exchange == "peer":
            return self._frame_peer(frame)
        e.cull_begin(frame)                                           # step 2
        self._mark("cull_begin")
        totals = torch.zeros(self.world, dtype=torch.int32, device=e.shard_total.device)
        if self.world > 1:                                            # step 3
            dist.all_gather_into_tensor(totals, e.shard_total, group=self.group)
        else:
            totals.copy_(e.shard_total)
        bases = torch.cumsum(totals, 0, dtype=torch.int32) - totals
        self._mark("totals")
        e.cull_finish(frame, bases[self.rank])                        # step 4
        self._mark("cull_finish")
        if gather and self.world > 1:                                 # step 5
            self._gather(totals, bases)
            self._mark("gather_grid")
        return totals, bases

    # ---- light exchange pipelined with the first kernels of the frame -------------------------------------------
    @staticmethod
    def chunk_bounds(n, chunks, tile=1024):
        """[(begin, end)] covering [0, n): whole tiles, the last chunk ends at n; at most `chunks` pieces."""
        tiles = max(1, -(-n // tile))
        chunks = max(1, min(chunks, tiles))
        per = -(-tiles // chunks)
        out, b = [], 0
        while b < n or not out:
            e = min(n, b + per * tile)
            out.append((b, e))
            if e >= n:
                break
            b = e
        return out

    def frame_pipelined(self, frame, n, chunks=8, wire="raw"):
        """steps 1-5 with step 1 cut into chunks: while chunk q+1 is on the wire, chunk q is transformed and binned
        (hvcf_cull_begin_range).  wire="raw": the 48 B LightData records travel and every rank transforms them
        (every rank ends up with the reference light buffer); wire="view": rank 0 transforms each chunk once and the
        16 B view-space culling records travel (a third of the bytes; the other ranks hold no LightData);
        wire="gather": lights are OWNED in shards -- rank r holds (and moves) the LightData of light tiles
        [r T/N, (r+1) T/N), transforms only those, and one in-place ncclAllGather of the 16 B records gives every rank
        the whole view buffer: no rank sends more than its share, so the exchange scales with the ranks.
        Peer exchange only."""
        if self.exchange != "peer":
            raise ValueError("frame_pipelined needs exchange='peer'")
        if wire not in ("raw", "view", "gather"):
            raise ValueError("wire must be 'raw', 'view' or 'gather'")
        e = self.engine
        T, B = 1, 2
        if e.num_lights != n:
            e.set_num_lights(n)
        bounds = self.chunk_bounds(n, chunks)
        self._mark("start")
        if self.world == 1:
            for b, en in bounds:
                e.cull_begin_range(frame, b, en, T | B)
        elif wire == "gather":
            view = e.lights_view()
            tiles = -(-max(n, 1) // 1024)
            per = -(-tiles // self.world)                     # tiles per rank; the buffer is padded to per * world tiles
            lo, hi = self.rank * per * 1024, (self.rank + 1) * per * 1024
            padded = tiles * 1024
            if per * self.world * 1024 * 16 > view.numel():
                raise ValueError("wire='gather' needs max_lights padded to world x whole tiles")
            if lo < padded:
                e.transform_lights(frame, lo, min(hi, max(n, lo)))   # the slice that reaches n also writes the padding records
            dist.all_gather_into_tensor(view[: 16 * per * self.world * 1024], view[16 * lo: 16 * hi], group=self.group)
            for b, en in bounds:
                e.cull_begin_range(frame, b, en, B)
        elif wire == "raw":
            works = [dist.broadcast(e.lights[48 * b: 48 * en], src=0, group=self.group, async_op=True) for b, en in bounds]
            for (b, en), w in zip(bounds, works):
                w.wait()                                   # the compute stream waits for this chunk only
                e.cull_begin_range(frame, b, en, T | B)
        else:
            view = e.lights_view()
            padded = -(-max(n, 1) // 1024) * 1024
            works = []
            for b, en in bounds:
                if self.rank == 0:
                    e.cull_begin_range(frame, b, en, T)
                hi = padded if en >= n else en
                works.append(dist.broadcast(view[16 * b: 16 * hi], src=0, group=self.group, async_op=True))
            for (b, en), w in zip(bounds, works):
                if self.rank != 0:
                    w.wait()
                e.cull_begin_range(frame, b, en, B)
            if self.rank == 0:
                for w in works:
                    w.wait()
        self._mark("lights+cull_begin")
        e.peer_exchange_totals()
        self._mark("totals")
        e.cull_finish_peer(frame)
        self._mark("cull_finish")
        e.peer_push_grid()
        self._mark("push_grid")
        e.peer_wait_grid()
        self._mark("wait_grid")
        totals = e.peer_totals[: self.world]
        bases = torch.cumsum(totals, 0, dtype=torch.int32) - totals
        return totals, bases

    def _frame_peer(self, frame):
        """steps 2-5 with the exchange done by the GPUs (hvcf_cull_sharded); with phase timing on, the same five
        stream-ordered calls are issued one by one so that events can be recorded between them."""
        e = self.engine
        if self._marks is None:
            e.cull_sharded(frame)
        else:
            e.cull_begin(frame)
            self._mark("cull_begin")
            e.peer_exchange_totals()
            self._mark("totals")
            e.cull_finish_peer(frame)
            self._mark("cull_finish")
            e.peer_push_grid()
            self._mark("push_grid")
            e.peer_wait_grid()
            self._mark("wait_grid")
        totals = e.peer_totals[: self.world]
        bases = torch.cumsum(totals, 0, dtype=torch.int32) - totals
        return totals, bases

    def _gather(self, totals, bases):
        e = self.engine
        if self.equal:
            a, b = self.cluster_ranges[self.rank]
            dist.all_gather_into_tensor(e.cells.view(-1), e.cells[a:b].reshape(-1).clone(), group=self.group)
        else:
            for r, (a, b) in enumerate(self.cluster_ranges):
                if b > a:
                    dist.broadcast(e.cells[a:b], src=r, group=self.group)
        # index segments are variable-sized: one broadcast per owner over its [base, base+total) range
        host = torch.stack([bases, totals]).cpu().numpy()
        for r in range(self.world):
            o, n = int(host[0, r]), int(host[1, r])
            if n:
                dist.broadcast(e.indices[o: o + n], src=r, group=self.group)


def merge_reference(cells_per_rank, indices_per_rank):
    """Pure-numpy statement of what steps 3-5 must produce (used by tests)."""
    totals = [int(c["count"].sum()) for c in cells_per_rank]
    bases = np.concatenate([[0], np.cumsum(totals)[:-1]]).astype(np.int64)
    cells = np.concatenate(cells_per_rank).copy()
    off = 0
    for c, b in zip(cells_per_rank, bases):
        cells["offset"][off: off + len(c)] = c["offset"] + b
        off += len(c)
    return cells, np.concatenate(indices_per_rank)


def slice_touch_counts(ubo: np.ndarray, lights: np.ndarray, stride: int = 1) -> np.ndarray:
    """[Z] how many lights' spheres reach each z-slice of the grid in `ubo` (host, float64, a cost proxy only: the
    exact routing test is the kernel's).  A light at view depth d with radius r reaches the slices whose depth
    interval [near (far/near)^(z/Z), near (far/near)^((z+1)/Z)) (AABBGenerator.comp:72-75) meets [d - r, d + r]."""
    u = np.frombuffer(np.ascontiguousarray(ubo).tobytes(), np.float32)
    view = u[16:32].astype(np.float64)             # column-major cameraView (UBOs.h:67-89)
    near, far = float(u[36]), float(u[37])
    Z = int(np.frombuffer(np.ascontiguousarray(ubo).tobytes(), np.uint32)[40])
    pos = lights["position"][::stride].astype(np.float64)
    r = lights["radius"][::stride].astype(np.float64)
    depth = -(view[2] * pos[:, 0] + view[6] * pos[:, 1] + view[10] * pos[:, 2] + view[14] * pos[:, 3])
    lo, hi = depth - r, depth + r
    keep = (hi > near) & (lo < far) & np.isfinite(lo) & np.isfinite(hi)
    lo, hi = np.maximum(lo[keep], near), np.minimum(hi[keep], far * (1 - 1e-12))
    k = Z / np.log(far / near)
    z0 = np.clip(np.floor(np.log(lo / near) * k).astype(np.int64), 0, Z - 1)
    z1 = np.clip(np.floor(np.log(hi / near) * k).astype(np.int64), 0, Z - 1)
    diff = np.zeros(Z + 1, np.int64)
    np.add.at(diff, z0, 1)
    np.add.at(diff, z1 + 1, -1)
    return np.cumsum(diff)[:Z] * stride


def cuts_from_cost(cost, world: int):
    """z_cuts [world + 1]: contiguous slabs of about equal summed cost, every rank at least one slice."""
    cost = np.asarray(cost, np.float64)
    Z = len(cost)
    if world > Z:
        raise ValueError(f"{world} ranks for {Z} slices")
    cum = np.concatenate([[0.0], np.cumsum(np.maximum(cost, 0.0))])
    cuts = [0]
    for k in range(1, world):
        target = cum[-1] * k / world
        z = int(np.searchsorted(cum, target, side="left"))
        if z > 0 and abs(cum[z - 1] - target) <= abs(cum[min(z, Z)] - target):
            z -= 1
        cuts.append(z)
    cuts.append(Z)
    for k in range(1, world):
        cuts[k] = max(cuts[k], cuts[k - 1] + 1)
    for k in range(world - 1, 0, -1):
        cuts[k] = min(cuts[k], cuts[k + 1] - 1)
    return cuts
